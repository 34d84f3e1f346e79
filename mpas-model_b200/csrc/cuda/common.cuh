// Shared definitions of the sm_100a dycore kernels: index macros (reference layout: Fortran order,
// k fastest, 1-based, trailing garbage column), the library stream, error plumbing and the
// host-pointer -> device-mirror translation used by every C-ABI entry point.
#pragma once
#include <cuda_runtime.h>

#include <cstddef>
#include <cstdint>
#include <string>

#include "../../../include/mpas_b200_args.h"

typedef mpas_real real;

// element (k, i) of an array dimensioned (ld, n+1), 1-based like the reference
#define IX2(k, i, ld) ((size_t)((i)-1) * (size_t)(ld) + (size_t)((k)-1))
#define IX3(a, b, i, la, lb) (((size_t)((i)-1) * (size_t)(lb) + (size_t)((b)-1)) * (size_t)(la) + (size_t)((a)-1))

namespace mpasb200 {

// src/framework/mpas_constants.F:25-36 (literals are RKIND in the reference's -fdefault-real-8 build)
#define MPAS_GRAVITY ((real)9.80616)
#define MPAS_RGAS ((real)287.0)
#define MPAS_CP ((real)(7. * 287.0 / 2.))
#define MPAS_CV ((real)(7. * 287.0 / 2. - 287.0))
#define MPAS_RVORD ((real)(461.6 / 287.0))
#define MPAS_PRANDTL ((real)1.0)
#define MPAS_NRELAXZONE 5  // mpas_atm_boundaries.F:19
#define MPAS_NSPECZONE 2   // mpas_atm_boundaries.F:20

// Fortran sign(a,b): |a| with the sign bit of b (so sign(1,+0)=+1, sign(1,-0)=-1)
__device__ __forceinline__ real fsign(real a, real b) { return copysign(a, b); }

__device__ __forceinline__ real flux4(real q_im2, real q_im1, real q_i, real q_ip1, real ua) {  // TI:4655
    return ua * ((real)7. * (q_i + q_im1) - (q_ip1 + q_im2)) / (real)12.0;
}
__device__ __forceinline__ real flux3(real q_im2, real q_im1, real q_i, real q_ip1, real ua, real coef3) {  // TI:4658
    return flux4(q_im2, q_im1, q_i, q_ip1, ua) + coef3 * fabs(ua) * ((q_ip1 - q_im2) - (real)3. * (q_i - q_im1)) / (real)12.0;
}

// exner = x ** rcv with rcv = rgas/(cp-rgas) (TI:3027, 5980).  The reference's gfortran build calls libm pow, which
// is (nearly always) correctly rounded; CUDA's pow is only within 2 ulp and differs from it in the last place
// for ~17% of the arguments, which is the ONLY arithmetic difference between this library and the oracle.
// In double precision rcv = fl(0.4) and 5*rcv - 2 = 2^-53 exactly, so r = x**rcv solves r^5 = x^2 * x^(2^-53):
// one Newton step on CUDA's pow with the residual r0^5 - x^2 (1 + 2^-53 ln x) evaluated in double-double
// (exact products through fma) gives the correctly rounded value except within ~1e-4 ulp of a rounding tie.
__device__ __forceinline__ real pow_rcv(real x, real rcv) {
#ifdef SINGLE_PRECISION
    return powf(x, rcv);
#else
    const double r0 = pow(x, rcv);
    const double eps = __fma_rn(5.0, rcv, -2.0);
    if (eps != 0x1p-53 || !(x > 0.0) || !(r0 > 0.0) || isinf(r0)) return r0;
    // r0^2, r0^4, r0^5 in double-double
    const double h2 = r0 * r0, l2 = __fma_rn(r0, r0, -h2);
    double p = h2 * h2, e = __fma_rn(h2, h2, -p);
    e = e + (h2 * l2 + l2 * h2);
    const double h4 = p + e, l4 = e - (h4 - p);
    p = h4 * r0;
    e = __fma_rn(h4, r0, -p);
    e = e + l4 * r0;
    const double h5 = p + e, l5 = e - (h5 - p);
    // T = x^2 (1 + eps ln x)
    const double xh = x * x, xl = __fma_rn(x, x, -xh);
    const double tsmall = xh * (eps * log(x));
    const double diff = (h5 - xh) + ((l5 - xl) - tsmall);
    const double delta = diff / xh;
    return r0 + (-r0 * delta / 5.0);
#endif
}

// ---- library state ---------------------------------------------------------------------------
cudaStream_t stream();                 // stream every kernel of the library is launched on
void set_stream(cudaStream_t s);
int fail(const char* where, cudaError_t e);  // records message, returns non-zero
int fail_msg(const std::string& msg);
long long launches();                  // kernels launched by the library so far
void count_launch(int n = 1);

// Persistent device workspace owned by the library (the reference allocates/deallocates its scratch around every
// phase, TI:592-626; here it lives for the life of the process).  Grows on demand; never call inside stream capture
// before the first eager step has sized it.
real* workspace(int slot, size_t count);

// Translate a pointer handed over the C ABI: a registered host address (anywhere inside a
// registered range) maps to its device mirror; anything else is taken to be a device pointer.
void* xlate(const void* p);
template <class T>
inline T* X(T* p) { return (T*)xlate((const void*)p); }
template <class T>
inline const T* X(const T* p) { return (const T*)xlate((const void*)p); }

inline int cdiv(long long a, int b) { return (int)((a + b - 1) / b); }

// optional per-kernel CUDA-event timing (mpas_b200_ktime_enable); no-ops when disabled or during stream capture
void ktime_begin(const char* kernel_name);
void ktime_end();

#define KLAUNCH(kernel, nthreads_total, block, ...)                                          \
    do {                                                                                     \
        long long _n = (long long)(nthreads_total);                                          \
        if (_n > 0) {                                                                        \
            mpasb200::ktime_begin(#kernel);                                                  \
            kernel<<<mpasb200::cdiv(_n, (block)), (block), 0, mpasb200::stream()>>>(__VA_ARGS__); \
            mpasb200::ktime_end();                                                           \
            mpasb200::count_launch();                                                        \
        }                                                                                    \
    } while (0)

#define KCHECK(where)                                          \
    do {                                                       \
        cudaError_t _e = cudaGetLastError();                   \
        if (_e != cudaSuccess) return mpasb200::fail(where, _e); \
    } while (0)

// warp-per-column mapping: WPB warps per block, one mesh element (cell / edge / vertex) per warp, lanes along k
#define WPB 8
#define WARP_ENTITY(first, count)                                   \
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;     \
    const int _ent0 = blockIdx.x * WPB + warp;                      \
    if (_ent0 >= (count)) return;                                   \
    const int ent = (first) + _ent0
#define WLAUNCH(kernel, count, shmem, ...)                                                              \
    do {                                                                                                \
        if ((count) > 0) {                                                                              \
            mpasb200::ktime_begin(#kernel);                                                             \
            kernel<<<mpasb200::cdiv((count), WPB), WPB * 32, (shmem), mpasb200::stream()>>>(__VA_ARGS__); \
            mpasb200::ktime_end();                                                                      \
            mpasb200::count_launch();                                                                   \
        }                                                                                               \
    } while (0)

// flattened (k, column) thread mapping: k fastest so that a warp reads contiguous levels
#define FLAT_KC(ncols, nk)                                                   \
    long long _t = (long long)blockIdx.x * blockDim.x + threadIdx.x;         \
    if (_t >= (long long)(ncols) * (long long)(nk)) return;                  \
    const int _c0 = (int)(_t / (nk));                                        \
    const int k = (int)(_t - (long long)_c0 * (nk)) + 1

}  // namespace mpasb200
