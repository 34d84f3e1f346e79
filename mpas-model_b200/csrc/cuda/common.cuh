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

// x / 12 without the generic division sequence: with r = RN(1/12), q0 = RN(x r), rem = x - 12 q0 (exact in one fma),
// RN(q0 + rem r) is the correctly rounded quotient (Markstein's correction step; for the divisor 3 * 2^2 the exact
// quotient is never closer than 1/6 ulp to a rounding tie, so the O(2^-53 ulp) error of the correction cannot change
// the rounding).  Outside [2^-962, 2^963) (underflow / overflow of the intermediates, zeros, NaN) the plain division
// is used, so the result equals x / 12.0 bit for bit for every input.  3 instructions instead of ~12.
__device__ __forceinline__ real div12(real x) {
#ifdef SINGLE_PRECISION
    return x / (real)12.0;
#else
    // biased exponent in [61, 1985] <=> 2^-962 <= |x| < 2^963: an integer test (zeros, subnormals, inf and NaN fail it)
    const unsigned ex = ((unsigned)__double2hiint(x) >> 20) & 0x7ffu;
    if (ex - 61u > 1924u) {
        if (x == 0.0) return x;
        return x / 12.0;
    }
    const double r = 1.0 / 12.0;
    const double q0 = x * r;
    const double rem = __fma_rn(-q0, 12.0, x);
    return __fma_rn(rem, r, q0);
#endif
}

__device__ __forceinline__ real flux4(real q_im2, real q_im1, real q_i, real q_ip1, real ua) {  // TI:4655
    return div12(ua * ((real)7. * (q_i + q_im1) - (q_ip1 + q_im2)));
}
__device__ __forceinline__ real flux3(real q_im2, real q_im1, real q_i, real q_ip1, real ua, real coef3) {  // TI:4658
    return flux4(q_im2, q_im1, q_i, q_ip1, ua) + div12(coef3 * fabs(ua) * ((q_ip1 - q_im2) - (real)3. * (q_i - q_im1)));
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

// Software prefetch of one column (nlev consecutive reals starting at col) into L2 / L1, issued by a whole warp:
// mode 1 = one prefetch.global.L2 per 128-byte line (lanes 0..), 2 = one bulk L2 prefetch by lane 0 (16-byte aligned
// window around the column), 3 = prefetch.global.L1 per line.  Fire-and-forget: no register, no scoreboard.
__device__ __forceinline__ void prefetch_col(const real* col, int nlev, int lane, int mode) {
    if (mode == 2) {
        if (lane == 0) {
            const uintptr_t a = (uintptr_t)col, a0 = a & ~(uintptr_t)15;
            const unsigned bytes = (unsigned)(((a + (size_t)nlev * sizeof(real) + 15) & ~(uintptr_t)15) - a0);
            asm volatile("cp.async.bulk.prefetch.L2.global [%0], %1;" ::"l"((const void*)a0), "r"(bytes) : "memory");
        }
    } else {
        const int per_line = 128 / (int)sizeof(real);
        if (lane * per_line < nlev + per_line - 1) {
            const real* q = col + min(lane * per_line, nlev - 1);
            if (mode == 3) asm volatile("prefetch.global.L1 [%0];" ::"l"(q));
            else asm volatile("prefetch.global.L2 [%0];" ::"l"(q));
        }
    }
}

// Asynchronous staging of whole columns into shared memory (LDGSTS): every lane copies its levels of the column without
// holding a register or a scoreboard, so a warp can have all the columns of its mesh element in flight at once and pays
// ONE memory latency per element instead of one per loop nest.  Columns are only sizeof(real)-aligned (odd nVertLevels),
// hence element-sized copies.
__device__ __forceinline__ void cp_async_real(real* smem_dst, const real* gsrc) {
    const unsigned s = (unsigned)__cvta_generic_to_shared(smem_dst);
    asm volatile("cp.async.ca.shared.global [%0], [%1], %2;" ::"r"(s), "l"(gsrc), "n"(sizeof(real)) : "memory");
}
__device__ __forceinline__ void stage_col(real* smem_dst, const real* gcol, int n, int lane) {
    for (int k = lane; k < n; k += 32) cp_async_real(smem_dst + k, gcol + k);
}
// same in straight-line form for the hot kernels: sdst_lane = shared-window address of element `lane` of the destination
// column, gsrc_lane = &column[lane]; two predicated copies cover n <= 64 levels, a loop the rest.
__device__ __forceinline__ void cp_async_s(unsigned sdst, const real* g) {
    asm volatile("cp.async.ca.shared.global [%0], [%1], %2;" ::"r"(sdst), "l"(g), "n"(sizeof(real)) : "memory");
}
__device__ __forceinline__ void stage_col_s(unsigned sdst_lane, const real* gsrc_lane, int n, int lane) {
    if (lane < n) cp_async_s(sdst_lane, gsrc_lane);
    if (lane + 32 < n) cp_async_s(sdst_lane + 32 * (unsigned)sizeof(real), gsrc_lane + 32);
    if (n > 64) {
#pragma unroll 1
        for (int d = 64; lane + d < n; d += 32) cp_async_s(sdst_lane + d * (unsigned)sizeof(real), gsrc_lane + d);
    }
}
__device__ __forceinline__ void stage_wait() {
    asm volatile("cp.async.wait_all;" ::: "memory");
    __syncwarp();
}

// ---- TMA column staging -------------------------------------------------------------------------------------------
// One cp.async.bulk (UBLKCP) per column, issued by lane 0 of the warp that owns the mesh element, completing on a
// per-warp mbarrier: the copy engine moves the column straight into shared memory, so staging costs the LSU / L1 data
// path nothing and the warp one instruction per column.  Bulk copies need 16-byte aligned addresses and sizes while a
// column is only sizeof(real)-aligned (odd nVertLevels): the copy covers the 16-byte aligned window around the column
// (at most 8 bytes before and 15 after it, inside the array or the allocation granule) and the column's first element
// lands at element `shift` = (address & 15) / sizeof(real) of its slot.  A slot is tma_slot(n) reals (multiple of 16 B).
__host__ __device__ inline int tma_slot(int n) { return (int)((((size_t)n * sizeof(real) + 16 + 15) & ~(size_t)15) / sizeof(real)); }
__device__ __forceinline__ void mbar_init(unsigned mbar, unsigned count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(mbar), "r"(count) : "memory");
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
}
__device__ __forceinline__ void mbar_arrive_expect_tx(unsigned mbar, unsigned bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(mbar), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_wait(unsigned mbar, unsigned parity) {
    unsigned ok;
    do {
        asm volatile(
            "{\n .reg .pred p;\n mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n selp.u32 %0, 1, 0, p;\n}"
            : "=r"(ok)
            : "r"(mbar), "r"(parity)
            : "memory");
    } while (!ok);
}
// One lane of a converged warp (elect.sync): the guard of the bulk-copy issue.  Unlike `lane == 0` it tells the compiler that
// exactly one thread runs the block, so a bulk copy with warp-uniform operands is a single predicated instruction instead
// of a loop over the active threads.
__device__ __forceinline__ bool elect_one() {
    unsigned pred;
    asm volatile("{\n .reg .pred p;\n elect.sync _|p, 0xffffffff;\n selp.u32 %0, 1, 0, p;\n}" : "=r"(pred));
    return pred != 0;
}
// Element shift of a column inside its slot (every lane), and the copy itself (issuing lane only; adds its size to total).
__device__ __forceinline__ int tma_shift(const real* gcol) { return (int)(((unsigned)(unsigned long long)gcol & 15u) / (unsigned)sizeof(real)); }
// size of that copy, for callers that post the expected byte count before they issue it
__device__ __forceinline__ unsigned tma_col_bytes(const real* gcol, int n) {
    const unsigned sh = (unsigned)(unsigned long long)gcol & 15u;
    return (sh + (unsigned)n * (unsigned)sizeof(real) + 15u) & ~15u;
}
__device__ __forceinline__ void tma_issue_col(unsigned sslot, const real* gcol, int n, unsigned mbar, unsigned& total) {
    const unsigned long long a = (unsigned long long)gcol;
    const unsigned sh = (unsigned)a & 15u;
    const unsigned bytes = (sh + (unsigned)n * (unsigned)sizeof(real) + 15u) & ~15u;
    asm volatile("cp.async.bulk.shared::cta.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(sslot), "l"(a - sh), "r"(bytes), "r"(mbar)
                 : "memory");
    total += bytes;
}
// &base[off32] in one IMAD.WIDE.U32 (the compiler otherwise spends four instructions on the 64-bit shift-and-add)
__device__ __forceinline__ const real* elem_ptr(const real* base, unsigned off) {
    unsigned long long r;
    asm("mad.wide.u32 %0, %1, %2, %3;" : "=l"(r) : "r"(off), "n"(sizeof(real)), "l"((unsigned long long)base));
    return (const real*)r;
}

// ---- library state ---------------------------------------------------------------------------
cudaStream_t stream();                 // stream every kernel of the library is launched on
void set_stream(cudaStream_t s);
int fail(const char* where, cudaError_t e);  // records message, returns non-zero
int fail_msg(const std::string& msg);
long long launches();                  // kernels launched by the library so far
void count_launch(int n = 1);

// Per-block persistent workspace keyed by (an address that identifies the block, slot); nullptr if it cannot be
// granted (out of memory, a first call inside a stream capture, or a provided view that is too small): the caller
// reports an error.  Slots: 0 = all-scalars monotonic transport (key scale_arr), 1 = dyn_tend edge fluxes (key tend_u).
real* keyed_workspace(const void* key, size_t count, int slot = 0);
// A device-resident Dycore hands the library a view of its own scratch arena instead (not owned, never freed here).
void provide_keyed_workspace(const void* key, int slot, real* p, size_t count);
void release_keyed_workspace(const void* key);

// Layout of the all-scalars monotonic-transport workspace (scalars.cu): [flux_arr | flux_upwind | wdtn | scalar_new | scale(2)],
// every piece num_scalars wide.  Shared with dycore.cu, which provides it as a view of the block's scratch arena.
struct MonoWsLayout {
    size_t n_edge, n_wd, n_cell, total;
    MonoWsLayout(int L, int ns, int nCells, int nEdges) {
        const size_t nel = (size_t)L * ns, nel1 = (size_t)(L + 1) * ns, nC1 = (size_t)nCells + 1, nE1 = (size_t)nEdges + 1;
        n_edge = (nel * nE1 + 1) & ~(size_t)1;
        n_wd = (nel1 * nC1 + 1) & ~(size_t)1;
        n_cell = (nel * nC1 + 1) & ~(size_t)1;
        total = 2 * n_edge + n_wd + 3 * n_cell;
    }
};

// Edge tiles for the shared-memory staged stencil gathers (tiles.cu): consecutive edges [first[t], first[t+1]) (offsets
// from the first edge of the plan's range) need the nc[t] <= ncmax distinct cells cells[t*ncmax ...] (1-based ids);
// lidx[16*e + j] is the position of advCellsForEdge(j+1, e) in its tile's list.
struct EdgeTilePlan {
    int nTiles, ncmax, max_stencil;  // max_stencil: the longest advCellsForEdge list in the range
    const int* first;
    const int* nc;
    const int* cells;
    const unsigned char* lidx;
};
// nullptr when the plan cannot be built now (inside a stream capture, out of memory): callers take the direct-gather kernel
const EdgeTilePlan* edge_tile_plan(const int* d_advCellsForEdge, const int* d_nAdvCellsForEdge, int edgeStart, int nEdgesLoop, int nCells, int ncmax, int temax);
void release_edge_tile_plans(const void* d_advCellsForEdge);  // nullptr: all

// Translate a pointer handed over the C ABI: a registered host address (anywhere inside a
// registered range) maps to its device mirror; anything else is taken to be a device pointer.
void* xlate(const void* p);
template <class T>
inline T* X(T* p) { return (T*)xlate((const void*)p); }
template <class T>
inline const T* X(const T* p) { return (const T*)xlate((const void*)p); }

// acoustic.cu: what the device-resident driver tells the next mpas_b200_advance_acoustic_step_work call of this thread.
//   FIRST (small_step 1): skip the edge pass ru_p = dts * tend_ru; the mpas_b200_divergence_damping_3d call on the same ru_p
//          that follows does it;
//   NEXT:  a substep follows: that damping call also runs the NEXT substep's edge pass (the driver exchanges rho_pp first);
//   DONE:  this call's edge pass has been run that way: skip it.
enum { ACOUSTIC_FOLD_FIRST = 1, ACOUSTIC_FOLD_NEXT = 2, ACOUSTIC_FOLD_DONE = 4 };
void acoustic_fold_next_call(int flags);

// opt a kernel in to more than 48 KB of dynamic shared memory (cached per kernel); false if it cannot be granted
bool ensure_dyn_smem(const void* kernel, size_t bytes);

// developer knob: integer read from the environment variable MPAS_B200_TUNE_<name> (kernel-variant sweeps in one bench run)
int tune(const char* name, int def);

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
    const int lane = threadIdx.x & 31, warp = __shfl_sync(0xffffffffu, threadIdx.x >> 5, 0); \
    const int _ent0 = blockIdx.x * WPB + warp;                      \
    if (_ent0 >= (count)) return;                                   \
    const int ent = (first) + _ent0
#define WLAUNCH(kernel, count, shmem, ...)                                                              \
    do {                                                                                                \
        if ((count) > 0 && !mpasb200::ensure_dyn_smem((const void*)kernel, (shmem)))                    \
            return mpasb200::fail_msg(std::string(#kernel) + ": nVertLevels too large for the shared-memory staging"); \
        if ((count) > 0) {                                                                              \
            mpasb200::ktime_begin(#kernel);                                                             \
            kernel<<<mpasb200::cdiv((count), WPB), WPB * 32, (shmem), mpasb200::stream()>>>(__VA_ARGS__); \
            mpasb200::ktime_end();                                                                      \
            mpasb200::count_launch();                                                                   \
        }                                                                                               \
    } while (0)

// Column index of a flat thread index without the ~25-instruction 64-bit division: the host precomputes
// magic = floor(2^64 / nk) + 1 for the (at most two) inner extents a parameter block is used with, and
// umul64hi(t, magic) is the exact quotient for t * nk < 2^64.  Any other extent takes the plain division.
struct KMagic {
    int n0 = 0, n1 = 0;
    unsigned long long m0 = 0, m1 = 0;
    static unsigned long long magic(int nk) { return 0xFFFFFFFFFFFFFFFFULL / (unsigned long long)nk + 1ULL; }
    void set(int a, int b = 0) {
        n0 = a > 1 ? a : 0; m0 = n0 ? magic(a) : 0;
        n1 = b > 1 ? b : 0; m1 = n1 ? magic(b) : 0;
    }
    __device__ __forceinline__ unsigned long long div(unsigned long long t, int nk) const {
        if (nk == n0) return __umul64hi(t, m0);
        if (nk == n1) return __umul64hi(t, m1);
        return t / (unsigned long long)nk;
    }
};

// flattened (k, column) thread mapping: k fastest so that a warp reads contiguous levels.  Expects the kernel's
// parameter block `p` to carry a KMagic member `km` (set on the host for the extents the block is launched with).
#define FLAT_KC(ncols, nk)                                                                   \
    const unsigned long long _t = (unsigned long long)blockIdx.x * blockDim.x + threadIdx.x;  \
    if (_t >= (unsigned long long)(ncols) * (unsigned long long)(nk)) return;                 \
    const int _c0 = (int)p.km.div(_t, (nk));                                                  \
    const int k = (int)(_t - (unsigned long long)_c0 * (unsigned long long)(nk)) + 1

// the same mapping with the plain 64-bit division (kernels whose integer pipe is the bottleneck: smlstep)
#define FLAT_KC_DIV(ncols, nk)                                               \
    long long _t = (long long)blockIdx.x * blockDim.x + threadIdx.x;         \
    if (_t >= (long long)(ncols) * (long long)(nk)) return;                  \
    const int _c0 = (int)(_t / (nk));                                        \
    const int k = (int)(_t - (long long)_c0 * (nk)) + 1

}  // namespace mpasb200
