// mpas_b200_halo_copy_no_alloc: the same-process halo copy (the reference's copyList leg of
// mpas_dmpar_exch_halo_field2d_real, pack into a message buffer, unpack out of it) for callers that cannot allocate:
// code under CUDA graph capture, tight serving loops, harnesses that forbid cudaMalloc.  The entry point has no counterpart
// in the reference.
//
// The exchange entry points of halo_abi.cu build a plan on the first exchange of an array (index tables and message buffers
// from cudaMalloc).  Here the caller passes the four index lists itself and the message buffer is a file-scope device array
// of fixed size: nothing is allocated, by this file or by a library routine, on any call.
//
//   * limit: n * dim1 <= MPAS_B200_NO_ALLOC_MAX_REALS (2^20 reals: 8 MiB in the double build, 4 MiB in the single build);
//     every GPU that loads the library holds that much until the process exits;
//   * the buffer is shared by all calls and keeps the last call's contents.  Nothing in it has to start at zero: the pack
//     kernel writes every slot the unpack kernel reads (slot lists are permutations of 1..n; a slot outside 1..n is skipped on
//     both sides, so stale contents are never unpacked for a valid list);
//   * calls are ordered on the library stream; calls that overlap on the GPU (several host threads with their own streams,
//     two processes' graphs replayed concurrently on one context) race on the buffer and are not supported.
#include <mutex>
#include <string>

#include "../../../include/mpas_b200.h"
#include "common.cuh"

#ifndef MPAS_B200_NO_ALLOC_MAX_REALS
#define MPAS_B200_NO_ALLOC_MAX_REALS (1u << 20)
#endif

namespace mpasb200 {
namespace {

static __device__ __align__(256) real g_copy_buffer[MPAS_B200_NO_ALLOC_MAX_REALS];

// In host code `g_copy_buffer` is only a handle for the runtime: the device address comes from cudaGetSymbolAddress, once per
// device, with that device current.
constexpr int kMaxDevices = 64;
real* g_buffer_addr[kMaxDevices] = {};
std::mutex g_buffer_mu;

int buffer_address(real** out) {
    int dev = -1;
    cudaError_t e = cudaGetDevice(&dev);
    if (e != cudaSuccess) return fail("mpas_b200_halo_copy_no_alloc: cudaGetDevice", e);
    if (dev < 0 || dev >= kMaxDevices) return fail_msg("mpas_b200_halo_copy_no_alloc: device ordinal beyond 63");
    std::lock_guard<std::mutex> lk(g_buffer_mu);
    if (!g_buffer_addr[dev]) {
        void* p = nullptr;
        e = cudaGetSymbolAddress(&p, g_copy_buffer);
        if (e != cudaSuccess || !p) return fail("mpas_b200_halo_copy_no_alloc: cudaGetSymbolAddress", e);
        g_buffer_addr[dev] = (real*)p;
    }
    *out = g_buffer_addr[dev];
    return 0;
}

// PACK:   buf[(slot[e]-1)*dim1 + j] = array[(col[e]-1)*dim1 + j]   (mpas_dmpar.F:5462: slot = destList of the send list)
// UNPACK: array[(col[e]-1)*dim1 + j] = buf[(slot[e]-1)*dim1 + j]   (mpas_dmpar.F:5521: slot = srcList of the recv list)
template <bool PACK>
__global__ void __launch_bounds__(256) copy_no_alloc_kernel(real* __restrict__ array, real* __restrict__ buf, const int* __restrict__ col,
                                                            const int* __restrict__ slot, long long n, int dim1) {
    const long long t = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (t >= n * dim1) return;
    const long long e = t / dim1;
    const int j = (int)(t - e * dim1);
    const long long s = slot[e], c = col[e];
    if (s < 1 || s > n || c < 1) return;
    const size_t ia = (size_t)(c - 1) * dim1 + j;
    const size_t ib = (size_t)(s - 1) * dim1 + j;
    if (PACK) buf[ib] = array[ia];
    else array[ia] = buf[ib];
}

int halo_copy_no_alloc(real* array, int dim1, int n, const int* sendSrc, const int* sendDest, const int* recvSrc, const int* recvDest) {
    if (dim1 < 1 || n < 0) return fail_msg("mpas_b200_halo_copy_no_alloc: dim1 must be >= 1 and n >= 0");
    const unsigned long long need = (unsigned long long)n * (unsigned long long)dim1;
    if (need > (unsigned long long)MPAS_B200_NO_ALLOC_MAX_REALS)
        return fail_msg("mpas_b200_halo_copy_no_alloc: n * dim1 = " + std::to_string(need) + " reals is beyond the fixed message buffer of " +
                        std::to_string((unsigned long long)MPAS_B200_NO_ALLOC_MAX_REALS) + " reals (MPAS_B200_NO_ALLOC_MAX_REALS); use mpas_b200_exch_halo_field2d");
    if (n == 0) return 0;
    if (!array || !sendSrc || !sendDest || !recvSrc || !recvDest) return fail_msg("mpas_b200_halo_copy_no_alloc: null array or list");
    real* buf = nullptr;
    if (int rc = buffer_address(&buf)) return rc;
    real* d = X(array);
    const int blocks = cdiv((long long)need, 256);
    copy_no_alloc_kernel<true><<<blocks, 256, 0, stream()>>>(d, buf, X(sendSrc), X(sendDest), n, dim1);
    copy_no_alloc_kernel<false><<<blocks, 256, 0, stream()>>>(d, buf, X(recvDest), X(recvSrc), n, dim1);
    count_launch(2);
    cudaError_t e = cudaGetLastError();
    if (e != cudaSuccess) return fail("mpas_b200_halo_copy_no_alloc: launch", e);
    return 0;
}

}  // namespace
}  // namespace mpasb200

extern "C" int mpas_b200_halo_copy_no_alloc(mpas_real* array, int dim1, int n, const int* sendSrc, const int* sendDest, const int* recvSrc,
                                            const int* recvDest) {
    try {
        return mpasb200::halo_copy_no_alloc(array, dim1, n, sendSrc, sendDest, recvSrc, recvDest);
    } catch (const std::exception& e) { return mpasb200::fail_msg(e.what()); }
}
