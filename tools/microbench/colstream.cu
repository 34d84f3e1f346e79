// Microbenchmark: how fast can "one warp per column, lanes along k" stream NA arrays of (ld, ncol) doubles compared with
// a flat one-thread-per-element kernel over the same bytes?  Build: nvcc -O3 -gencode arch=compute_100a,code=sm_100a colstream.cu -o colstream
#include <cstdio>
#include <cstdlib>
#include <cuda_runtime.h>
#define NA 12
struct Arr { const double* a[NA]; };

__global__ void __launch_bounds__(256) flat_kernel(Arr A, double* out, long long n) {
    long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    double s = 0;
#pragma unroll
    for (int a = 0; a < NA; ++a) s += A.a[a][i];
    out[i] = s;
}
// warp per column, simple loop over k
__global__ void __launch_bounds__(256) col_kernel(Arr A, double* out, int ncol, int L, int ld) {
    int lane = threadIdx.x & 31, c = blockIdx.x * 8 + (threadIdx.x >> 5);
    if (c >= ncol) return;
    size_t o = (size_t)c * ld;
    for (int k = lane; k < L; k += 32) {
        double s = 0;
#pragma unroll
        for (int a = 0; a < NA; ++a) s += A.a[a][o + k];
        out[o + k] = s;
    }
}
// warp per column, two levels per lane (all 2*NA loads issued together)
__global__ void __launch_bounds__(256) col2_kernel(Arr A, double* out, int ncol, int L, int ld) {
    int lane = threadIdx.x & 31, c = blockIdx.x * 8 + (threadIdx.x >> 5);
    if (c >= ncol) return;
    size_t o = (size_t)c * ld;
    for (int k = lane; k < L; k += 64) {
        bool vb = k + 32 < L;
        int bo = vb ? 32 : 0;
        double s = 0, t = 0;
#pragma unroll
        for (int a = 0; a < NA; ++a) { s += A.a[a][o + k]; t += A.a[a][o + k + bo]; }
        out[o + k] = s;
        if (vb) out[o + k + 32] = t;
    }
}
// warp per column with a dependent chain: array a+1 is read only after array a arrived (latency-chain model), depth D
template <int D>
__global__ void __launch_bounds__(256) chain_kernel(Arr A, double* out, int ncol, int L, int ld) {
    int lane = threadIdx.x & 31, c = blockIdx.x * 8 + (threadIdx.x >> 5);
    if (c >= ncol) return;
    size_t o = (size_t)c * ld;
    for (int k = lane; k < L; k += 32) {
        double s = 0;
#pragma unroll
        for (int g = 0; g < NA; g += NA / D) {
            double u = 0;
#pragma unroll
            for (int a = g; a < g + NA / D; ++a) u += A.a[a][o + k + (s != 123.456 ? 0 : 1)];  // address depends on s
            s += u;
        }
        out[o + k] = s;
    }
}
// gather: NA/2 arrays at the own column, NA/2 arrays at neighbour columns c + off[j]
__global__ void __launch_bounds__(256) gather_kernel(Arr A, double* out, int ncol, int L, int ld) {
    int lane = threadIdx.x & 31, c = blockIdx.x * 8 + (threadIdx.x >> 5);
    if (c >= ncol) return;
    size_t o = (size_t)c * ld;
    for (int k = lane; k < L; k += 32) {
        double s = 0;
#pragma unroll
        for (int a = 0; a < NA / 2; ++a) s += A.a[a][o + k];
#pragma unroll
        for (int a = NA / 2; a < NA; ++a) {
            int cc = c + (a - NA / 2) * 37 - 100;
            cc = cc < 0 ? 0 : (cc >= ncol ? ncol - 1 : cc);
            s += A.a[0][(size_t)cc * ld + k];
        }
        out[o + k] = s;
    }
}
// chain depth 12 preceded by a software prefetch of all 12 columns (mode 1: prefetch.global.L2 per 128-B line, 2: bulk L2
// prefetch, 3: prefetch.global.L1)
template <int MODE>
__global__ void __launch_bounds__(256) chain_pf_kernel(Arr A, double* out, int ncol, int L, int ld) {
    int lane = threadIdx.x & 31, c = blockIdx.x * 8 + (threadIdx.x >> 5);
    if (c >= ncol) return;
    size_t o = (size_t)c * ld;
#pragma unroll
    for (int a = 0; a < NA; ++a) {
        const double* col = A.a[a] + o;
        if (MODE == 2) {
            if (lane == 0) {
                unsigned long long ad = (unsigned long long)col, a0 = ad & ~15ull;
                unsigned bytes = (unsigned)(((ad + L * 8 + 15) & ~15ull) - a0);
                asm volatile("cp.async.bulk.prefetch.L2.global [%0], %1;" ::"l"(a0), "r"(bytes) : "memory");
            }
        } else if (lane * 16 < L + 15) {
            const double* q = col + min(lane * 16, L - 1);
            if (MODE == 3) asm volatile("prefetch.global.L1 [%0];" ::"l"(q));
            else asm volatile("prefetch.global.L2 [%0];" ::"l"(q));
        }
    }
    for (int k = lane; k < L; k += 32) {
        double s = 0;
#pragma unroll
        for (int a = 0; a < NA; ++a) s += A.a[a][o + k + (s != 123.456 ? 0 : 1)];
        out[o + k] = s;
    }
}
// all 12 columns staged into shared memory with cp.async (8-byte copies), one wait, then the same dependent sum
__global__ void __launch_bounds__(256) chain_ldgsts_kernel(Arr A, double* out, int ncol, int L, int ld) {
    extern __shared__ double sm[];
    int lane = threadIdx.x & 31, w = threadIdx.x >> 5, c = blockIdx.x * 8 + w;
    if (c >= ncol) return;
    size_t o = (size_t)c * ld;
    double* sc = sm + (size_t)w * NA * 64;
    unsigned s0 = (unsigned)__cvta_generic_to_shared(sc + lane);
#pragma unroll
    for (int a = 0; a < NA; ++a) {
        const double* g = A.a[a] + o + lane;
        if (lane < L) asm volatile("cp.async.ca.shared.global [%0], [%1], 8;" ::"r"(s0 + a * 512), "l"(g) : "memory");
        if (lane + 32 < L) asm volatile("cp.async.ca.shared.global [%0], [%1], 8;" ::"r"(s0 + a * 512 + 256), "l"(g + 32) : "memory");
    }
    asm volatile("cp.async.wait_all;" ::: "memory");
    __syncwarp();
    for (int k = lane; k < L; k += 32) {
        double s = 0;
#pragma unroll
        for (int a = 0; a < NA; ++a) s += sc[a * 64 + k + (s != 123.456 ? 0 : 1)];
        out[o + k] = s;
    }
}
// the same with one bulk (TMA) copy per column on a per-warp mbarrier
__global__ void __launch_bounds__(256) chain_tma_kernel(Arr A, double* out, int ncol, int L, int ld) {
    extern __shared__ __align__(16) double sm[];
    __shared__ __align__(8) unsigned long long bar[8];
    int lane = threadIdx.x & 31, w = threadIdx.x >> 5, c = blockIdx.x * 8 + w;
    if (c >= ncol) return;
    size_t o = (size_t)c * ld;
    double* sc = sm + (size_t)w * NA * 64;
    unsigned mb = (unsigned)__cvta_generic_to_shared(&bar[w]);
    int sh[NA];
#pragma unroll
    for (int a = 0; a < NA; ++a) sh[a] = (int)(((unsigned long long)(A.a[a] + o) & 15ull) >> 3);
    if (lane == 0) {
        asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(mb) : "memory");
        asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
        unsigned s0 = (unsigned)__cvta_generic_to_shared(sc), total = 0;
#pragma unroll
        for (int a = 0; a < NA; ++a) {
            unsigned long long ad = (unsigned long long)(A.a[a] + o), a0 = ad & ~15ull;
            unsigned bytes = (unsigned)(((ad + L * 8 + 15) & ~15ull) - a0);
            asm volatile("cp.async.bulk.shared::cta.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(s0 + a * 512), "l"(a0), "r"(bytes), "r"(mb) : "memory");
            total += bytes;
        }
        asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(mb), "r"(total) : "memory");
    }
    __syncwarp();
    unsigned ok;
    do {
        asm volatile("{\n .reg .pred p;\n mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n selp.u32 %0, 1, 0, p;\n}" : "=r"(ok) : "r"(mb), "r"(0) : "memory");
    } while (!ok);
    for (int k = lane; k < L; k += 32) {
        double s = 0;
#pragma unroll
        for (int a = 0; a < NA; ++a) s += sc[a * 64 + sh[a] + k + (s != 123.456 ? 0 : 1)];
        out[o + k] = s;
    }
}
template <class F>
float timeit(F f, int reps = 5) {
    cudaEvent_t e0, e1;
    cudaEventCreate(&e0); cudaEventCreate(&e1);
    f(); cudaDeviceSynchronize();
    cudaEventRecord(e0);
    for (int r = 0; r < reps; ++r) f();
    cudaEventRecord(e1); cudaEventSynchronize(e1);
    float ms; cudaEventElapsedTime(&ms, e0, e1);
    return ms / reps;
}
int main() {
    const int ncol = 491520;
    for (int cfg = 0; cfg < 2; ++cfg) {
        int L = 55, ld = cfg == 0 ? 55 : (cfg == 1 ? 56 : 64);
        size_t n = (size_t)ncol * ld + 64;
        Arr A; double* out;
        for (int a = 0; a < NA; ++a) { double* p; cudaMalloc(&p, n * 8); cudaMemset(p, 0, n * 8); A.a[a] = p; }
        cudaMalloc(&out, n * 8);
        double bytes = (double)ncol * L * 8 * (NA + 1);
        int nb = (ncol + 7) / 8;
        if (cfg == 0) {
            long long ne = (long long)ncol * L;
            float ms = timeit([&] { flat_kernel<<<(unsigned)((ne + 255) / 256), 256>>>(A, out, ne); });
            printf("flat            ld=%d  %.1f us  %.0f GB/s\n", ld, ms * 1e3, bytes / ms * 1e-6);
        }
        float ms = timeit([&] { col_kernel<<<nb, 256>>>(A, out, ncol, L, ld); });
        printf("col             ld=%d  %.1f us  %.0f GB/s\n", ld, ms * 1e3, bytes / ms * 1e-6);
        ms = timeit([&] { col2_kernel<<<nb, 256>>>(A, out, ncol, L, ld); });
        printf("col2            ld=%d  %.1f us  %.0f GB/s\n", ld, ms * 1e3, bytes / ms * 1e-6);
        ms = timeit([&] { chain_kernel<2><<<nb, 256>>>(A, out, ncol, L, ld); });
        printf("chain depth 2   ld=%d  %.1f us  %.0f GB/s\n", ld, ms * 1e3, bytes / ms * 1e-6);
        ms = timeit([&] { chain_kernel<4><<<nb, 256>>>(A, out, ncol, L, ld); });
        printf("chain depth 4   ld=%d  %.1f us  %.0f GB/s\n", ld, ms * 1e3, bytes / ms * 1e-6);
        ms = timeit([&] { chain_kernel<12><<<nb, 256>>>(A, out, ncol, L, ld); });
        printf("chain depth 12  ld=%d  %.1f us  %.0f GB/s\n", ld, ms * 1e3, bytes / ms * 1e-6);
        ms = timeit([&] { chain_pf_kernel<1><<<nb, 256>>>(A, out, ncol, L, ld); });
        printf("chain 12 + prefetch.L2      ld=%d  %.1f us  %.0f GB/s\n", ld, ms * 1e3, bytes / ms * 1e-6);
        ms = timeit([&] { chain_pf_kernel<2><<<nb, 256>>>(A, out, ncol, L, ld); });
        printf("chain 12 + bulk prefetch L2 ld=%d  %.1f us  %.0f GB/s\n", ld, ms * 1e3, bytes / ms * 1e-6);
        ms = timeit([&] { chain_pf_kernel<3><<<nb, 256>>>(A, out, ncol, L, ld); });
        printf("chain 12 + prefetch.L1      ld=%d  %.1f us  %.0f GB/s\n", ld, ms * 1e3, bytes / ms * 1e-6);
        cudaFuncSetAttribute(chain_ldgsts_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 8 * NA * 64 * 8);
        cudaFuncSetAttribute(chain_tma_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 8 * NA * 64 * 8);
        ms = timeit([&] { chain_ldgsts_kernel<<<nb, 256, 8 * NA * 64 * 8>>>(A, out, ncol, L, ld); });
        printf("chain 12 via cp.async smem  ld=%d  %.1f us  %.0f GB/s\n", ld, ms * 1e3, bytes / ms * 1e-6);
        ms = timeit([&] { chain_tma_kernel<<<nb, 256, 8 * NA * 64 * 8>>>(A, out, ncol, L, ld); });
        printf("chain 12 via TMA bulk smem  ld=%d  %.1f us  %.0f GB/s\n", ld, ms * 1e3, bytes / ms * 1e-6);
        double gbytes = (double)ncol * L * 8 * (NA / 2 + 1 + 1);
        ms = timeit([&] { gather_kernel<<<nb, 256>>>(A, out, ncol, L, ld); });
        printf("gather (6 own + 6 nbr of one array) ld=%d  %.1f us  %.0f GB/s (unique bytes)\n", ld, ms * 1e3, gbytes / ms * 1e-6);
        for (int a = 0; a < NA; ++a) cudaFree((void*)A.a[a]);
        cudaFree(out);
    }
    printf("%s\n", cudaGetErrorString(cudaGetLastError()));
    return 0;
}
