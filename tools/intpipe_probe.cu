// Integer-pipe issue-rate probe for sm_100a: per-SM lanes/clk of the instructions the NTT/bootstrapping kernels use.
// nvcc -O3 -gencode arch=compute_100a,code=sm_100a -o tools/intpipe_probe tools/intpipe_probe.cu
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>
typedef uint32_t u32;
typedef uint64_t u64;
#define CHAINS 8
#define BODY(NAME, STMT)                                                                  \
    __global__ void NAME(u32* out, u32 iters, u32 seed) {                                 \
        u32 a[CHAINS];                                                                    \
        for (int i = 0; i < CHAINS; ++i) a[i] = seed * (i + 3) + threadIdx.x;             \
        const u32 m = seed | 1u, c = seed ^ 0x9e3779b9u;                                  \
        for (u32 it = 0; it < iters; ++it) {                                              \
            _Pragma("unroll") for (int u = 0; u < 8; ++u) {                               \
                _Pragma("unroll") for (int i = 0; i < CHAINS; ++i) { STMT; }              \
            }                                                                             \
        }                                                                                 \
        u32 r = 0;                                                                        \
        for (int i = 0; i < CHAINS; ++i) r ^= a[i];                                       \
        out[blockIdx.x * blockDim.x + threadIdx.x] = r;                                   \
    }
BODY(k_imad, a[i] = a[i] * m + c)
BODY(k_imadhi, a[i] = __umulhi(a[i], m) + c)
BODY(k_iadd3, asm volatile("add.u32 %0, %0, %1;" : "+r"(a[i]) : "r"(c)); asm volatile("xor.b32 %0, %0, %1;" : "+r"(a[i]) : "r"(m)))
BODY(k_shf, a[i] = __funnelshift_l(a[i], c, 7) ^ m)
BODY(k_lop, a[i] = (a[i] & m) ^ c)
BODY(k_min, a[i] = min(a[i] + c, a[i] - m))
__global__ void k_wide(u32* out, u32 iters, u32 seed) {
    u64 a[CHAINS];
    for (int i = 0; i < CHAINS; ++i) a[i] = seed * (i + 3) + threadIdx.x;
    const u32 m = seed | 1u;
    for (u32 it = 0; it < iters; ++it) {
#pragma unroll
        for (int u = 0; u < 8; ++u) {
#pragma unroll
            for (int i = 0; i < CHAINS; ++i) a[i] = (u64)(u32)a[i] * m + a[i];
        }
    }
    u64 r = 0;
    for (int i = 0; i < CHAINS; ++i) r ^= a[i];
    out[blockIdx.x * blockDim.x + threadIdx.x] = (u32)r ^ (u32)(r >> 32);
}
// mixed: Shoup butterfly (IMAD.HI + 2 IMAD + 2 adds)
__global__ void k_bfly(u32* out, u32 iters, u32 seed) {
    u32 x[8];
    for (int u = 0; u < 8; ++u) x[u] = seed * (u + 1) + threadIdx.x;
    const u32 Q = 134215681u, Q2 = 2 * Q, w = (seed % Q) | 1u, ws = (u32)(((u64)w << 32) / Q);
    for (u32 i = 0; i < iters; ++i) {
#pragma unroll
        for (int r = 0; r < 4; ++r) {
#pragma unroll
            for (int u = 0; u < 8; u += 2) {
                u32 q = __umulhi(x[u + 1], ws);
                u32 t = x[u + 1] * w - q * Q;
                x[u + 1] = x[u] - t + Q2;
                x[u] = x[u] + t;
            }
        }
    }
    u32 r = 0;
    for (int u = 0; u < 8; ++u) r ^= x[u];
    out[blockIdx.x * blockDim.x + threadIdx.x] = r;
}
template <typename F>
double run(F k, int sms, double ops_per_iter, u32* d) {
    const int blocks = sms * 8, threads = 256;
    const u32 iters = 2048;
    cudaEvent_t e0, e1;
    cudaEventCreate(&e0); cudaEventCreate(&e1);
    float best = 1e30f;
    for (int rep = 0; rep < 4; ++rep) {
        cudaEventRecord(e0);
        k<<<blocks, threads>>>(d, iters, 12345u + rep);
        cudaEventRecord(e1);
        cudaEventSynchronize(e1);
        float ms; cudaEventElapsedTime(&ms, e0, e1);
        if (rep && ms < best) best = ms;
    }
    return (double)blocks * threads * iters * ops_per_iter / (best * 1e-3);
}
int main() {
    cudaDeviceProp p; cudaGetDeviceProperties(&p, 0);
    int clk_khz = 0; cudaDeviceGetAttribute(&clk_khz, cudaDevAttrClockRate, 0);
    u32* d; cudaMalloc(&d, (size_t)p.multiProcessorCount * 8 * 256 * 4);
    const double per_clk = 1.0 / (p.multiProcessorCount * (double)clk_khz * 1e3);
    struct { const char* n; double v; } r[] = {
        {"IMAD (mad.lo.u32)", run(k_imad, p.multiProcessorCount, 64, d)},
        {"IMAD.HI + IADD (umulhi+add)", run(k_imadhi, p.multiProcessorCount, 64, d)},
        {"IMAD.WIDE (32x32+64)", run(k_wide, p.multiProcessorCount, 64, d)},
        {"IADD+LOP pair (counted as 2)", run(k_iadd3, p.multiProcessorCount, 128, d)},
        {"SHF+LOP pair (counted as 2)", run(k_shf, p.multiProcessorCount, 128, d)},
        {"LOP3 (and-xor fused)", run(k_lop, p.multiProcessorCount, 64, d)},
        {"IADD,IADD,MIN triple (counted as 3)", run(k_min, p.multiProcessorCount, 192, d)},
        {"Shoup butterfly (per butterfly)", run(k_bfly, p.multiProcessorCount, 16, d)},
    };
    printf("%s, %d SMs, max clock %.0f MHz\n", p.name, p.multiProcessorCount, clk_khz / 1e3);
    for (auto& x : r) printf("%-38s %8.2f Tops/s  %7.1f lanes/clk/SM (at max clock)\n", x.n, x.v / 1e12, x.v * per_clk);
    return 0;
}
