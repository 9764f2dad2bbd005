// Butterfly throughput vs resident warps per SM sub-partition (u32 Shoup butterfly, 16 independent per stage).
// nvcc -O3 -gencode arch=compute_100a,code=sm_100a -o occ_probe tools/occ_probe.cu
// B200 r1: 1 warp/SMSP 9.3, 2: 14.0, 3: 15.6, 4: 15.8, 8: 15.9 lanes/clk/SM (ideal 16) - 3 warps already fill the IMAD pipe.
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>
typedef uint32_t u32; typedef uint64_t u64;
// 16 independent butterflies per stage like the kernels (32 registers of data), 4 "stages" per iteration
template <int NX>
__global__ void k_bfly(u32* out, u32 iters, u32 seed) {
    u32 x[NX];
    for (int u = 0; u < NX; ++u) x[u] = seed * (u + 1) + threadIdx.x;
    const u32 Q = 134215681u, Q2 = 2 * Q;
    u32 w[4], ws[4];
    for (int k = 0; k < 4; ++k) { w[k] = ((seed + k * 977) % Q) | 1u; ws[k] = (u32)(((u64)w[k] << 32) / Q); }
    for (u32 i = 0; i < iters; ++i) {
#pragma unroll
        for (int s = 0; s < 4; ++s) {
#pragma unroll
            for (int u = 0; u < NX; ++u) {
                if (u & (1 << s)) continue;
                const int v = u | (1 << s);
                u32 q = __umulhi(x[v], ws[s]);
                u32 t = x[v] * w[s] - q * Q;
                x[v] = x[u] - t + Q2;
                x[u] = x[u] + t;
            }
        }
    }
    u32 r = 0;
    for (int u = 0; u < NX; ++u) r ^= x[u];
    out[blockIdx.x * blockDim.x + threadIdx.x] = r;
}
int main() {
    cudaDeviceProp p; cudaGetDeviceProperties(&p, 0);
    int clk; cudaDeviceGetAttribute(&clk, cudaDevAttrClockRate, 0);
    u32* d; cudaMalloc(&d, (size_t)p.multiProcessorCount * 2048 * 4);
    cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
    const u32 iters = 2048;
    for (int threads : {128, 256, 384, 512, 768, 1024}) {
        float best = 1e30f;
        for (int rep = 0; rep < 4; ++rep) {
            cudaEventRecord(e0);
            k_bfly<32><<<p.multiProcessorCount, threads>>>(d, iters, 12345u + rep);
            cudaEventRecord(e1); cudaEventSynchronize(e1);
            float ms; cudaEventElapsedTime(&ms, e0, e1);
            if (rep && ms < best) best = ms;
        }
        double bf = (double)p.multiProcessorCount * threads * iters * 64.0 / (best * 1e-3);
        printf("warps/SMSP %2d: %.2f T butterflies/s = %.1f lanes/clk/SM (ideal 16)\n", threads / 128, bf / 1e12,
               bf / (p.multiProcessorCount * (double)clk * 1e3));
    }
    return 0;
}
