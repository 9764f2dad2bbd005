// What the shared-memory transpose and per-butterfly twiddle loads cost the IMAD pipe (12 warps/SM, u32 Shoup butterflies).
// B200 r1: register twiddles 15.4, LDS.64 twiddles 14.8, + one 32x32 transpose per 4 stages 11.6 / 11.3 lanes/clk/SM (ideal 16).
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>
typedef uint32_t u32; typedef uint64_t u64;
// V: 0 shared twiddles in regs; 1 twiddles via LDS.64 per butterfly; 2 = 0 + smem transpose each 4 stages; 3 = 1 + transpose
template <int V>
__global__ void __launch_bounds__(384) k_bfly(u32* out, u32 iters, u32 seed, const uint2* gtw) {
    __shared__ uint2 tw[32 * 32];
    extern __shared__ u32 scr_raw[]; u32 (*scr)[1024] = reinterpret_cast<u32 (*)[1024]>(scr_raw);
    for (int i = threadIdx.x; i < 1024; i += blockDim.x) tw[i] = gtw[i];
    __syncthreads();
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    u32 x[32];
    for (int u = 0; u < 32; ++u) x[u] = seed * (u + 1) + threadIdx.x;
    const u32 Q = 134215681u, Q2 = 2 * Q;
    u32 w[4], ws[4];
    for (int k = 0; k < 4; ++k) { w[k] = ((seed + k * 977) % Q) | 1u; ws[k] = (u32)(((u64)w[k] << 32) / Q); }
    for (u32 i = 0; i < iters; ++i) {
#pragma unroll
        for (int s = 0; s < 4; ++s) {
            int bi = 0;
#pragma unroll
            for (int u = 0; u < 32; ++u) {
                if (u & (1 << s)) continue;
                const int v = u | (1 << s);
                u32 ww, wws;
                if (V & 1) { const uint2 t = tw[((s * 4 + (bi & 3)) & 31) * 32 + lane]; ww = t.x; wws = t.y; }
                else { ww = w[s]; wws = ws[s]; }
                ++bi;
                u32 q = __umulhi(x[v], wws);
                u32 t = x[v] * ww - q * Q;
                x[v] = x[u] - t + Q2;
                x[u] = x[u] + t;
            }
        }
        if (V & 2) {
            u32* sm = scr[warp];
            __syncwarp();
#pragma unroll
            for (int r = 0; r < 32; ++r) sm[((lane + 32 * r) ^ r)] = x[r];
            __syncwarp();
#pragma unroll
            for (int r = 0; r < 32; ++r) x[r] = sm[(32 * lane + r) ^ lane];
        }
    }
    u32 r = 0;
    for (int u = 0; u < 32; ++u) r ^= x[u];
    out[blockIdx.x * blockDim.x + threadIdx.x] = r;
}
template <int V> void run(const char* name, int sms, int clk, u32* d, uint2* gtw) {
    cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
    const u32 iters = 1024; float best = 1e30f;
    for (int rep = 0; rep < 4; ++rep) {
        cudaEventRecord(e0);
        cudaFuncSetAttribute(k_bfly<V>, cudaFuncAttributeMaxDynamicSharedMemorySize, 49152); k_bfly<V><<<sms, 384, 49152>>>(d, iters, 12345u + rep, gtw);
        cudaEventRecord(e1); cudaEventSynchronize(e1);
        float ms; cudaEventElapsedTime(&ms, e0, e1);
        if (rep && ms < best) best = ms;
    }
    double bf = (double)sms * 384 * iters * 64.0 / (best * 1e-3);
    printf("%-44s %.2f T/s = %.1f lanes/clk/SM (ideal 16) %s\n", name, bf / 1e12, bf / (sms * (double)clk * 1e3), cudaGetErrorString(cudaGetLastError()));
}
int main() {
    cudaDeviceProp p; cudaGetDeviceProperties(&p, 0);
    int clk; cudaDeviceGetAttribute(&clk, cudaDevAttrClockRate, 0);
    u32* d; cudaMalloc(&d, (size_t)p.multiProcessorCount * 2048 * 4);
    uint2* g; cudaMalloc(&g, 1024 * 8); cudaMemset(g, 0x11, 1024 * 8);
    run<0>("regs twiddles", p.multiProcessorCount, clk, d, g);
    run<1>("LDS.64 twiddle per butterfly", p.multiProcessorCount, clk, d, g);
    run<2>("regs twiddles + transpose / 4 stages", p.multiProcessorCount, clk, d, g);
    run<3>("LDS twiddles + transpose / 4 stages", p.multiProcessorCount, clk, d, g);
    return 0;
}
