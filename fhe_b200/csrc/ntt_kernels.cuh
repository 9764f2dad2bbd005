// ntt_kernels.cuh — standalone batched NTT/INTT (PolyImpl::SwitchFormat, poly-impl.h:420-440).
//   * u32 path (Q < 2^30, 256 <= N <= 2048): one warp per polynomial, ntt_warp.cuh; HBM-bound.
//   * u64 path (Q < 2^62, 2^10 <= N <= 2^17): two-pass decomposition, see ntt64 kernels below.
#pragma once
#include "ntt_warp.cuh"

struct Ntt32Params {
    NttConsts C;
    const uint2* twB;
    const uint2* twiB;
    const uint2* twfull;
    const uint2* twifull;
    u32* data;       // [count][L][N], this launch handles limb `limb`
    u64 count;       // number of polynomials for this limb
    u32 L, limb;
};

// natural-order global <-> register layouts; forward: coefficient (natural) in, EVALUATION (bit-reversed
// index order, as the reference stores it) out; both are plain index order in memory.
template <int LOGN, int LOGH, bool INVERSE, int WARPS>
__global__ void __launch_bounds__(32 * WARPS) ntt32_kernel(const __grid_constant__ Ntt32Params P) {
    using W = WNtt<LOGN, LOGH>;
    constexpr int N = W::N, E = W::E;
    extern __shared__ __align__(16) unsigned char smem_raw[];
    uint2* tws = reinterpret_cast<uint2*>(smem_raw);                      // [NB*32]
    u32* scr   = reinterpret_cast<u32*>(smem_raw + sizeof(uint2) * W::NB * 32);  // [WARPS][N]
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const uint2* src = INVERSE ? P.twiB : P.twB;
    for (int i = threadIdx.x; i < W::NB * 32; i += blockDim.x) tws[i] = src[i];
    __syncthreads();
    u32* sm = scr + warp * N;
    for (u64 poly = (u64)blockIdx.x * WARPS + warp; poly < P.count; poly += (u64)gridDim.x * WARPS) {
        u32* g = P.data + (poly * P.L + P.limb) * N;
        u32 x[E];
#pragma unroll
        for (int r = 0; r < E; ++r) x[r] = g[lane + 32 * r];
        if (!INVERSE) {
            W::template forward<1>(x, P.C, sm, tws, P.twfull, lane);
#pragma unroll
            for (int r = 0; r < E; ++r) x[r] = csub(x[r], P.C.Q);
            __syncwarp();
            W::store_B(sm, x, lane);
            __syncwarp();
            W::load_A(sm, x, lane);
        } else {
            __syncwarp();
            W::store_A(sm, x, lane);
            __syncwarp();
            W::load_B(sm, x, lane);
            W::inverse(x, P.C, sm, tws, P.twifull, lane);
        }
#pragma unroll
        for (int r = 0; r < E; ++r) g[lane + 32 * r] = x[r];
        __syncwarp();
    }
}

