// br_launch.cuh — launch configuration of the warp-pair blind-rotation kernels (br_kernels.cuh, br_sched.cuh; N = 512 / 1024).
// N = 2048 is br_quad.cuh, launched from br11.cu.
#pragma once
#include "api_common.h"
#include "br_sched.cuh"

template <int LOGN, int LOGH, int ROWS>
static int launch_sched_t(luxb200_ctx* c, const BrParams& P) {
#ifdef LB_SCHED_CTS   // A/B builds: ciphertexts per CTA of the scheduled kernel (2 warps each)
    constexpr int CTS = ROWS > 4 ? 4 : LB_SCHED_CTS;
#else
    constexpr int CTS = ROWS > 4 ? 4 : 6;   // digitsG-1 = 3: more registers per thread, fewer ciphertexts per SM
#endif
    using S = BrSmem<LOGN, LOGH, ROWS, CTS>;
    const unsigned grid = (unsigned)((P.count + CTS - 1) / CTS);
    auto k = br_sched_kernel<LOGN, LOGH, ROWS, CTS>;
    LB_CUDA(cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sizeof(S)));
    k<<<grid, 64 * CTS, sizeof(S), c->stream>>>(P);
    c->launches++;
    LB_CUDA(cudaGetLastError());
    return 0;
}
template <int LOGN, int LOGH, int ROWS>
static int launch_br_t(luxb200_ctx* c, const BrParams& P) {
    if (c->p.method == LUXB200_GINX) {
        constexpr int CTS = ROWS > 4 ? 4 : 6;  // 12 warps (8 when digitsG-1 = 3), one CTA per SM
        using S = typename std::conditional<ROWS == 4, GinxSmem<LOGN, LOGH, ROWS, CTS>, BrSmem<LOGN, LOGH, ROWS, CTS>>::type;
        static_assert(sizeof(S) <= 232448, "GINX shared memory exceeds the 227 KB opt-in limit");
        const unsigned grid = (unsigned)((P.count + CTS - 1) / CTS);
        auto k = br_ginx_kernel<LOGN, LOGH, ROWS, CTS>;
        LB_CUDA(cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sizeof(S)));
        k<<<grid, 64 * CTS, sizeof(S), c->stream>>>(P);
    } else {
        return launch_sched_t<LOGN, LOGH, ROWS>(c, P);
    }
    c->launches++;
    LB_CUDA(cudaGetLastError());
    return 0;
}

