// br11.cu — N = 2048 (log2 N = 11) on the register-resident code: four warps per ciphertext (br_quad.cuh).  A translation unit of
// its own so that it compiles beside api.cu.  Sets: every STD256* / STD256Q* name (GINX, AP, LMKCDEY) and hand-made sets with a Q
// below 2^30 and 2 to 4 gadget digits.
#include "api_common.h"
#include "br_quad.cuh"

namespace {

template <int LOGH, int ROWS>
int launch_quad_t(luxb200_ctx* c, const BrParams& B, int sched_only) {
    constexpr int CTS = ROWS > 4 ? 2 : 3;   // 12 warps at 168 registers, 8 with three or four digits (like the N = 1024 kernels)
    using S = QuadSmem<LOGH, ROWS, CTS>;
    static_assert(sizeof(S) <= 232448, "quad shared memory exceeds the 227 KB opt-in limit");
    QuadParams P;
    memset(&P, 0, sizeof(P));
    P.B = B;
    P.C1 = c->ntt1.C;
    P.twB1 = c->ntt1.twB.as<uint2>(); P.twiB1 = c->ntt1.twiB.as<uint2>();
    P.xw = c->xw; P.xws = c->xws; P.xwi = c->xwi; P.xwis = c->xwis;
    const unsigned grid = (unsigned)((B.count + CTS - 1) / CTS);
    if (c->p.method == LUXB200_GINX && !sched_only) {
        auto k = br_quad_kernel<LOGH, ROWS, CTS, true>;
        LB_CUDA(cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sizeof(S)));
        k<<<grid, 128 * CTS, sizeof(S), c->stream>>>(P);
    } else {
        auto k = br_quad_kernel<LOGH, ROWS, CTS, false>;
        LB_CUDA(cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sizeof(S)));
        k<<<grid, 128 * CTS, sizeof(S), c->stream>>>(P);
    }
    c->launches++;
    LB_CUDA(cudaGetLastError());
    return 0;
}

}  // namespace

// sched_only: execute B.sched op lists (external products only) whatever the context's method (ExternalProductBatch)
int lb_launch_quad(luxb200_ctx* c, const BrParams& B, int sched_only) {
    const int key = (int)c->ntt.logH * 10 + (int)c->rows;
    switch (key) {
        case 54: return launch_quad_t<5, 4>(c, B, sched_only);
        case 44: return launch_quad_t<4, 4>(c, B, sched_only);
        case 34: return launch_quad_t<3, 4>(c, B, sched_only);   // STD256 (29-bit Q)
        case 24: return launch_quad_t<2, 4>(c, B, sched_only);
        case 56: return launch_quad_t<5, 6>(c, B, sched_only);   // STD256Q (27)
        case 46: return launch_quad_t<4, 6>(c, B, sched_only);
        case 36: return launch_quad_t<3, 6>(c, B, sched_only);   // STD256_3, STD256_4 (29)
        case 26: return launch_quad_t<2, 6>(c, B, sched_only);
        case 58: return launch_quad_t<5, 8>(c, B, sched_only);   // four digits below 2^27: STD256Q_3, STD256Q_4, STD256Q_4_LMKCDEY
        default:
            return lb_fail(LUXB200_ERR_UNSUPPORTED, "no N = 2048 blind-rotation kernel for a %u-bit Q with %u digit rows", 32 - c->ntt.logH, c->rows);
    }
}
