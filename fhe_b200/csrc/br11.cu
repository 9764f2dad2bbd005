// br11.cu — the N = 2048, two-digit instantiations of the register-resident blind rotation (64 coefficients per lane, 255
// registers, 4 ciphertexts = 8 warps per SM, keys read through L2): STD256 (29-bit Q), STD256_LMKCDEY (30), STD256Q_LMKCDEY (28).
#include "br_launch.cuh"

int lb_launch_br11(luxb200_ctx* c, const BrParams& P, int sched_only) {
    switch (c->ntt.logH) {
        case 4: return sched_only ? launch_sched_t<11, 4, 4>(c, P) : launch_br_t<11, 4, 4>(c, P);
        case 3: return sched_only ? launch_sched_t<11, 3, 4>(c, P) : launch_br_t<11, 3, 4>(c, P);
        case 2: return sched_only ? launch_sched_t<11, 2, 4>(c, P) : launch_br_t<11, 2, 4>(c, P);
        default: return lb_fail(LUXB200_ERR_UNSUPPORTED, "no N = 2048 blind-rotation kernel for a %u-bit Q", 32 - c->ntt.logH);
    }
}
