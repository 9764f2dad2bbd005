// br_quad.cuh — register-resident blind rotation at N = 2048 (GINX, AP, LMKCDEY; Q < 2^30; 2 to 4 gadget digits): one ciphertext =
// FOUR warps.  Warp (c, h) owns half h (coefficients / evaluation points 1024 h .. 1024 h + 1023) of accumulator component c,
// 32 values per lane like the N = 1024 kernels of br_kernels.cuh, so the step body keeps their code size and register budget
// (an earlier form with one warp per polynomial and 64 values per lane — 255 registers, a 10 k-instruction step body that did
// not fit the instruction caches — reached 0.32 of the IMAD peak; this one 0.46).
//
// The in-place Cooley-Tukey transform of the reference (transformnat-impl.h:327-351) on 2048 points is ONE stage across the
// two halves — (x[k], x[k+1024]) -> (x[k] + w x[k+1024], x[k] - w x[k+1024]), w = table[1] — followed by two independent
// 1024-point transforms whose stage-m twiddles are table[2m + h m + i] (h = 0 top, 1 bottom); the Gentleman-Sande inverse
// (:539-598) mirrors it.  Each warp therefore runs the existing WNtt<10> on its half with its half's twiddle set (N^-1 =
// 2048^-1 folded into the sub-transform's last stage) and the cross stage exchanges 32 registers with the partner warp
// through shared memory under a 64-thread named barrier.  Evaluation index p = 1024 h + p' sits where the reference puts it;
// the device layout of a polynomial is [h][sigma order of p'], so the multiply-accumulate helpers of br_kernels.cuh work on
// one half with a row stride of 2048.
#pragma once
#include "br_sched.cuh"

struct QuadParams {
    BrParams B;          // B.C, B.twB, B.twiB: twiddle set of half 0; everything else as for the warp-pair kernels
    NttConsts C1;        // half 1
    const uint2* twB1;
    const uint2* twiB1;
    u32 xw, xws;         // cross stage: forward table[1] and its Shoup companion
    u32 xwi, xwis;       // inverse table[1]
};

template <int LOGH, int ROWS, int CTS>
struct QuadSmem {
    using W = WNtt<10, LOGH>;
    uint2 twB[2][W::NB * 32];
    uint2 twiB[2][W::NB * 32];
    u32 scratch[4 * CTS][1024];     // per-warp transpose scratch, also the cross-stage mailbox
    uint4 exch[CTS][ROWS][512];     // NTT'd digit polynomials: [row 2k+c][h][sigma(1024)]
    u32 bsum[CTS];
};

// this warp's 32 registers (layout A: index lane + 32 r of its half) <-> the partner warp's, through the two scratch areas.
// TWO = true: a second barrier, so that either scratch may be overwritten right after (the caller has no later pair barrier
// before its next use of the scratch)
template <bool TWO>
__device__ __forceinline__ void quad_swap(const u32 (&mine)[32], u32 (&other)[32], u32* scr, const u32* pscr, int bar, int lane) {
    // any (register, lane) -> word mapping works as long as both warps use the same one: 16-byte vectors, conflict-free
    uint4* w4 = reinterpret_cast<uint4*>(scr);
    const uint4* r4 = reinterpret_cast<const uint4*>(pscr);
#pragma unroll
    for (int v = 0; v < 8; ++v) w4[v * 32 + lane] = make_uint4(mine[4 * v], mine[4 * v + 1], mine[4 * v + 2], mine[4 * v + 3]);
    pair_barrier(bar);
#pragma unroll
    for (int v = 0; v < 8; ++v) {
        const uint4 t = r4[v * 32 + lane];
        other[4 * v] = t.x; other[4 * v + 1] = t.y; other[4 * v + 2] = t.z; other[4 * v + 3] = t.w;
    }
    if (TWO) pair_barrier(bar);
}

// canonical sub-inverse outputs (a in half 0, b in half 1) -> canonical coefficients (a + b, (a - b) w^-1)
template <bool TWO>
__device__ __forceinline__ void quad_inv_cross(u32 (&x)[32], const QuadParams& P, u32* scr, const u32* pscr, int h, int bar, int lane) {
    const u32 Q = P.B.C.Q;
    u32 o[32];
    quad_swap<TWO>(x, o, scr, pscr, bar, lane);
    if (h == 0) {
#pragma unroll
        for (int r = 0; r < 32; ++r) x[r] = csub(x[r] + o[r], Q);
    } else {
#pragma unroll
        for (int r = 0; r < 32; ++r) x[r] = csub(shoup_mul(o[r] - x[r] + Q, P.xwi, P.xwis, Q), Q);
    }
}

// ---- digit pipeline: INTT -> decompose -> NTT -> publish (rows 2k+c, half h) -----------------------------------------------
template <int LOGH, int ROWS, bool MULTI>
__device__ __forceinline__ void quad_digits(const u32 (&acc)[32], u32 (&dg)[ROWS / 2][32], const QuadParams& P, const NttConsts& C,
                                            u32* scr, const u32* pscr, const uint2* twB, const uint2* twiB,
                                            uint4* exs /* [ROWS][512] */, int c, int h, int bar, int lane) {
    using W = WNtt<10, LOGH>;
    constexpr int ND = ROWS / 2;
    const u32 Q = P.B.C.Q, Q2 = P.B.C.Q2;
    u32 x[32];
#pragma unroll
    for (int r = 0; r < 32; ++r) x[r] = acc[r];
    W::inverse(x, C, scr, twiB, nullptr, lane);
    quad_inv_cross<false>(x, P, scr, pscr, h, bar, lane);   // the barriers below cover the scratch reuse
#pragma unroll
    for (int r = 0; r < 32; ++r) {
        u32 d[ND];
        decompose_coeff<ND>(x[r], Q, P.B.qhalf, (int)P.B.gbits, d);
#pragma unroll
        for (int k = 0; k < ND; ++k) dg[k][r] = d[k];
    }
    // forward cross stage of all digit polynomials at once, through the half-rows this pair is about to publish to
    u32* myrow[ND];
    const u32* prow[ND];
#pragma unroll
    for (int k = 0; k < ND; ++k) {
        myrow[k] = reinterpret_cast<u32*>(exs + (2 * k + c) * 512 + h * 256);
        prow[k]  = reinterpret_cast<const u32*>(exs + (2 * k + c) * 512 + (1 - h) * 256);
    }
    if (h) {
#pragma unroll
        for (int k = 0; k < ND; ++k)
#pragma unroll
            for (int r = 0; r < 32; ++r) dg[k][r] = shoup_mul(dg[k][r], P.xw, P.xws, Q);   // w b, [0, 2Q)
    }
#pragma unroll
    for (int k = 0; k < ND; ++k)
#pragma unroll
        for (int v = 0; v < 8; ++v)
            reinterpret_cast<uint4*>(myrow[k])[v * 32 + lane] = make_uint4(dg[k][4 * v], dg[k][4 * v + 1], dg[k][4 * v + 2], dg[k][4 * v + 3]);
    pair_barrier(bar);
#pragma unroll
    for (int k = 0; k < ND; ++k)
#pragma unroll
        for (int v = 0; v < 8; ++v) {
            const uint4 t = reinterpret_cast<const uint4*>(prow[k])[v * 32 + lane];
#pragma unroll
            for (int e = 0; e < 4; ++e) {
                const u32 o = comp4(t, e);
                dg[k][4 * v + e] = h ? o + Q2 - dg[k][4 * v + e] : dg[k][4 * v + e] + o;   // a - w b + 2Q | a + w b : below 3Q
            }
        }
    pair_barrier(bar);   // both halves are read before either is overwritten (transpose scratch of polynomial k >= 1, publish)
    if (MULTI) {
        u32* smp[ND];
        smp[0] = scr;
#pragma unroll
        for (int k = 1; k < ND; ++k) smp[k] = myrow[k];
        W::template forward_multi<3, ND>(dg, C, smp, twB, nullptr, lane);
        __syncwarp();
    }
#pragma unroll
    for (int k = 0; k < ND; ++k) {
        if (!MULTI) W::template forward<3>(dg[k], C, scr, twB, nullptr, lane);
        uint4* dst = reinterpret_cast<uint4*>(myrow[k]);
#pragma unroll
        for (int v = 0; v < 8; ++v)
            dst[v * 32 + lane] = make_uint4(dg[k][4 * v], dg[k][4 * v + 1], dg[k][4 * v + 2], dg[k][4 * v + 3]);
    }
}

// EVALUATION-format automorphism of one polynomial held by a warp pair (poly-impl.h:345-356): out[p] = in[map[p]], map in the
// two-half sigma order with natural source indices 0 .. 2047
template <int LOGH>
__device__ __forceinline__ void quad_permute(u32 (&x)[32], u32* scr, const u32* pscr, const u32* __restrict__ map, int h, int bar, int lane) {
    using W = WNtt<10, LOGH>;
    __syncwarp();
    W::store_B(scr, x, lane);
    pair_barrier(bar);
    const u32* half[2] = {h ? pscr : scr, h ? scr : pscr};
#pragma unroll
    for (int v = 0; v < 8; ++v) {
        const uint4 s = __ldg(reinterpret_cast<const uint4*>(map) + v * 32 + lane);
        x[4 * v]     = half[s.x >> 10][W::phys(s.x & 1023u)];
        x[4 * v + 1] = half[s.y >> 10][W::phys(s.y & 1023u)];
        x[4 * v + 2] = half[s.z >> 10][W::phys(s.z & 1023u)];
        x[4 * v + 3] = half[s.w >> 10][W::phys(s.w & 1023u)];
    }
    pair_barrier(bar);   // both halves are read before either scratch is reused
}

// GINX = true: RingGSWAccumulatorCGGI::EvalAcc (rgsw-acc-cggi.cpp:59-68); false: the op lists of br_sched.cuh
// (AP, rgsw-acc-dm.cpp:118-143; LMKCDEY with its automorphism rounds, rgsw-acc-lmkcdey.cpp:228-287; ExternalProductBatch)
template <int LOGH, int ROWS, int CTS, bool GINX>
__global__ void __launch_bounds__(128 * CTS, 1) br_quad_kernel(const __grid_constant__ QuadParams P) {
    using W = WNtt<10, LOGH>;
    using S = QuadSmem<LOGH, ROWS, CTS>;
    constexpr int NP = 2048;
    constexpr int XRED = (LOGH == 2) ? 2 : 1;   // 30-bit Q: two-word reduction; below: wide-constant Barrett (bound checked on the host)
    extern __shared__ __align__(16) unsigned char smem_raw[];
    S& sm = *reinterpret_cast<S*>(smem_raw);
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int slot = warp >> 2, c = (warp >> 1) & 1, h = warp & 1;
    const int bar = 1 + (warp >> 1);
    for (int i = threadIdx.x; i < W::NB * 32; i += blockDim.x) {
        sm.twB[0][i]  = P.B.twB[i];
        sm.twB[1][i]  = P.twB1[i];
        sm.twiB[0][i] = P.B.twiB[i];
        sm.twiB[1][i] = P.twiB1[i];
    }
    __syncthreads();
    const NttConsts& C = h ? P.C1 : P.B.C;
    const uint2* twB   = sm.twB[h];
    const uint2* twiB  = sm.twiB[h];
    u32* scr           = sm.scratch[warp];
    const u32* pscr    = sm.scratch[warp ^ 1];
    uint4* exs         = &sm.exch[slot][0][0];
    const u32 Q = P.B.C.Q, Q2 = P.B.C.Q2;
    const u32 ct_raw = blockIdx.x * CTS + slot;
    const bool live  = ct_raw < P.B.count;
    const u32 ct     = live ? ct_raw : P.B.count - 1;   // idle slots shadow the last ciphertext (keeps the barriers uniform)

    u32 acc[32];
    if (P.B.mode == 0) {
        // acc = (0, NTT(m)), m[j * factor] = lut[(b - j) mod ctmod]   (BootstrapGateCore / FuncCore)
        if (c == 0) {
#pragma unroll
            for (int r = 0; r < 32; ++r) acc[r] = 0;
        } else {
            const u32 b = P.B.b[ct], half = P.B.ctmod >> 1;
            const u32* lut = P.B.lutsel ? P.B.lut + P.B.lutsel[ct] : P.B.lut;
            u32 o[32];
#pragma unroll
            for (int r = 0; r < 32; ++r) {
                const u32 j = 1024 * h + lane + 32 * r;
                u32 val     = 0;
                if (j % P.B.factor == 0) {
                    const u32 jj = j / P.B.factor;
                    if (jj < half) val = lut[(b + P.B.ctmod - jj) % P.B.ctmod];
                }
                acc[r] = h ? shoup_mul(val, P.xw, P.xws, Q) : val;
            }
            quad_swap<true>(acc, o, scr, pscr, bar, lane);
#pragma unroll
            for (int r = 0; r < 32; ++r) acc[r] = h ? o[r] + Q2 - acc[r] : acc[r] + o[r];
            W::template forward<3>(acc, C, scr, twB, nullptr, lane);
        }
    } else {
        const uint4* src = reinterpret_cast<const uint4*>(P.B.acc_io + ((size_t)ct * 2 + c) * NP + 1024 * h);
#pragma unroll
        for (int v = 0; v < 8; ++v) {
            const uint4 t  = src[v * 32 + lane];
            acc[4 * v]     = t.x;
            acc[4 * v + 1] = t.y;
            acc[4 * v + 2] = t.z;
            acc[4 * v + 3] = t.w;
        }
    }

    const size_t kblock = (size_t)ROWS * 2 * NP;
    if (GINX) {
        const u32* arow = P.B.a + (size_t)ct * P.B.n;
#pragma unroll 1
        for (u32 i = 0; i < P.B.n; ++i) {
            const u32 ai = arow[i];
            const u32 ap = ((P.B.ctmod - ai) % P.B.ctmod) * P.B.amul;   // (-a_i mod q) * (2N/q)
            u32 dg[ROWS / 2][32];
            quad_digits<LOGH, ROWS, true>(acc, dg, P, C, scr, pscr, twB, twiB, exs, c, h, bar, lane);
            __syncthreads();
            const u32* ek1 = P.B.bsk + (size_t)(2 * i) * kblock + 1024 * h;
            const u32* mp  = P.B.mono + (size_t)ap * NP + 1024 * h;
            const u32* mn  = P.B.mono + (size_t)((2 * NP - ap) & (2 * NP - 1)) * NP + 1024 * h;
            mac_ginx<10, ROWS, false, NP, XRED>(acc, dg, exs + 256 * h, P.B, ek1, ek1 + kblock, mp, mn, c, lane);
            __syncthreads();
        }
    } else {
        const u32* ops = P.B.sched + (size_t)ct * P.B.sched_stride;
        bool done = !live;
        u32 pc    = 0;
        u32 op_next = done ? (u32)OP_END : ops[0];
#pragma unroll 1
        for (;;) {
            const u32 op = op_next;
            if ((op & OP_MASK) == OP_END) done = true;
            op_next = done ? (u32)OP_END : ops[pc + 1];
            if (__syncthreads_and(done)) break;
            const u32 ty = done ? (u32)OP_END : (op & OP_MASK), arg = op & ~OP_MASK;
            u32 dg[ROWS / 2][32];
            if (ty == OP_AUTO || (ty == OP_PERM1 && c == 1))
                quad_permute<LOGH>(acc, scr, pscr, P.B.automap + (size_t)arg * NP + 1024 * h, h, bar, lane);
            // external product: every warp; automorphism: only component 0 is key-switched (one inlined copy of the pipeline)
            if (ty == OP_EXT || (ty == OP_AUTO && c == 0))
                quad_digits<LOGH, ROWS, false>(acc, dg, P, C, scr, pscr, twB, twiB, exs, c, h, bar, lane);
            __syncthreads();
            if (ty == OP_EXT) {
                mac_replace<10, ROWS, NP, XRED>(acc, dg, exs + 256 * h, P.B, P.B.bsk + (size_t)arg * kblock + 1024 * h, c, lane);
            } else if (ty == OP_AUTO) {
                constexpr int ND = ROWS / 2;
                const u32* ak = P.B.bsk + P.B.auto_off + (size_t)arg * ((size_t)ND * 2 * NP) + 1024 * h;
                if (c == 0) mac_auto_split<10, ND, 0, ND, false, NP, XRED>(acc, dg, exs + 256 * h, P.B, ak, 0, lane);
                else        mac_auto_split<10, ND, 0, 0, true, NP, XRED>(acc, dg, exs + 256 * h, P.B, ak, 1, lane);
            }
            if (!done) ++pc;
        }
    }
    if (P.B.mode == 0) {
        // a' = INTT(Transpose(acc0)): a'[0] = a[0], a'[N-j] = -a[j];  b = N^-1 sum_p acc1[p] + b_add   (every slot runs the
        // barriers, only live ones store)
        u32* out = P.B.ext + (size_t)ct * (NP + 1);
        if (c == 0) {
            W::inverse(acc, C, scr, twiB, nullptr, lane);
            quad_inv_cross<true>(acc, P, scr, pscr, h, bar, lane);
            if (live) {
#pragma unroll
                for (int r = 0; r < 32; ++r) {
                    const u32 j = 1024 * h + lane + 32 * r;
                    const u32 v = acc[r];
                    if (j == 0) out[0] = v;
                    else out[NP - j] = v ? Q - v : 0;
                }
            }
        } else {
            u64 s = 0;
#pragma unroll
            for (int r = 0; r < 32; ++r) s += acc[r];
#pragma unroll
            for (int o = 16; o >= 1; o >>= 1) s += __shfl_xor_sync(0xffffffffu, s, o);
            if (h == 1 && lane == 0) sm.bsum[slot] = (u32)(s % Q);
            pair_barrier(bar);
            if (h == 0 && lane == 0 && live) {
                const u32 t = (u32)((s % Q + sm.bsum[slot]) % Q);
                const u32 v = csub(shoup_mul(t, C.ninv, C.ninv_s, Q), Q);   // the sub-tables carry 2048^-1
                out[NP]     = csub(v + P.B.b_add, Q);
            }
        }
    } else if (live) {
        uint4* dst = reinterpret_cast<uint4*>(P.B.acc_io + ((size_t)ct * 2 + c) * NP + 1024 * h);
#pragma unroll
        for (int v = 0; v < 8; ++v)
            dst[v * 32 + lane] = make_uint4(csub(acc[4 * v], Q), csub(acc[4 * v + 1], Q), csub(acc[4 * v + 2], Q), csub(acc[4 * v + 3], Q));
    }
}
