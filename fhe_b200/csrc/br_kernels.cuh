// br_kernels.cuh — fused blind rotation (RingGSWAccumulator*::EvalAcc) + test-vector prologue
// (BootstrapGateCore / BootstrapFuncCore, binfhe-base-scheme.cpp:525-626) + extraction epilogue
// (Transpose, INTT, b = acc1[0] + offset; binfhe-base-scheme.cpp:111-121, :634-639).
#pragma once
#include <type_traits>

#include "br_common.cuh"

enum { M_AP = 1, M_GINX = 2, M_LMKCDEY = 3 };

template <int LOGN, int LOGH, int ROWS, int CTS>
struct BrSmem {
    using W = WNtt<LOGN, LOGH>;
    uint2 twB[W::NB * 32];
    uint2 twiB[W::NB * 32];
    u32 scratch[2 * CTS][W::N];        // per-warp transpose scratch
    uint4 exch[CTS][ROWS][W::N / 4];   // NTT'd digit polynomials, sigma order, shared by the warp pair
};
// GINX with digitsG-1 = 2: every ciphertext of the CTA multiplies by the same key pair (ek+_i, ek-_i) in step i, and the
// 64 KB pair fits beside the rest (224.5 of 227 KB), so one bulk async copy per step stages it while the transforms
// run; the multiply-accumulate then reads keys with LDS.128 instead of waiting on L2 (ncu v4: long_scoreboard 5 %,
// all of it on the IMAD.WIDE that consume the key loads).
template <int LOGN, int LOGH, int ROWS, int CTS>
struct GinxSmem : BrSmem<LOGN, LOGH, ROWS, CTS> {
    uint4 keys[2][ROWS * 2][(1 << LOGN) / 4];   // [ek+ | ek-][row*2 + component][sigma(N)]
    u64 mbar;
};

// ---- digit pipeline: INTT -> decompose -> NTT -> publish (rows 2k+c) -------------------------------------
template <int LOGN, int LOGH, int ROWS, bool MULTI = false>
__device__ __forceinline__ void digits_from_acc(const u32 (&acc)[(1 << LOGN) / 32], u32 (&dg)[ROWS / 2][(1 << LOGN) / 32],
                                                const BrParams& P, u32* scratch, const uint2* twB, const uint2* twiB,
                                                uint4* exch /* [ROWS][N/4] */, int c, int lane) {
    using W         = WNtt<LOGN, LOGH>;
    constexpr int E = W::E;
    u32 x[E];
#pragma unroll
    for (int r = 0; r < E; ++r) x[r] = acc[r];
    W::inverse(x, P.C, scratch, twiB, P.twifull, lane);
#pragma unroll
    for (int r = 0; r < E; ++r) {
        u32 d[ROWS / 2];
        decompose_coeff<ROWS / 2>(x[r], P.C.Q, P.qhalf, (int)P.gbits, d);
#pragma unroll
        for (int k = 0; k < ROWS / 2; ++k) dg[k][r] = d[k];
    }
    if (MULTI) {
        // all digit polynomials are transformed together (WNtt::forward_multi: +2.9 % for GINX; the scheduled AP/LMKCDEY
        // kernel measured 1-3 % slower with it and keeps the one-by-one form); polynomial k >= 1 borrows the exchange
        // row it is about to be published to (2k+c, this warp's own, idle until the barrier) as its transpose scratch
        u32* smp[ROWS / 2];
        smp[0] = scratch;
#pragma unroll
        for (int k = 1; k < ROWS / 2; ++k) smp[k] = reinterpret_cast<u32*>(exch + (2 * k + c) * (W::N / 4));
        W::template forward_multi<1, ROWS / 2>(dg, P.C, smp, twB, P.twfull, lane);
        __syncwarp();   // every lane is past its transpose loads before the rows are overwritten
    }
#pragma unroll
    for (int k = 0; k < ROWS / 2; ++k) {
        if (!MULTI) W::template forward<1>(dg[k], P.C, scratch, twB, P.twfull, lane);
        uint4* dst = exch + (2 * k + c) * (W::N / 4);
#pragma unroll
        for (int v = 0; v < E / 4; ++v)
            dst[v * 32 + lane] = make_uint4(dg[k][4 * v], dg[k][4 * v + 1], dg[k][4 * v + 2], dg[k][4 * v + 3]);
    }
}

__device__ __forceinline__ u32 comp4(const uint4& v, int e) {
    return e == 0 ? v.x : e == 1 ? v.y : e == 2 ? v.z : v.w;
}

// ---- GINX accumulator update (rgsw-acc-cggi.cpp:132-150) -----------------------------------------------
// acc[c] += (sum_row dct[row] * ek1[row][c]) * mono + (sum_row dct[row] * ek2[row][c]) * monoNeg
// RS: elements between consecutive polynomials of a key block / exchange rows (the quad kernel of br_quad.cuh works on one
// 1024-point half of 2048-element rows); XRED: the wide reduction constant (29- / 30-bit Q at N = 2048)
template <int LOGN, int ROWS, bool SMEM_KEYS = false, int RS = (1 << LOGN), int RED = 0>
__device__ __forceinline__ void mac_ginx(u32 (&acc)[(1 << LOGN) / 32], const u32 (&dg)[ROWS / 2][(1 << LOGN) / 32],
                                         const uint4* exch, const BrParams& P, const u32* ek1, const u32* ek2,
                                         const u32* mp, const u32* mn, int c, int lane) {
    constexpr int N = 1 << LOGN, E = N / 32, ND = ROWS / 2;
    const u32 Q = P.C.Q, Q2 = P.C.Q2;
    // this warp's digit polynomials are rows 2k+c, the partner's rows 2k+1-c: the (runtime) warp parity goes into the key
    // row ADDRESS, not into a per-operand select of the digit
    const uint4* k1o = reinterpret_cast<const uint4*>(ek1) + (size_t)(c * 2 + c) * (RS / 4);         // row c,   component c
    const uint4* k1p = reinterpret_cast<const uint4*>(ek1) + (size_t)((1 - c) * 2 + c) * (RS / 4);   // row 1-c, component c
    const uint4* k2o = reinterpret_cast<const uint4*>(ek2) + (size_t)(c * 2 + c) * (RS / 4);
    const uint4* k2p = reinterpret_cast<const uint4*>(ek2) + (size_t)((1 - c) * 2 + c) * (RS / 4);
    const uint4* xp  = exch + (1 - c) * (RS / 4);
#pragma unroll
    for (int v = 0; v < E / 4; ++v) {
        uint4 o1[ND], p1[ND], o2[ND], p2[ND], dx[ND];
#pragma unroll
        for (int k = 0; k < ND; ++k) {   // rows advance by 2 = 4 polynomials of N/4 vectors
            if (SMEM_KEYS) {             // ek1 / ek2 point into shared memory (GinxSmem::keys)
                o1[k] = k1o[k * RS + v * 32 + lane]; p1[k] = k1p[k * RS + v * 32 + lane];
                o2[k] = k2o[k * RS + v * 32 + lane]; p2[k] = k2p[k * RS + v * 32 + lane];
            } else {
                o1[k] = __ldg(k1o + k * RS + v * 32 + lane); p1[k] = __ldg(k1p + k * RS + v * 32 + lane);
                o2[k] = __ldg(k2o + k * RS + v * 32 + lane); p2[k] = __ldg(k2p + k * RS + v * 32 + lane);
            }
        }
        const uint4 m1 = __ldg(reinterpret_cast<const uint4*>(mp) + v * 32 + lane);
        const uint4 m2 = __ldg(reinterpret_cast<const uint4*>(mn) + v * 32 + lane);
#pragma unroll
        for (int k = 0; k < ND; ++k) dx[k] = xp[2 * k * (RS / 4) + v * 32 + lane];
#pragma unroll
        for (int e = 0; e < 4; ++e) {
            u64 s1 = 0, s2 = 0;
#pragma unroll
            for (int k = 0; k < ND; ++k) {
                const u32 d = dg[k][4 * v + e], x = comp4(dx[k], e);
                s1 = mad_wide(d, comp4(o1[k], e), s1);
                s1 = mad_wide(x, comp4(p1[k], e), s1);
                s2 = mad_wide(d, comp4(o2[k], e), s2);
                s2 = mad_wide(x, comp4(p2[k], e), s2);
            }
            const u32 r1 = red64<RED>(s1, P, Q);
            const u32 r2 = red64<RED>(s2, P, Q);
            u64 t        = mad_wide(r1, comp4(m1, e), (u64)acc[4 * v + e]);
            t            = mad_wide(r2, comp4(m2, e), t);
            acc[4 * v + e] = csub(red64<RED>(t, P, Q), Q2);
        }
    }
}

// ---- AP / LMKCDEY accumulator update: acc[c] = sum_row dct[row] * ek[row][c]  (replace) ---------------
template <int LOGN, int ROWS, int RS = (1 << LOGN), int RED = 0>
__device__ __forceinline__ void mac_replace(u32 (&acc)[(1 << LOGN) / 32], const u32 (&dg)[ROWS / 2][(1 << LOGN) / 32],
                                            const uint4* exch, const BrParams& P, const u32* ek, int c, int lane) {
    constexpr int N = 1 << LOGN, E = N / 32, ND = ROWS / 2;
    const u32 Q = P.C.Q, Q2 = P.C.Q2;
    const uint4* ko = reinterpret_cast<const uint4*>(ek) + (size_t)(c * 2 + c) * (RS / 4);         // own rows 2k+c
    const uint4* kp = reinterpret_cast<const uint4*>(ek) + (size_t)((1 - c) * 2 + c) * (RS / 4);   // partner rows 2k+1-c
    const uint4* xp = exch + (1 - c) * (RS / 4);
#pragma unroll
    for (int v = 0; v < E / 4; ++v) {
        uint4 o1[ND], p1[ND], dx[ND];
#pragma unroll
        for (int k = 0; k < ND; ++k) {
            o1[k] = __ldg(ko + k * RS + v * 32 + lane);
            p1[k] = __ldg(kp + k * RS + v * 32 + lane);
            dx[k] = xp[2 * k * (RS / 4) + v * 32 + lane];
        }
#pragma unroll
        for (int e = 0; e < 4; ++e) {
            u64 s1 = 0;
#pragma unroll
            for (int k = 0; k < ND; ++k) {
                s1 = mad_wide(dg[k][4 * v + e], comp4(o1[k], e), s1);
                s1 = mad_wide(comp4(dx[k], e), comp4(p1[k], e), s1);
            }
            acc[4 * v + e] = csub(red64<RED>(s1, P, Q), Q2);
        }
    }
}

// ---- prologue: test vector -> acc = (0, NTT(m)) --------------------------------------------------------
template <int LOGN, int LOGH>
__device__ __forceinline__ void prologue(u32 (&acc)[(1 << LOGN) / 32], const BrParams& P, u32 ct, u32* scratch,
                                         const uint2* twB, int c, int lane) {
    using W         = WNtt<LOGN, LOGH>;
    constexpr int E = W::E;
    if (c == 0) {
#pragma unroll
        for (int r = 0; r < E; ++r) acc[r] = 0;
        return;
    }
    const u32 b = P.b[ct], half = P.ctmod >> 1;
    const u32* lut = P.lutsel ? P.lut + P.lutsel[ct] : P.lut;
    u32 m[E];
#pragma unroll
    for (int r = 0; r < E; ++r) {
        const u32 j = lane + 32 * r;
        u32 val     = 0;
        if (j % P.factor == 0) {
            const u32 jj = j / P.factor;
            if (jj < half) val = lut[(b + P.ctmod - jj) % P.ctmod];
        }
        m[r] = val;
    }
    W::template forward<1>(m, P.C, scratch, twB, P.twfull, lane);
#pragma unroll
    for (int r = 0; r < E; ++r) acc[r] = m[r];
}

// ---- epilogue: a' = Transpose(acc0) in coefficient form, b = coeff0(acc1) + b_add -----------------------
template <int LOGN, int LOGH>
__device__ __forceinline__ void epilogue(u32 (&acc)[(1 << LOGN) / 32], const BrParams& P, u32 ct, u32* scratch,
                                         const uint2* twiB, int c, int lane) {
    using W         = WNtt<LOGN, LOGH>;
    constexpr int N = W::N, E = W::E;
    const u32 Q     = P.C.Q;
    u32* out        = P.ext + (size_t)ct * (N + 1);
    if (c == 0) {
        // INTT(Transpose(x)) == coefficient automorphism X -> X^-1 of INTT(x): a'[0] = a[0], a'[N-j] = -a[j]
        W::inverse(acc, P.C, scratch, twiB, P.twifull, lane);
#pragma unroll
        for (int r = 0; r < E; ++r) {
            const u32 j = lane + 32 * r;
            const u32 v = acc[r];
            if (j == 0) out[0] = v;
            else out[N - j] = v ? Q - v : 0;
        }
    } else {
        // coefficient 0 of INTT(x) is N^-1 * sum_p x[p]
        u64 s = 0;
#pragma unroll
        for (int r = 0; r < E; ++r) s += acc[r];
#pragma unroll
        for (int o = 16; o >= 1; o >>= 1) s += __shfl_xor_sync(0xffffffffu, s, o);
        if (lane == 0) {
            u32 sm = (u32)(s % Q);
            u32 v  = csub(shoup_mul(sm, P.C.ninv, P.C.ninv_s, Q), Q);
            out[N] = csub(v + P.b_add, Q);
        }
    }
}

// ---- GINX kernel (RingGSWAccumulatorCGGI::EvalAcc, rgsw-acc-cggi.cpp:59-68) -----------------------------
// One CTA per SM with CTS ciphertexts (2*CTS warps).  All warps run the same ~5k-instruction step body; CTA-wide
// barriers at the two exchange points keep the warps that share an SM sub-partition within the same code window,
// so they share instruction-cache lines instead of streaming 3 private copies of the loop (ncu r1: 39 % of
// issue-stall samples were "no_instructions" with pair-only barriers).
template <int LOGN, int LOGH, int ROWS, int CTS>
__global__ void __launch_bounds__(64 * CTS, 1)
br_ginx_kernel(const __grid_constant__ BrParams P) {
    using W         = WNtt<LOGN, LOGH>;
    constexpr bool STAGE = (ROWS == 4);   // see GinxSmem
    using S         = typename std::conditional<STAGE, GinxSmem<LOGN, LOGH, ROWS, CTS>, BrSmem<LOGN, LOGH, ROWS, CTS>>::type;
    constexpr int N = W::N, E = W::E;
    extern __shared__ __align__(16) unsigned char smem_raw[];
    S& sm = *reinterpret_cast<S*>(smem_raw);
    GinxSmem<LOGN, LOGH, ROWS, CTS>& ks = *reinterpret_cast<GinxSmem<LOGN, LOGH, ROWS, CTS>*>(smem_raw);  // STAGE only

    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int slot = warp >> 1, c = warp & 1;
    for (int i = threadIdx.x; i < W::NB * 32; i += blockDim.x) {
        sm.twB[i]  = P.twB[i];
        sm.twiB[i] = P.twiB[i];
    }
    constexpr u32 KEY_BYTES = 2u * ROWS * 2 * N * 4;   // ek+_i and ek-_i are adjacent in the device BSK
    if (STAGE && threadIdx.x == 0) {
        brtma::mbar_init(&ks.mbar, 1);
        if (P.n) brtma::bulk_load(&ks.keys[0][0][0], P.bsk, KEY_BYTES, &ks.mbar);
    }
    __syncthreads();
    const u32 ct_raw = blockIdx.x * CTS + slot;
    const bool live  = ct_raw < P.count;
    const u32 ct     = live ? ct_raw : P.count - 1;  // idle slots shadow the last ciphertext (keeps barriers uniform)

    u32* scratch = sm.scratch[warp];
    uint4* exch  = &sm.exch[slot][0][0];
    u32 acc[E];
    if (P.mode == 0) {
        prologue<LOGN, LOGH>(acc, P, ct, scratch, sm.twB, c, lane);
    } else {
        const uint4* src = reinterpret_cast<const uint4*>(P.acc_io + ((size_t)ct * 2 + c) * N);
#pragma unroll
        for (int v = 0; v < E / 4; ++v) {
            uint4 t        = src[v * 32 + lane];
            acc[4 * v]     = t.x;
            acc[4 * v + 1] = t.y;
            acc[4 * v + 2] = t.z;
            acc[4 * v + 3] = t.w;
        }
    }
    const u32* arow     = P.a + (size_t)ct * P.n;
    const size_t kblock = (size_t)ROWS * 2 * N;
#pragma unroll 1
    for (u32 i = 0; i < P.n; ++i) {
        const u32 ai = arow[i];
        const u32 ap = ((P.ctmod - ai) % P.ctmod) * P.amul;  // (-a_i mod q) * (2N/q); ap = 0 -> monomials are 0
        u32 dg[ROWS / 2][E];
        digits_from_acc<LOGN, LOGH, ROWS, true>(acc, dg, P, scratch, sm.twB, sm.twiB, exch, c, lane);
        __syncthreads();
        const u32* ek1 = P.bsk + (size_t)(2 * i) * kblock;
        const u32* mp  = P.mono + (size_t)ap * N;
        const u32* mn  = P.mono + (size_t)((2 * N - ap) & (2 * N - 1)) * N;
        if (STAGE) {
            brtma::mbar_wait(&ks.mbar, i & 1);
            mac_ginx<LOGN, ROWS, true>(acc, dg, exch, P, reinterpret_cast<const u32*>(&ks.keys[0][0][0]),
                                       reinterpret_cast<const u32*>(&ks.keys[1][0][0]), mp, mn, c, lane);
        } else {
            mac_ginx<LOGN, ROWS>(acc, dg, exch, P, ek1, ek1 + kblock, mp, mn, c, lane);
        }
        __syncthreads();
        if (STAGE && threadIdx.x == 0 && i + 1 < P.n) {   // every warp is past its key reads: refill for step i+1
            brtma::fence_async();
            brtma::bulk_load(&ks.keys[0][0][0], ek1 + 2 * kblock, KEY_BYTES, &ks.mbar);
        }
    }
    if (!live) return;
    if (P.mode == 0) {
        epilogue<LOGN, LOGH>(acc, P, ct, scratch, sm.twiB, c, lane);
    } else {
        uint4* dst = reinterpret_cast<uint4*>(P.acc_io + ((size_t)ct * 2 + c) * N);
#pragma unroll
        for (int v = 0; v < E / 4; ++v)
            dst[v * 32 + lane] = make_uint4(csub(acc[4 * v], P.C.Q), csub(acc[4 * v + 1], P.C.Q),
                                            csub(acc[4 * v + 2], P.C.Q), csub(acc[4 * v + 3], P.C.Q));
    }
}
