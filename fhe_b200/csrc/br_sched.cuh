// br_sched.cuh — AP (RingGSWAccumulatorDM::EvalAcc, rgsw-acc-dm.cpp:62-77) and LMKCDEY
// (RingGSWAccumulatorLMKCDEY::EvalAcc, rgsw-acc-lmkcdey.cpp:70-158) blind rotation.
// Both walk a per-ciphertext op list that a small pre-kernel derives from the LWE mask exactly as the
// reference's loops do; the fused kernel then executes it with the accumulator resident in registers.
#pragma once
#include "br_kernels.cuh"

enum : u32 { OP_END = 0u, OP_EXT = 1u << 30, OP_AUTO = 2u << 30, OP_PERM1 = 3u << 30, OP_MASK = 3u << 30 };

// ---- schedules -------------------------------------------------------------------------------------------------
// AP: for i < n, k < digitsR: a0 = k-th base-baseR digit of (0 - a_i mod q) [u64 arithmetic exactly as
// NativeInteger::ModSubFast + Mod + /=]; if a0 != 0 -> ext-product with ek[i][a0][k]
static __global__ void sched_ap_kernel(const u32* __restrict__ a, u32* __restrict__ sched, u32 n, u32 stride, u64 q, u32 baseR,
                                u32 digitsR, u64 count) {
    const u64 ct = (u64)blockIdx.x * blockDim.x + threadIdx.x;
    if (ct >= count) return;
    const u32* ar = a + ct * n;
    u32* s = sched + ct * stride;
    u32 pos = 0;
    for (u32 i = 0; i < n; ++i) {
        const u64 ai = ar[i];
        u64 aI = (0 < ai) ? (u64)0 + q - ai : 0;  // ModSubFast(0, a, q), wraps like the reference if a > q
        for (u32 k = 0; k < digitsR; ++k, aI /= baseR) {
            const u32 a0 = (u32)(aI % baseR);
            if (a0) s[pos++] = OP_EXT | ((i * baseR + a0) * digitsR + k);
        }
    }
    s[pos] = OP_END;
}

// LMKCDEY: one warp per ciphertext; buckets by logGen[(-a_i mod 2N) | 1] (ascending i inside a bucket), then the
// two sweeps i = N/2-1 .. 1 with automorphisms every numAutoKeys skips (rgsw-acc-lmkcdey.cpp:83-157).
static __global__ void __launch_bounds__(32) sched_lmkcdey_kernel(const u32* __restrict__ a, const i32* __restrict__ logGen,
                                                           u32* __restrict__ sched, u32 n, u32 stride, u32 N, u32 w) {
    extern __shared__ i32 sm_s[];
    const u32 Nh = N / 2, M = 2 * N;
    i32* first = sm_s;               // [2*Nh + 1] : key + Nh, key M -> slot 2*Nh
    i32* next  = sm_s + 2 * Nh + 1;  // [n]
    i32* keyof = next + n;           // [n]
    const u64 ct = blockIdx.x;
    const u32* ar = a + ct * n;
    for (u32 i = threadIdx.x; i < 2 * Nh + 1; i += 32) first[i] = -1;
    for (u32 i = threadIdx.x; i < n; i += 32) {
        const u32 ai = ar[i];
        const u32 aodd = ((ai == 0) ? 0u : (M - ai)) | 1u;
        const i32 lg = logGen[aodd];
        keyof[i] = (lg == (i32)M) ? (i32)(2 * Nh) : lg + (i32)Nh;
    }
    __syncwarp();
    if (threadIdx.x != 0) return;
    for (i32 i = (i32)n - 1; i >= 0; --i) {
        next[i] = first[keyof[i]];
        first[keyof[i]] = i;
    }
    u32* s = sched + ct * stride;
    u32 pos = 0, nSkips = 0;
    s[pos++] = OP_PERM1 | 0u;  // acc[1] <- AutomorphismTransform(2N - 5)
    for (int pass = 0; pass < 2; ++pass) {
        for (u32 i = Nh - 1; i > 0; --i) {
            const i32 slot = pass == 0 ? (i32)Nh - (i32)i : (i32)Nh + (i32)i;
            if (first[slot] >= 0) {
                if (nSkips != 0) { s[pos++] = OP_AUTO | nSkips; nSkips = 0; }
                for (i32 j = first[slot]; j >= 0; j = next[j]) s[pos++] = OP_EXT | (u32)j;
            }
            nSkips++;
            if (nSkips == w || i == 1) { s[pos++] = OP_AUTO | nSkips; nSkips = 0; }
        }
        if (pass == 0) {
            for (i32 j = first[2 * Nh]; j >= 0; j = next[j]) s[pos++] = OP_EXT | (u32)j;  // a_j = -1
            s[pos++] = OP_AUTO | 0u;                                                        // by 2N - 5
        } else {
            for (i32 j = first[Nh]; j >= 0; j = next[j]) s[pos++] = OP_EXT | (u32)j;       // a_j = 5^0
        }
    }
    s[pos] = OP_END;
}

// ---- EVALUATION-format automorphism of a register-resident polynomial (poly-impl.h:345-356) --------------------
// map: sigma-ordered source indices; out[p] = in[map[p]]
template <int LOGN, int LOGH>
__device__ __forceinline__ void permute_eval(u32 (&x)[(1 << LOGN) / 32], u32* scratch, const u32* __restrict__ map, int lane) {
    using W = WNtt<LOGN, LOGH>;
    constexpr int E = W::E;
    __syncwarp();
    W::store_B(scratch, x, lane);
    __syncwarp();
#pragma unroll
    for (int v = 0; v < E / 4; ++v) {
        const uint4 s = __ldg(reinterpret_cast<const uint4*>(map) + v * 32 + lane);
        x[4 * v]     = scratch[W::phys(s.x)];
        x[4 * v + 1] = scratch[W::phys(s.y)];
        x[4 * v + 2] = scratch[W::phys(s.z)];
        x[4 * v + 3] = scratch[W::phys(s.w)];
    }
    __syncwarp();
}

// acc[c] (+)= sum_d dig[d] * ak[d][c]; digits LO <= d < HI are this warp's own (registers), the others come from exch
template <int LOGN, int ND, int LO, int HI, bool ADD, int RS = (1 << LOGN), int RED = 0>
__device__ __forceinline__ void mac_auto_split(u32 (&acc)[(1 << LOGN) / 32], const u32 (&dg)[ND][(1 << LOGN) / 32],
                                               const uint4* exch, const BrParams& P, const u32* ak, int c, int lane) {
    constexpr int N = 1 << LOGN, E = N / 32;
    const u32 Q = P.C.Q, Q2 = P.C.Q2;
#pragma unroll
    for (int v = 0; v < E / 4; ++v) {
        uint4 k1[ND], dx[ND];
#pragma unroll
        for (int d = 0; d < ND; ++d) {
            k1[d] = __ldg(reinterpret_cast<const uint4*>(ak + (size_t)(d * 2 + c) * RS) + v * 32 + lane);
            if (d < LO || d >= HI) dx[d] = exch[2 * d * (RS / 4) + v * 32 + lane];   // warp 0 publishes digit d in row 2d
        }
#pragma unroll
        for (int e = 0; e < 4; ++e) {
            u64 s = ADD ? (u64)acc[4 * v + e] : 0ull;
#pragma unroll
            for (int d = 0; d < ND; ++d)
                s = mad_wide((d >= LO && d < HI) ? dg[d][4 * v + e] : comp4(dx[d], e), comp4(k1[d], e), s);
            acc[4 * v + e] = csub(red64<RED>(s, P, Q), Q2);
        }
    }
}
// ---- scheduled kernel ------------------------------------------------------------------------------------------
// One 2*CTS-warp CTA per SM, like br_ginx_kernel.  Ciphertexts walk their own op lists, but the CTA advances in
// rounds with exactly two CTA-wide barriers per round whatever the op type, so the warps of an SM sub-partition stay
// in the same code window (instruction-cache sharing) and the exchange buffer is race-free.  A ciphertext whose
// list is exhausted idles at the barriers until every list in the CTA is done (op counts differ by a few percent).
template <int LOGN, int LOGH, int ROWS, int CTS>
__global__ void __launch_bounds__(64 * CTS, CTS <= 3 ? 2 : 1)
br_sched_kernel(const __grid_constant__ BrParams P) {
    using W         = WNtt<LOGN, LOGH>;
    using S         = BrSmem<LOGN, LOGH, ROWS, CTS>;
    constexpr int N = W::N, E = W::E, ND = ROWS / 2;
    extern __shared__ __align__(16) unsigned char smem_raw[];
    S& sm = *reinterpret_cast<S*>(smem_raw);
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    // Warp w issues on SM sub-partition w % 4, so ciphertext slots 0, 2, 4 share sub-partitions {0, 1} and slots 1, 3, 5 share
    // {2, 3}.  In an LMKCDEY automorphism round only component 0 runs the digit pipeline; with c = w & 1 every component-0
    // warp would sit on sub-partitions 0 and 2.  Swapping the roles in every other slot pair spreads them 2 : 1.
    const int slot = warp >> 1, c = (warp & 1) ^ ((slot >> 1) & 1);
    for (int i = threadIdx.x; i < W::NB * 32; i += blockDim.x) {
        sm.twB[i]  = P.twB[i];
        sm.twiB[i] = P.twiB[i];
    }
    __syncthreads();
    const u32 ct_raw = blockIdx.x * CTS + slot;
    const bool live  = ct_raw < P.count;
    const u32 ct     = live ? ct_raw : P.count - 1;
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
    const u32* ops      = P.sched + (size_t)ct * P.sched_stride;
    const size_t kblock = (size_t)ROWS * 2 * N;
    const size_t ablock = (size_t)ND * 2 * N;
    bool done = !live;
    u32 pc    = 0;
    // the op word of the NEXT round is fetched a round ahead, so no round starts by waiting on a global load
    // (lists end with OP_END: pc + 1 is readable whenever the current op is not the end)
    u32 op_next = done ? (u32)OP_END : ops[0];
#pragma unroll 1
    for (;;) {
        const u32 op = op_next;
        if ((op & OP_MASK) == OP_END) done = true;
        op_next = done ? (u32)OP_END : ops[pc + 1];
#ifdef LB_SCHED_PAIRBAR   // A/B build: warp pairs advance on their own (named barrier per ciphertext slot)
        pair_barrier(1 + slot);
        if (done) break;
#else
        if (__syncthreads_and(done)) break;
#endif
        const u32 ty = done ? (u32)OP_END : (op & OP_MASK), arg = op & ~OP_MASK;
        u32 dg[ND][E];
        if (ty == OP_AUTO || (ty == OP_PERM1 && c == 1))   // EVALUATION-format automorphism of the accumulator
            permute_eval<LOGN, LOGH>(acc, scratch, P.automap + (size_t)arg * N, lane);
        // ONE inlined copy of the digit pipeline (INTT -> decompose -> NTTs -> publish rows 2k+c) for both op kinds, so all
        // warps of the CTA walk the same instruction window: with a separate automorphism pipeline (warp 1 transforming
        // the upper digit polynomial) the loop body was 15.7 k instructions and 27 % of the issue-stall samples were
        // no_instructions (profiles/r1_sched_lmk_v0.txt; the L1.5 instruction cache holds 2 k instructions).
        // External product: both warps.  Automorphism (rgsw-acc-lmkcdey.cpp:257-287): only component 0 is key-switched;
        // warp 1 waits at the barrier and then adds sum_d digit_d * ak[d][1] from warp 0's published rows.
        if (ty == OP_EXT || (ty == OP_AUTO && c == 0))
            digits_from_acc<LOGN, LOGH, ROWS>(acc, dg, P, scratch, sm.twB, sm.twiB, exch, c, lane);
#ifdef LB_SCHED_PAIRBAR
        pair_barrier(1 + slot);
#else
        __syncthreads();
#endif
        if (ty == OP_EXT) {
            mac_replace<LOGN, ROWS>(acc, dg, exch, P, P.bsk + (size_t)arg * kblock, c, lane);
        } else if (ty == OP_AUTO) {
            const u32* ak = P.bsk + P.auto_off + (size_t)arg * ablock;
            if (c == 0) mac_auto_split<LOGN, ND, 0, ND, false>(acc, dg, exch, P, ak, 0, lane);   // own digits, registers
            else        mac_auto_split<LOGN, ND, 0, 0, true>(acc, dg, exch, P, ak, 1, lane);     // warp 0's rows 2d
        }
        if (!done) ++pc;
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
