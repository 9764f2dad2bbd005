// ntt_warp.cuh — one-warp-per-polynomial negacyclic NTT/INTT for N = 2^LOGN (256 <= N <= 2048), u32 lazy.
//
// Computes exactly what NumberTheoreticTransformNat::ForwardTransformToBitReverseInPlace
// (transformnat-impl.h:302-373) and ::InverseTransformFromBitReverseInPlace (:511-624) compute:
// Cooley-Tukey, natural order in -> bit-reversed out with twiddle Table[brv(i)] = psi^i (:777-831), and
// Gentleman-Sande bit-reversed in -> natural out with the N^-1 scaling folded into the last stage.
//
// B200 mapping: each lane keeps E = N/32 coefficients in registers.
//   layout A: reg r of lane l holds index j = l + 32 r   (strides >= 32 are in-thread)
//   layout B: reg r of lane l holds index p = E l + r    (strides <  E  are in-thread)
// A 2^LOGN transform is: phase A (in-register, twiddles uniform across the warp -> read from the kernel
// parameter constant bank), [shuffle stages when E < 32], ONE swizzled shared-memory transpose, phase B
// (in-register, per-lane twiddles from a lane-major shared-memory table).  So a whole NTT costs one
// shared-memory round trip of the data instead of log2(N).
#pragma once
#include "modarith.cuh"

// uniform (warp-wide) constants, passed by value as a __grid_constant__ kernel parameter
struct NttConsts {
    uint2 twA[64];   // forward  table entries 1..E-1 : (w, floor(w 2^32/Q))
    uint2 twiA[64];  // inverse  table entries 1..E-1
    u32 Q, Q2, mu32;
    u32 ninv, ninv_s;      // N^-1 and its Shoup companion
    u32 w1ninv, w1ninv_s;  // TableI[1] * N^-1 (transformnat-impl.h:537-538)
    u32 pad;
    u32 QM[8];             // Q << k: read as constant-bank operands (IADD3 / VIADDMNMX take c[][] directly), so the lazy
                           // offsets cost no IMAD/LEA to rematerialise and no register to keep
};
__host__ __device__ constexpr int ntt_ilog2(int v) { return v <= 1 ? 0 : 1 + ntt_ilog2(v >> 1); }

// x < bound*Q -> [0, 2Q) by a ladder of conditional subtractions (VIADDMNMX on the ALU pipe) instead of a Barrett
// step (IMAD.HI + IMAD on the IMAD pipe, which is the pipe these kernels saturate: ncu r1, fmaheavy 69 %).
// `bound` is a compile-time constant after unrolling, so the ladder folds to ceil(log2(bound)) - 1 instructions.
__device__ __forceinline__ u32 ladder_2q(u32 x, int bound, const NttConsts& C) {
#pragma unroll
    for (int k = 4; k >= 0; --k)
        if ((2 << k) < bound) x = csub(x, C.QM[k + 1]);
    return x;
}
// x < from*Q -> x < to*Q (to a power of two): ceil(log2(from/to)) conditional subtractions
__device__ __forceinline__ u32 ladder_to(u32 x, int from, int to, const NttConsts& C) {
#pragma unroll
    for (int k = 5; k >= 0; --k)
        if ((to << k) < from) x = csub(x, C.QM[ntt_ilog2(to) + k]);
    return x;
}
__host__ __device__ constexpr int ntt_clog2(int v) { return ntt_ilog2(v) + (((1 << ntt_ilog2(v)) < v) ? 1 : 0); }

template <int LOGN, int LOGH>
struct WNtt {
    static constexpr int N    = 1 << LOGN;
    static constexpr int E    = N / 32;
    static constexpr int LOGE = LOGN - 5;
    static constexpr int TB   = (E / 2 < 16) ? E / 2 : 16;  // largest in-thread stride of phase B
    static constexpr int LOGTB = (TB == 16) ? 4 : (TB == 8) ? 3 : (TB == 4) ? 2 : 1;
    static constexpr int H    = 1 << LOGH;                  // Q < 2^(32-LOGH): values < H*Q fit in 32 bits

    __host__ __device__ static constexpr int offB(int t) {  // first entry of stage t in the lane-major table
        int o = 0;
        for (int tt = TB; tt > t; tt >>= 1) o += E / (2 * tt);
        return o;
    }
    static constexpr int NB = offB(1) + E / 2;  // phase-B twiddles per lane

    // ---- A <-> B transpose through a per-warp N-word scratch ------------------------------------------------------
    // The scratch is organised by layout-B rows (lane l owns words E*l .. E*l+E-1 = E/4 16-byte chunks); chunk v of row
    // l is stored at chunk v ^ SB(l).  The layout-B side then moves whole chunks (STS.128 / LDS.128, conflict-free
    // because 8 consecutive lanes get 8 different bank groups), and on the layout-A side index lane + 32r lives at word
    // 32r + (lane ^ 4*SA(r)): a permutation of one 128-byte line per r, so it is conflict-free too and its address is
    // one of at most 8 precomputed per-lane pointers plus the immediate 128*r — no per-access address arithmetic.
    // (tools/mix_probe.cu: the scalar-both-sides transpose cost 25 % of the IMAD pipe, this one 12 %.)
    static constexpr int NSW = (E >= 32) ? 8 : (E == 16) ? 4 : 2;                 // distinct swizzle values
    __host__ __device__ static constexpr int SA(int r) { return (E >= 64) ? ((r >> 1) & 7) : (r & (NSW - 1)); }
    __device__ static __forceinline__ u32 SB(u32 l) { return (E >= 32) ? (l & 7u) : (E == 16) ? ((l >> 1) & 3u) : ((l >> 2) & 1u); }
    // word address of index j (generic; used by gathers such as the EVALUATION-format automorphism)
    __device__ static __forceinline__ u32 phys(u32 j) {
        const u32 l = j >> LOGE, rr = j & (E - 1);
        return (l << LOGE) + ((((rr >> 2) ^ SB(l)) << 2) | (rr & 3u));
    }
    __device__ static __forceinline__ void store_A(u32* sm, const u32 (&x)[E], int lane) {
        u32* pk[NSW];
#pragma unroll
        for (int k = 0; k < NSW; ++k) pk[k] = sm + (lane ^ (4 * k));
#pragma unroll
        for (int r = 0; r < E; ++r) pk[SA(r)][32 * r] = x[r];
    }
    __device__ static __forceinline__ void load_A(const u32* sm, u32 (&x)[E], int lane) {
        const u32* pk[NSW];
#pragma unroll
        for (int k = 0; k < NSW; ++k) pk[k] = sm + (lane ^ (4 * k));
#pragma unroll
        for (int r = 0; r < E; ++r) x[r] = pk[SA(r)][32 * r];
    }
    __device__ static __forceinline__ void store_B(u32* sm, const u32 (&x)[E], int lane) {
        uint4* row   = reinterpret_cast<uint4*>(sm + E * lane);
        const u32 sb = SB((u32)lane);
#pragma unroll
        for (int v = 0; v < E / 4; ++v) row[v ^ sb] = make_uint4(x[4 * v], x[4 * v + 1], x[4 * v + 2], x[4 * v + 3]);
    }
    __device__ static __forceinline__ void load_B(const u32* sm, u32 (&x)[E], int lane) {
        const uint4* row = reinterpret_cast<const uint4*>(sm + E * lane);
        const u32 sb     = SB((u32)lane);
#pragma unroll
        for (int v = 0; v < E / 4; ++v) {
            const uint4 t = row[v ^ sb];
            x[4 * v] = t.x; x[4 * v + 1] = t.y; x[4 * v + 2] = t.z; x[4 * v + 3] = t.w;
        }
    }

    __device__ static __forceinline__ void ct(u32& a, u32& b, u32 w, u32 ws, u32 Q, u32 Q2) {
        u32 u = shoup_mul(b, w, ws, Q);
        b     = a - u + Q2;
        a     = a + u;
    }

    // ---- forward: layout A in (values < B0*Q), layout B out in [0, 2Q) ------------------------------------
    // sm: N u32 of scratch private to this warp; twB: lane-major phase-B table (shared memory);
    // twfull: full (w, ws) table [N] in global memory, used only by the shuffle stages (E < 32).
    template <int B0>
    __device__ static __forceinline__ void forward(u32 (&x)[E], const NttConsts& C, u32* sm, const uint2* twB,
                                                   const uint2* __restrict__ twfull, int lane) {
        const u32 Q = C.Q, Q2 = C.Q2;
        int bound = B0;
#pragma unroll
        for (int s = LOGE - 1; s >= 0; --s) {  // in-register stride tr = 2^s  (global half-size t = 32 tr)
            if (bound + 2 > H) {
#pragma unroll
                for (int r = 0; r < E; ++r) x[r] = ladder_2q(x[r], bound, C);
                bound = 2;
            }
#pragma unroll
            for (int qq = 0; qq < E / 2; ++qq) {
                const int tr = 1 << s, i = qq >> s, r = (i << (s + 1)) | (qq & (tr - 1));
                const uint2 w = C.twA[(E >> (s + 1)) + i];
                ct(x[r], x[r + tr], w.x, w.y, Q, Q2);
            }
            bound += 2;
        }
        if (E < 32) {
#pragma unroll
            for (int t = 16; t >= E; t >>= 1) {
                if (bound + 2 > H) {
#pragma unroll
                    for (int r = 0; r < E; ++r) x[r] = ladder_2q(x[r], bound, C);
                    bound = 2;
                }
                const bool hi = lane & t;
#pragma unroll
                for (int r = 0; r < E; ++r) {
                    const uint2 w = twfull[N / (2 * t) + (32 * r + lane) / (2 * t)];
                    u32 u         = shoup_mul(x[r], w.x, w.y, Q);
                    u32 v         = __shfl_xor_sync(0xffffffffu, hi ? u : x[r], t);
                    x[r]          = hi ? (v - u + Q2) : (x[r] + v);
                }
                bound += 2;
            }
        }
        __syncwarp();
        store_A(sm, x, lane);
        __syncwarp();
        load_B(sm, x, lane);
#pragma unroll
        for (int s = LOGTB; s >= 0; --s) {  // in-register stride t = 2^s
            if (bound + 2 > H) {
#pragma unroll
                for (int r = 0; r < E; ++r) x[r] = ladder_2q(x[r], bound, C);
                bound = 2;
            }
            if (s == 0 && bound > 2) {
                // last stage: bring only the sum operand to [0, 2Q); both results are then < 4Q, and the final
                // conditional subtraction below makes them < 2Q (half as many ladders as reducing every output)
#pragma unroll
                for (int qq = 0; qq < E / 2; ++qq) x[2 * qq] = ladder_2q(x[2 * qq], bound, C);
                bound = 2;
            }
#pragma unroll
            for (int qq = 0; qq < E / 2; ++qq) {
                const int t = 1 << s, i = qq >> s, r = (i << (s + 1)) | (qq & (t - 1));
                const uint2 w = twB[(offB(t) + i) * 32 + lane];
                ct(x[r], x[r + t], w.x, w.y, Q, Q2);
            }
            bound += 2;
        }
        if (bound > 2) {
#pragma unroll
            for (int r = 0; r < E; ++r) x[r] = ladder_2q(x[r], bound, C);
        }
    }

    // ---- forward of NP polynomials at once (same contract as forward<B0> for each) ---------------------------------
    // The NP transforms advance butterfly by butterfly together: every twiddle is fetched once for all of them and each
    // dependent chain has NP-1 independent ones to hide behind (3 warps per scheduler leave ~20 % of issue slots waiting
    // on fixed latencies otherwise).  sm[p]: N u32 of scratch per polynomial, private to this warp.
    template <int B0, int NP>
    __device__ static __forceinline__ void forward_multi(u32 (&x)[NP][E], const NttConsts& C, u32* const (&sm)[NP],
                                                         const uint2* twB, const uint2* __restrict__ twfull, int lane) {
        const u32 Q = C.Q, Q2 = C.Q2;
        int bound = B0;
#pragma unroll
        for (int s = LOGE - 1; s >= 0; --s) {
            if (bound + 2 > H) {
#pragma unroll
                for (int p = 0; p < NP; ++p)
#pragma unroll
                    for (int r = 0; r < E; ++r) x[p][r] = ladder_2q(x[p][r], bound, C);
                bound = 2;
            }
#pragma unroll
            for (int qq = 0; qq < E / 2; ++qq) {
                const int tr = 1 << s, i = qq >> s, r = (i << (s + 1)) | (qq & (tr - 1));
                const uint2 w = C.twA[(E >> (s + 1)) + i];
#pragma unroll
                for (int p = 0; p < NP; ++p) ct(x[p][r], x[p][r + tr], w.x, w.y, Q, Q2);
            }
            bound += 2;
        }
        if (E < 32) {
#pragma unroll
            for (int t = 16; t >= E; t >>= 1) {
                if (bound + 2 > H) {
#pragma unroll
                    for (int p = 0; p < NP; ++p)
#pragma unroll
                        for (int r = 0; r < E; ++r) x[p][r] = ladder_2q(x[p][r], bound, C);
                    bound = 2;
                }
                const bool hi = lane & t;
#pragma unroll
                for (int r = 0; r < E; ++r) {
                    const uint2 w = twfull[N / (2 * t) + (32 * r + lane) / (2 * t)];
#pragma unroll
                    for (int p = 0; p < NP; ++p) {
                        u32 u   = shoup_mul(x[p][r], w.x, w.y, Q);
                        u32 v   = __shfl_xor_sync(0xffffffffu, hi ? u : x[p][r], t);
                        x[p][r] = hi ? (v - u + Q2) : (x[p][r] + v);
                    }
                }
                bound += 2;
            }
        }
        __syncwarp();
#pragma unroll
        for (int p = 0; p < NP; ++p) store_A(sm[p], x[p], lane);
        __syncwarp();
#pragma unroll
        for (int p = 0; p < NP; ++p) load_B(sm[p], x[p], lane);
#pragma unroll
        for (int s = LOGTB; s >= 0; --s) {
            if (bound + 2 > H) {
#pragma unroll
                for (int p = 0; p < NP; ++p)
#pragma unroll
                    for (int r = 0; r < E; ++r) x[p][r] = ladder_2q(x[p][r], bound, C);
                bound = 2;
            }
            if (s == 0 && bound > 2) {
#pragma unroll
                for (int p = 0; p < NP; ++p)
#pragma unroll
                    for (int qq = 0; qq < E / 2; ++qq) x[p][2 * qq] = ladder_2q(x[p][2 * qq], bound, C);
                bound = 2;
            }
#pragma unroll
            for (int qq = 0; qq < E / 2; ++qq) {
                const int t = 1 << s, i = qq >> s, r = (i << (s + 1)) | (qq & (t - 1));
                const uint2 w = twB[(offB(t) + i) * 32 + lane];
#pragma unroll
                for (int p = 0; p < NP; ++p) ct(x[p][r], x[p][r + t], w.x, w.y, Q, Q2);
            }
            bound += 2;
        }
        if (bound > 2) {
#pragma unroll
            for (int p = 0; p < NP; ++p)
#pragma unroll
                for (int r = 0; r < E; ++r) x[p][r] = ladder_2q(x[p][r], bound, C);
        }
    }

    // ---- inverse: layout B in (values in [0, 2Q)), layout A out, canonical [0, Q) --------------------------
    // Gentleman-Sande: a' = a + b doubles the bound of the sum path, b' = (a - b + kQ) * w comes back < 2Q.  Bounds are
    // tracked PER REGISTER at compile time: after j stages only the registers whose low j index bits are 0 carry the
    // maximal bound, half of them are < 2Q, a quarter < 4Q, ... so only the few registers that would overflow are
    // laddered (37 conditional subtractions per N=1024 transform at H = 32 instead of 128 with one bound per stage).
    __device__ static __forceinline__ void gs_pair(u32& a, u32& b, int& ba, int& bb, u32 w, u32 ws, const NttConsts& C) {
        if (ba + bb > H) {   // the sum must stay below H*Q < 2^32
            a = ladder_to(a, ba, H / 2, C); ba = ba < H / 2 ? ba : H / 2;
            b = ladder_to(b, bb, H / 2, C); bb = bb < H / 2 ? bb : H / 2;
        }
        const u32 d = a - b + C.QM[ntt_clog2(bb)];
        a  = a + b;
        ba = ba + bb;
        b  = shoup_mul(d, w, ws, C.Q);
        bb = 2;
    }
    __device__ static __forceinline__ void inverse(u32 (&x)[E], const NttConsts& C, u32* sm, const uint2* twiB,
                                                   const uint2* __restrict__ twifull, int lane) {
        const u32 Q = C.Q;
        int bnd[E];
#pragma unroll
        for (int r = 0; r < E; ++r) bnd[r] = 2;
#pragma unroll
        for (int s = 0; s <= LOGTB; ++s) {
#pragma unroll
            for (int qq = 0; qq < E / 2; ++qq) {
                const int t = 1 << s, i = qq >> s, r = (i << (s + 1)) | (qq & (t - 1));
                const uint2 w = twiB[(offB(t) + i) * 32 + lane];
                gs_pair(x[r], x[r + t], bnd[r], bnd[r + t], w.x, w.y, C);
            }
        }
        // after the transpose a register's history depends on the lane, so phase A starts from one bound: [0, 2Q)
#pragma unroll
        for (int r = 0; r < E; ++r) {
            x[r]   = ladder_to(x[r], bnd[r], 2, C);
            bnd[r] = 2;
        }
        __syncwarp();
        store_B(sm, x, lane);
        __syncwarp();
        load_A(sm, x, lane);
        if (E < 32) {
            int bound = 2;
#pragma unroll
            for (int t = E; t <= 16; t <<= 1) {
                const bool red = 4 * bound > H;
                const bool hi  = lane & t;
#pragma unroll
                for (int r = 0; r < E; ++r) {
                    const uint2 w = twifull[N / (2 * t) + (32 * r + lane) / (2 * t)];
                    u32 v         = __shfl_xor_sync(0xffffffffu, x[r], t);
                    u32 s         = x[r] + v;
                    if (red) s = ladder_2q(s, 2 * bound, C);
                    u32 d = shoup_mul(v - x[r] + C.QM[ntt_ilog2(bound)], w.x, w.y, Q);
                    x[r]  = hi ? d : s;
                }
                bound = red ? 2 : 2 * bound;
            }
#pragma unroll
            for (int r = 0; r < E; ++r) bnd[r] = bound;
        }
#pragma unroll
        for (int s = 0; s < LOGE - 1; ++s) {
#pragma unroll
            for (int qq = 0; qq < E / 2; ++qq) {
                const int tr = 1 << s, i = qq >> s, r = (i << (s + 1)) | (qq & (tr - 1));
                const uint2 w = C.twiA[(E >> (s + 1)) + i];
                gs_pair(x[r], x[r + tr], bnd[r], bnd[r + tr], w.x, w.y, C);
            }
        }
        // last stage (t = N/2) with N^-1 folded in; outputs canonical
        {
#pragma unroll
            for (int r = 0; r < E / 2; ++r) {
                u32 a = x[r], b = x[r + E / 2];
                int ba = bnd[r], bb = bnd[r + E / 2];
                if (ba + bb > H) {
                    a = ladder_to(a, ba, H / 2, C); ba = ba < H / 2 ? ba : H / 2;
                    b = ladder_to(b, bb, H / 2, C); bb = bb < H / 2 ? bb : H / 2;
                }
                x[r]         = csub(shoup_mul(a + b, C.ninv, C.ninv_s, Q), Q);
                x[r + E / 2] = csub(shoup_mul(a - b + C.QM[ntt_clog2(bb)], C.w1ninv, C.w1ninv_s, Q), Q);
            }
        }
    }
};
