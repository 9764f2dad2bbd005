// br_wide.cuh — blind rotation for every parameter set the register-resident kernels (br_kernels.cuh, br_sched.cuh) do not
// cover: N = 2048 (STD192*, STD256*, STD128Q_3/4 and their LMKCDEY variants, binfhecontext.cpp:121-158), Q above 2^30
// (34..54 bits: 64-bit residues), the large-precision contexts behind EvalSign / EvalDecomp (Q = LastPrime(54, 2N),
// binfhecontext.cpp:53-104) and gadget decompositions with 1, 4 or 5 digits.
//
// Same algorithm, same op lists (GINX steps / AP and LMKCDEY schedules of br_sched.cuh), other mapping: one CTA per
// ciphertext, the accumulator (2 polynomials) and the digit polynomials (2 (digitsG-1) of them) live in shared memory as
// T = u32 or u64 residues; every transform is a radix-16 register-blocked pass over shared memory (3 passes for N = 2048);
// RGSW key rows stream from L2/HBM in natural EVALUATION order — for u64 that is the canonical host layout itself, so keys
// load with one copy and no repacking.  Pointwise multiply-accumulate sums exact 2B-bit products and reduces once per
// output; every value stored between phases is the canonical residue, like the reference's NativeVector entries.
//   RingGSWAccumulatorCGGI::EvalAcc / AddToAccCGGI      rgsw-acc-cggi.cpp:59-68, :102-151
//   RingGSWAccumulatorDM::EvalAcc / AddToAccDM          rgsw-acc-dm.cpp:62-77, :118-143
//   RingGSWAccumulatorLMKCDEY::EvalAcc / Automorphism   rgsw-acc-lmkcdey.cpp:70-158, :228-287
//   SignedDigitDecompose                                 rgsw-acc.cpp:54-119
//   BootstrapGateCore / BootstrapFuncCore, extraction    binfhe-base-scheme.cpp:525-648, :111-121
#pragma once
#include "br_sched.cuh"

template <typename T> struct WideTraits;
template <> struct WideTraits<u32> {
    typedef u64 Wide;
    typedef uint2 Tw;
    typedef i32 Signed;
    static constexpr int BITS = 32;
};
template <> struct WideTraits<u64> {
    typedef unsigned __int128 Wide;
    typedef ulonglong2 Tw;
    typedef long long Signed;
    static constexpr int BITS = 64;
};

__device__ __forceinline__ u32 w_mulhi(u32 a, u32 b) { return __umulhi(a, b); }
__device__ __forceinline__ u64 w_mulhi(u64 a, u64 b) { return __umul64hi(a, b); }
// Shoup multiplication by a constant w (ws = floor(w 2^B / Q)): any x < 2^B -> [0, 2Q)
template <typename T> __device__ __forceinline__ T w_shoup(T x, T w, T ws, T Q) { return x * w - w_mulhi(x, ws) * Q; }
template <typename T> __device__ __forceinline__ T w_csub(T x, T m) { return x >= m ? x - m : x; }
// 32-bit: x - m wraps above x exactly when x < m, so the conditional subtraction is one VIADDMNMX (15 % of brw_kernel<u32>'s
// instructions were the compare + select form, ncu r2_brw_std256_v0)
template <> __device__ __forceinline__ u32 w_csub<u32>(u32 x, u32 m) { return min(x, x - m); }

template <typename T>
struct BrwParams {
    typedef typename WideTraits<T>::Tw Tw;
    T Q, Q2, qhalf;
    T ninv, ninv_s;        // N^-1 and its Shoup companion
    T mu;                  // floor(2^B / Q)
    T r, rs;               // 2^B mod Q and its Shoup companion: folds the high word of a 2B-bit sum
    T b_add;               // constant added to b after extraction
    const Tw* tw;          // [N] (psi^brv(i), Shoup companion)   transformnat-impl.h:777-831
    const Tw* twi;         // inverse table
    const T* pw;           // [2N] psi^k: the GINX monomial X^a - 1 at evaluation point p is pw[a (2 brv(p) + 1) mod 2N] - 1
    const T* bsk;          // key blocks [rows][2][N], natural EVALUATION order
    const u32* a;          // [count][n] LWE mask mod ctmod
    const u32* b;          // [count]
    const T* lut;          // [ctmod] test-vector table
    const u32* lutsel;     // optional per-ciphertext table offset
    const u32* sched;      // AP / LMKCDEY op lists (br_sched.cuh)
    const u32* automap;    // LMKCDEY: [numAutoKeys+1][N] natural-order source indices, out[p] = in[map[p]]
    T* ext;                // [count][N+1]
    T* acc_io;             // mode 1: [count][2][N] EVALUATION format in/out
    unsigned long long auto_off;   // element offset of the automorphism keys inside bsk
    u32 n, count, ctmod, factor, amul;
    u32 gbits, nd, logN, method, mode, sched_stride;
};

__device__ __forceinline__ u32 wphys(u32 e) { return e + (e >> 4); }   // one pad word per 16: stride-16 accesses hit 16 banks

// 2B-bit value -> canonical residue.  hi * 2^B is folded with a Shoup multiplication by r = 2^B mod Q, lo with a Barrett
// step (mu = floor(2^B / Q) is at most one short): both land in [0, 2Q), the sum in [0, 4Q) (4Q < 2^B).
template <typename T>
__device__ __forceinline__ T w_reduce(typename WideTraits<T>::Wide x, const BrwParams<T>& P) {
    const T lo = (T)x, hi = (T)(x >> WideTraits<T>::BITS);
    const T t = w_shoup(hi, P.r, P.rs, P.Q);
    const T l = lo - w_mulhi(lo, P.mu) * P.Q;
    return w_csub(w_csub(t + l, P.Q2), P.Q);
}

// K consecutive NTT stages on `lines` polynomials held in shared memory (line stride PL, element e at wphys(e)).
// Forward (Cooley-Tukey, transformnat-impl.h:327-351): stages with half-size tlo * 2^(K-1) .. tlo; values stay in [0, 4Q).
// Inverse (Gentleman-Sande, :539-598): half-size tlo .. tlo * 2^(K-1); values stay in [0, 2Q).
// LAST: the group's write-back canonicalises (forward) or applies N^-1 and canonicalises (inverse, :599-623).
template <typename T, bool INVERSE, int K, bool LAST>
__device__ __forceinline__ void w_stage_group(T* tile, u32 lines, u32 PL, u32 logN, u32 tlo, const BrwParams<T>& P) {
    typedef typename WideTraits<T>::Tw Tw;
    constexpr int R = 1 << K;
    const u32 N = 1u << logN;
    const u32 per = N >> K, work = lines * per;
    const u32 ltlo = 31 - __clz(tlo);
    const Tw* __restrict__ tw = INVERSE ? P.twi : P.tw;
    const T Q = P.Q, Q2 = P.Q2;
    for (u32 wi = threadIdx.x; wi < work; wi += blockDim.x) {
        const u32 grp = wi % per, line = wi / per;
        const u32 e0 = (grp >> ltlo) * (tlo * R) + (grp & (tlo - 1));
        T* base = tile + (size_t)line * PL;
        T x[R];
#pragma unroll
        for (int u = 0; u < R; ++u) x[u] = base[wphys(e0 + u * tlo)];
#pragma unroll
        for (int ss = 0; ss < K; ++ss) {
            const int s  = INVERSE ? ss : (K - 1 - ss);
            const u32 sh = ltlo + s + 1;                       // log2(2 t)
            const u32 i0 = (1u << (logN - sh)) + (e0 >> sh);
#pragma unroll
            for (int g = 0; g < (R >> (s + 1)); ++g) {
                const Tw w = __ldg(tw + i0 + g);
#pragma unroll
                for (int v = 0; v < (1 << s); ++v) {
                    const int u = (g << (s + 1)) + v;
                    T& a = x[u];
                    T& b = x[u + (1 << s)];
                    if (INVERSE) {
                        const T d = a - b + Q2;
                        a = w_csub(a + b, Q2);
                        b = w_shoup(d, (T)w.x, (T)w.y, Q);
                    } else {
                        a = w_csub(a, Q2);
                        const T m = w_shoup(b, (T)w.x, (T)w.y, Q);
                        b = a - m + Q2;
                        a = a + m;
                    }
                }
            }
        }
#pragma unroll
        for (int u = 0; u < R; ++u) {
            T v = x[u];
            if (LAST) v = INVERSE ? w_csub(w_shoup(v, P.ninv, P.ninv_s, Q), Q) : w_csub(w_csub(v, Q2), Q);
            base[wphys(e0 + u * tlo)] = v;
        }
    }
    __syncthreads();
}

template <typename T, bool INVERSE, bool LAST>
__device__ __forceinline__ void w_stage_group_k(int K, T* tile, u32 lines, u32 PL, u32 logN, u32 tlo, const BrwParams<T>& P) {
    switch (K) {
        case 4: w_stage_group<T, INVERSE, 4, LAST>(tile, lines, PL, logN, tlo, P); break;
        case 3: w_stage_group<T, INVERSE, 3, LAST>(tile, lines, PL, logN, tlo, P); break;
        case 2: w_stage_group<T, INVERSE, 2, LAST>(tile, lines, PL, logN, tlo, P); break;
        default: w_stage_group<T, INVERSE, 1, LAST>(tile, lines, PL, logN, tlo, P); break;
    }
}

// canonical coefficients -> canonical evaluations (bit-reversed order, as the reference stores them), in place
template <typename T>
__device__ __forceinline__ void w_ntt_fwd(T* tile, u32 lines, u32 PL, u32 logN, const BrwParams<T>& P) {
    u32 done = 0;
    while (done < logN) {
        const u32 K = (logN - done) >= 4 ? 4 : (logN - done);
        const u32 tlo = (1u << logN) >> (done + K);
        if (done + K == logN) w_stage_group_k<T, false, true>((int)K, tile, lines, PL, logN, tlo, P);
        else w_stage_group_k<T, false, false>((int)K, tile, lines, PL, logN, tlo, P);
        done += K;
    }
}
template <typename T>
__device__ __forceinline__ void w_ntt_inv(T* tile, u32 lines, u32 PL, u32 logN, const BrwParams<T>& P) {
    u32 done = 0;
    while (done < logN) {
        const u32 K = (logN - done) >= 4 ? 4 : (logN - done);
        const u32 tlo = 1u << done;
        if (done + K == logN) w_stage_group_k<T, true, true>((int)K, tile, lines, PL, logN, tlo, P);
        else w_stage_group_k<T, true, false>((int)K, tile, lines, PL, logN, tlo, P);
        done += K;
    }
}

// SignedDigitDecompose of one canonical coefficient (rgsw-acc.cpp:62-89): centre, drop the lowest digit, then nd balanced
// base-2^gbits digits, negative digits stored as r + Q.  out advances by `ostride` elements per digit.
template <typename T>
__device__ __forceinline__ void w_decompose(T t, const BrwParams<T>& P, T* out, u32 ostride) {
    typedef typename WideTraits<T>::Signed S;
    const int sh = WideTraits<T>::BITS - (int)P.gbits;
    S d = (t < P.qhalf) ? (S)t : (S)t - (S)P.Q;
    S r = (S)((T)d << sh) >> sh;
    d = (d - r) >> P.gbits;
    for (u32 k = 0; k < P.nd; ++k) {
        r = (S)((T)d << sh) >> sh;
        d = (d - r) >> P.gbits;
        out[(size_t)k * ostride] = r < 0 ? (T)(r + (S)P.Q) : (T)r;
    }
}

// acc (2 lines) -> INTT -> decompose -> NTT of the 2 nd digit polynomials: dig row 2k+c = digit k of component c
template <typename T>
__device__ __forceinline__ void w_digits2(const T* acc, T* dig, u32 PL, const BrwParams<T>& P) {
    const u32 N = 1u << P.logN;
    for (u32 i = threadIdx.x; i < 2 * PL; i += blockDim.x) dig[i] = acc[i];
    __syncthreads();
    w_ntt_inv(dig, 2, PL, P.logN, P);
    for (u32 i = threadIdx.x; i < 2 * N; i += blockDim.x) {
        const u32 c = i >> P.logN, j = i & (N - 1);
        T* cell = dig + (size_t)c * PL + wphys(j);
        w_decompose(*cell, P, cell, 2 * PL);
    }
    __syncthreads();
    w_ntt_fwd(dig, 2 * P.nd, PL, P.logN, P);
}

// (a 64-register cap for u32 — four 52 KB CTAs per SM instead of three — moved the two-digit sets by +3 % and the three-digit ones
// by -2 %: not kept, r2 job 20)
template <typename T, int NT>
__global__ void __launch_bounds__(NT) brw_kernel(const __grid_constant__ BrwParams<T> P) {
    typedef typename WideTraits<T>::Wide Wide;
    extern __shared__ __align__(16) unsigned char brw_smem[];
    const u32 N = 1u << P.logN, PL = N + (N >> 4), rows = 2 * P.nd;
    T* acc = reinterpret_cast<T*>(brw_smem);   // [2][PL]
    T* dig = acc + 2 * PL;                     // [rows][PL]
    const u32 ct = blockIdx.x;
    if (ct >= P.count) return;
    const size_t kblock = (size_t)rows * 2 * N;

    // ---- prologue: acc = (0, NTT(m)), m[j * factor] = lut[(b - j) mod ctmod] for j < ctmod / 2 --------------------------
    if (P.mode == 0) {
        const u32 b = P.b[ct], half = P.ctmod >> 1;
        const T* lut = P.lutsel ? P.lut + P.lutsel[ct] : P.lut;
        for (u32 j = threadIdx.x; j < N; j += blockDim.x) {
            T val = 0;
            if (j % P.factor == 0) {
                const u32 jj = j / P.factor;
                if (jj < half) val = lut[(b + P.ctmod - jj) % P.ctmod];
            }
            acc[wphys(j)]      = 0;
            acc[PL + wphys(j)] = val;
        }
        __syncthreads();
        w_ntt_fwd(acc + PL, 1, PL, P.logN, P);
    } else {
        const T* src = P.acc_io + (size_t)ct * 2 * N;
        for (u32 i = threadIdx.x; i < 2 * N; i += blockDim.x) acc[(size_t)(i >> P.logN) * PL + wphys(i & (N - 1))] = src[i];
        __syncthreads();
    }

    // ---- blind rotation ---------------------------------------------------------------------------------------------------
    if (P.method == M_GINX) {
        const u32* arow = P.a + (size_t)ct * P.n;
        const u32 M = 2 * N;
        for (u32 i = 0; i < P.n; ++i) {
            const u32 ai = arow[i];
            const u32 ap = ((P.ctmod - ai) % P.ctmod) * P.amul;   // (-a_i mod q) (2N / q)
            if (ap == 0) continue;   // X^0 - 1 = 0: the step adds zero polynomials (exact arithmetic: skipping is identical)
            w_digits2(acc, dig, PL, P);
            const T* ek1 = P.bsk + (size_t)(2 * i) * kblock;
            const T* ek2 = ek1 + kblock;
            const u32 an = M - ap;
            for (u32 p = threadIdx.x; p < N; p += blockDim.x) {
                const u32 odd = 2 * (__brev(p) >> (32 - P.logN)) + 1;
                const T m1 = P.pw[(ap * odd) & (M - 1)], m2 = P.pw[(an * odd) & (M - 1)];
                const T mono1 = m1 ? m1 - 1 : P.Q - 1, mono2 = m2 ? m2 - 1 : P.Q - 1;
                Wide s10 = 0, s11 = 0, s20 = 0, s21 = 0;
                for (u32 r = 0; r < rows; ++r) {
                    const T d = dig[(size_t)r * PL + wphys(p)];
                    const size_t ko = ((size_t)r * 2) * N + p;
                    s10 += (Wide)d * __ldg(ek1 + ko);
                    s11 += (Wide)d * __ldg(ek1 + ko + N);
                    s20 += (Wide)d * __ldg(ek2 + ko);
                    s21 += (Wide)d * __ldg(ek2 + ko + N);
                }
                const T r10 = w_reduce<T>(s10, P), r11 = w_reduce<T>(s11, P), r20 = w_reduce<T>(s20, P), r21 = w_reduce<T>(s21, P);
                T& a0 = acc[wphys(p)];
                T& a1 = acc[PL + wphys(p)];
                a0 = w_reduce<T>((Wide)a0 + (Wide)r10 * mono1 + (Wide)r20 * mono2, P);
                a1 = w_reduce<T>((Wide)a1 + (Wide)r11 * mono1 + (Wide)r21 * mono2, P);
            }
            __syncthreads();
        }
    } else {
        const u32* ops = P.sched + (size_t)ct * P.sched_stride;
        for (u32 pc = 0;; ++pc) {
            const u32 op = ops[pc], ty = op & OP_MASK, arg = op & ~OP_MASK;
            if (ty == OP_END) break;
            if (ty == OP_EXT) {   // acc = sum_r dct[r] * ek[r]   (AddToAccDM / AddToAccLMKCDEY)
                w_digits2(acc, dig, PL, P);
                const T* ek = P.bsk + (size_t)arg * kblock;
                for (u32 p = threadIdx.x; p < N; p += blockDim.x) {
                    Wide s0 = 0, s1 = 0;
                    for (u32 r = 0; r < rows; ++r) {
                        const T d = dig[(size_t)r * PL + wphys(p)];
                        const size_t ko = ((size_t)r * 2) * N + p;
                        s0 += (Wide)d * __ldg(ek + ko);
                        s1 += (Wide)d * __ldg(ek + ko + N);
                    }
                    acc[wphys(p)]      = w_reduce<T>(s0, P);
                    acc[PL + wphys(p)] = w_reduce<T>(s1, P);
                }
                __syncthreads();
            } else {
                // EVALUATION-format automorphism (poly-impl.h:345-356): out[p] = in[map[p]]
                const u32* map = P.automap + (size_t)arg * N;
                for (u32 p = threadIdx.x; p < N; p += blockDim.x) dig[PL + wphys(p)] = acc[PL + wphys(map[p])];
                if (ty == OP_AUTO)
                    for (u32 p = threadIdx.x; p < N; p += blockDim.x) dig[wphys(p)] = acc[wphys(map[p])];
                __syncthreads();
                for (u32 p = threadIdx.x; p < N; p += blockDim.x) acc[PL + wphys(p)] = dig[PL + wphys(p)];
                __syncthreads();
                if (ty == OP_AUTO) {   // key switch of the permuted component 0 with the automorphism key (:257-287)
                    w_ntt_inv(dig, 1, PL, P.logN, P);
                    for (u32 j = threadIdx.x; j < N; j += blockDim.x) {
                        T* cell = dig + wphys(j);
                        w_decompose(*cell, P, cell, PL);
                    }
                    __syncthreads();
                    w_ntt_fwd(dig, P.nd, PL, P.logN, P);
                    const T* ak = P.bsk + P.auto_off + (size_t)arg * ((size_t)P.nd * 2 * N);
                    for (u32 p = threadIdx.x; p < N; p += blockDim.x) {
                        Wide s0 = 0, s1 = (Wide)acc[PL + wphys(p)];
                        for (u32 d = 0; d < P.nd; ++d) {
                            const T x = dig[(size_t)d * PL + wphys(p)];
                            const size_t ko = ((size_t)d * 2) * N + p;
                            s0 += (Wide)x * __ldg(ak + ko);
                            s1 += (Wide)x * __ldg(ak + ko + N);
                        }
                        acc[wphys(p)]      = w_reduce<T>(s0, P);
                        acc[PL + wphys(p)] = w_reduce<T>(s1, P);
                    }
                    __syncthreads();
                }
            }
        }
    }

    // ---- epilogue ---------------------------------------------------------------------------------------------------------
    if (P.mode == 0) {
        // a' = INTT(Transpose(acc0)) = coefficient map a'[0] = a[0], a'[N-j] = -a[j]; b = INTT(acc1)[0] + b_add
        w_ntt_inv(acc, 2, PL, P.logN, P);
        T* out = P.ext + (size_t)ct * (N + 1);
        for (u32 j = threadIdx.x; j < N; j += blockDim.x) {
            const T v = acc[wphys(j)];
            if (j == 0) out[0] = v;
            else out[N - j] = v ? P.Q - v : 0;
        }
        if (threadIdx.x == 0) out[N] = w_csub(acc[PL] + P.b_add, P.Q);
    } else {
        T* dst = P.acc_io + (size_t)ct * 2 * N;
        for (u32 i = threadIdx.x; i < 2 * N; i += blockDim.x) dst[i] = acc[(size_t)(i >> P.logN) * PL + wphys(i & (N - 1))];
    }
}
