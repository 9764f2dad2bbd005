// br_common.cuh — shared pieces of the fused blind-rotation kernels (GINX / AP / LMKCDEY).
//
// One ciphertext = one warp pair: warp c (c = 0,1) owns component c of the RLWE accumulator
// (RLWECiphertextImpl, rlwe-ciphertext.h) for the whole blind rotation; the accumulator never leaves
// registers between steps.  A step is the reference's AddToAcc* (rgsw-acc-cggi.cpp:102-151,
// rgsw-acc-dm.cpp:118-143, rgsw-acc-lmkcdey.cpp:228-254):
//   INTT(acc[c]) -> SignedDigitDecompose (rgsw-acc.cpp:54-91) -> NTT of the digit polynomials ->
//   multiply-accumulate against the RGSW key rows -> accumulator update.
#pragma once
#include "ntt_warp.cuh"

struct BrParams {
    NttConsts C;
    const uint2* twB;      // lane-major phase-B twiddles, forward   [NB][32]
    const uint2* twiB;     // inverse
    const uint2* twfull;   // full (w, ws) tables [N] (shuffle stages, N < 1024 only)
    const uint2* twifull;
    const u32* bsk;        // device BSK, u32, per key block [rows][2][sigma(N)]
    const u32* mono;       // GINX: [2N][sigma(N)]  (X^a - 1 in EVALUATION format)
    const u32* a;          // [count][n]  LWE mask mod ctmod
    const u32* b;          // [count]
    const u32* lut;        // [ctmod] test-vector table: m[j*factor] = lut[(b - j) mod ctmod]
    const u32* lutsel;     // optional [count]: element offset of ciphertext ct's table inside lut (gate circuits mix gate kinds)
    const u32* sched;      // AP/LMKCDEY: per-ciphertext op list [count][sched_stride]
    const u32* automap;    // LMKCDEY: [numAutoKeys+1][sigma(N)] source indices of the EVALUATION-format automorphisms
    u32* ext;              // [count][N+1] : a' (N coefficients mod Q) then b
    u32* acc_io;           // optional [count][2][N] EVALUATION-format accumulator in/out (BlindRotateBatch)
    u32 n, count, ctmod, factor;
    u32 gbits, qhalf;
    u32 red_s, red_mu;     // reduce64 constants
    u32 b_add;             // constant added to b after extraction (gates: Q/(2p)+1, func: 0)
    u32 sched_stride;
    u32 mode;              // 0: lut prologue + extract epilogue, 1: acc_io in/out (raw EvalAcc)
    u32 amul;              // GINX: 2N / ctmod
    u32 auto_off;          // LMKCDEY: u32 offset of the automorphism keys inside bsk
    u32 red_mu_hi;         // bits 32.. of the reduce64 constant (reduce64x, N = 2048)
    u32 red_r, red_rs;     // 2^32 mod Q and its Shoup companion (reduce64w, N = 2048 with a 30-bit Q)
};

// sigma: storage position of EVALUATION index p = E*l + r, chosen so that a lane's 4 consecutive
// registers are one 16-byte vector and a warp's vector load is 512 contiguous bytes.
template <int E>
__device__ __forceinline__ u32 sigma_vec(int v, int lane) {  // in uint4 units
    return v * 32 + lane;
}

// SignedDigitDecompose of one canonical coefficient (rgsw-acc.cpp:62-89): centre, drop the lowest digit,
// emit ND balanced base-2^gbits digits, negative digits stored as r + Q.
template <int ND>
__device__ __forceinline__ void decompose_coeff(u32 t, u32 Q, u32 qhalf, int gbits, u32 (&out)[ND]) {
    i32 d  = (t < qhalf) ? (i32)t : (i32)(t - Q);
    i32 r0 = sext(d, gbits);
    d      = (d - r0) >> gbits;
#pragma unroll
    for (int k = 0; k < ND; ++k) {
        r0 = sext(d, gbits);
        d  = (d - r0) >> gbits;
        out[k] = min((u32)r0, (u32)r0 + Q);   // r0 < 0 ? r0 + Q : r0 as one VIADDMNMX.U32 (a negative r0 is > r0 + Q unsigned)
    }
}

// 64-bit sum -> [0, 4Q).  RED 0: reduce64 (2^s < Q: every N <= 1024 set); 1: reduce64x (N = 2048, Q up to 29 bits, bound checked
// on the host); 2: reduce64w (30-bit Q)
template <int RED>
__device__ __forceinline__ u32 red64(u64 x, const BrParams& P, u32 Q) {
    if (RED == 2) return reduce64w(x, P.red_r, P.red_rs, P.C.mu32, Q);
    if (RED == 1) return reduce64x(x, P.red_s, P.red_mu, P.red_mu_hi, Q);
    return reduce64(x, P.red_s, P.red_mu, Q);
}

__device__ __forceinline__ void pair_barrier(int id) {
    asm volatile("bar.sync %0, 64;" ::"r"(id) : "memory");
}

// ---- bulk async copy (TMA engine, no tensor map) + mbarrier: stages a step's RGSW key block in shared memory -------
namespace brtma {
__device__ __forceinline__ u32 saddr(const void* p) { return (u32)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void mbar_init(u64* bar, u32 cnt) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(saddr(bar)), "r"(cnt) : "memory");
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
}
__device__ __forceinline__ void bulk_load(void* dst, const void* src, u32 bytes, u64* bar) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(saddr(bar)), "r"(bytes) : "memory");
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                 ::"r"(saddr(dst)), "l"(src), "r"(bytes), "r"(saddr(bar)) : "memory");
}
__device__ __forceinline__ void mbar_wait(u64* bar, u32 parity) {
    u32 ok;
    const u32 a = saddr(bar);
    do {
        asm volatile("{\n.reg .pred p;\nmbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\nselp.u32 %0, 1, 0, p;\n}"
                     : "=r"(ok) : "r"(a), "r"(parity) : "memory");
    } while (!ok);
}
// generic-proxy reads of a buffer (ordered by the preceding barrier) before the async proxy overwrites it
__device__ __forceinline__ void fence_async() { asm volatile("fence.proxy.async.shared::cta;" ::: "memory"); }
}  // namespace brtma
