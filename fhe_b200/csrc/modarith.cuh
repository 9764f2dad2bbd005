// modarith.cuh — 32-bit lazy modular arithmetic for sm_100a (Q < 2^30).
//
// Every reference operation on this path (ModMulFastConst ubintnat.h:1464-1469, ModMulFast Barrett
// :1348-1361, ModAddFast/ModSubFast) returns the canonical residue in [0,Q); any exact modular algorithm
// therefore reproduces the reference bit for bit as long as the value finally stored is canonical.
// Inside kernels values are kept "lazy" in [0, k*Q) with k tracked at compile time.
#pragma once
#include <cstdint>
#include <cuda_runtime.h>

typedef uint32_t u32;
typedef uint64_t u64;
typedef int32_t i32;

// Shoup/Harvey multiplication by a constant w with ws = floor(w * 2^32 / Q):
// valid for ANY x < 2^32, result in [0, 2Q).  1 IMAD.HI + 2 IMAD.
__device__ __forceinline__ u32 shoup_mul(u32 x, u32 w, u32 ws, u32 Q) {
    u32 q = __umulhi(x, ws);
    return x * w - q * Q;
}
// x < 2^32  ->  [0, 2Q)   (mu32 = floor(2^32 / Q))
__device__ __forceinline__ u32 reduce_2q(u32 x, u32 mu32, u32 Q) {
    u32 q = __umulhi(x, mu32);
    return x - q * Q;
}
// conditional subtract: x in [0, 2m) -> [0, m)
__device__ __forceinline__ u32 csub(u32 x, u32 m) {
    return min(x, x - m);
}
// 64-bit -> [0, 4Q): x < 2^(s+32), mu = floor(2^(s+32) / Q) < 2^32, s <= bits(Q)-1
__device__ __forceinline__ u32 reduce64(u64 x, u32 s, u32 mu, u32 Q) {
    u32 xh = (u32)(x >> s);
    u32 q  = __umulhi(xh, mu);
    return (u32)x - q * Q;
}
// The same with mu = floor(2^(s+32) / Q) above 32 bits (2^s > Q): only the low 32 bits of the quotient estimate enter
// x - q Q mod 2^32, so the high part of mu is one more multiply-add.  The estimate is short by less than
// x / 2^(s+32) + 2^s / Q + 1, so the result stays below 4Q only while that sum is at most 4: the host checks it per context
// (29-bit Q at N = 2048: 3.0 with two digits, 3.75 with three; a 30-bit Q gives 4.0000 and takes reduce64w instead).
__device__ __forceinline__ u32 reduce64x(u64 x, u32 s, u32 mu, u32 mu_hi, u32 Q) {
    u32 xh = (u32)(x >> s);
    u32 q  = __umulhi(xh, mu) + xh * mu_hi;
    return (u32)x - q * Q;
}
// Any x < 2^64 with Q < 2^30 -> [0, 4Q), for the moduli where 2^s would exceed Q and the truncated Barrett step above loses its
// error bound (28- to 30-bit Q at N = 2048): the high word is folded with a Shoup multiplication by R = 2^32 mod Q ([0, 2Q),
// exact for any 32-bit operand), the low word takes a 32-bit Barrett step with mu32 = floor(2^32 / Q) ([0, 2Q)).
__device__ __forceinline__ u32 reduce64w(u64 x, u32 R, u32 Rs, u32 mu32, u32 Q) {
    const u32 hi = (u32)(x >> 32), lo = (u32)x;
    const u32 t = hi * R - __umulhi(hi, Rs) * Q;
    const u32 l = lo - __umulhi(lo, mu32) * Q;
    return t + l;
}
__device__ __forceinline__ u64 mad_wide(u32 a, u32 b, u64 c) {
    return (u64)a * b + c;
}
// sign-extend the low `bits` bits (1 <= bits <= 31): szext.wrap lowers to one SGXT.W; bfe.s32 with a register length
// made ptxas re-derive the clamped length with a PRMT per use (3.4 % of the blind-rotation instructions)
__device__ __forceinline__ i32 sext(i32 v, int bits) {
    i32 r;
    asm("szext.wrap.s32 %0, %1, %2;" : "=r"(r) : "r"(v), "r"(bits));
    return r;
}

// RoundqQ (lwe-pke.cpp:41-46): floor(0.5 + double(v) * double(q) / double(Q)) mod q, IEEE double,
// round-to-nearest, no FMA contraction — the same operation sequence as the x86-64 reference build.
__device__ __forceinline__ u64 round_qQ(u64 v, u64 q, u64 Q) {
    double prod = __dmul_rn((double)v, (double)q);
    double quo  = __ddiv_rn(prod, (double)Q);
    double r    = floor(__dadd_rn(0.5, quo));
    return (u64)r % q;
}

// 64-bit modular helpers for the u64 NTT path (Q < 2^62)
__device__ __forceinline__ u64 shoup_mul64(u64 x, u64 w, u64 ws, u64 Q) {
    u64 q = __umul64hi(x, ws);
    return x * w - q * Q;  // [0, 2Q)
}
__device__ __forceinline__ u64 csub64(u64 x, u64 m) {
    return x >= m ? x - m : x;
}
