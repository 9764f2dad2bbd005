// host_math.h — host-side number theory for table generation (one-time, per context / plan).
// Reproduces the *values* the reference derives: LastPrime / PreviousPrime (nbtheory-impl.h:350-393), the
// minimal primitive root RootOfUnity (nbtheory-impl.h:183-231), and the bit-reversed twiddle tables of
// ChineseRemainderTransformFTTNat::PreCompute (transformnat-impl.h:777-831).
#pragma once
#include <cmath>
#include <cstdint>
#include <vector>

namespace lb {
typedef unsigned __int128 u128_t;

inline uint64_t mulmod(uint64_t a, uint64_t b, uint64_t q) { return (uint64_t)((u128_t)a * b % q); }
inline uint64_t addmod(uint64_t a, uint64_t b, uint64_t q) { uint64_t s = a + b; return s >= q ? s - q : s; }
inline uint64_t submod(uint64_t a, uint64_t b, uint64_t q) { return a >= b ? a - b : a + q - b; }
inline uint64_t powmod(uint64_t a, uint64_t e, uint64_t q) {
    uint64_t r = 1 % q;
    a %= q;
    for (; e; e >>= 1) {
        if (e & 1) r = mulmod(r, a, q);
        a = mulmod(a, a, q);
    }
    return r;
}
inline uint64_t invmod_prime(uint64_t a, uint64_t q) { return powmod(a, q - 2, q); }
inline uint32_t ilog2(uint64_t x) { uint32_t r = 0; while (x >>= 1) ++r; return r; }
inline uint32_t bits_of(uint64_t x) { return x ? ilog2(x) + 1 : 0; }
inline uint32_t bitrev(uint32_t x, uint32_t bits) {
    uint32_t r = 0;
    for (uint32_t i = 0; i < bits; ++i) r |= ((x >> i) & 1u) << (bits - 1 - i);
    return r;
}
inline bool is_prime(uint64_t n) {  // deterministic Miller-Rabin, valid for all 64-bit n
    if (n < 2) return false;
    static const uint64_t sp[] = {2, 3, 5, 7, 11, 13, 17, 19, 23, 29, 31, 37};
    for (uint64_t p : sp) {
        if (n == p) return true;
        if (n % p == 0) return false;
    }
    uint64_t d = n - 1;
    int s = 0;
    while (!(d & 1)) { d >>= 1; ++s; }
    for (uint64_t a : sp) {
        uint64_t x = powmod(a, d, n);
        if (x == 1 || x == n - 1) continue;
        bool comp = true;
        for (int r = 1; r < s && comp; ++r) {
            x = mulmod(x, x, n);
            if (x == n - 1) comp = false;
        }
        if (comp) return false;
    }
    return true;
}
inline uint64_t last_prime(uint32_t bits, uint64_t m) {
    uint64_t q = (uint64_t)1 << bits, r = q % m, qn = q + 1 - r;
    if (r < 2) qn -= m;
    while (!is_prime(qn)) qn -= m;
    return qn;
}
inline uint64_t previous_prime(uint64_t q, uint64_t m) {
    uint64_t qn = q - m;
    while (!is_prime(qn)) qn -= m;
    return qn;
}
// minimal primitive m-th root of unity mod prime q, m a power of two dividing q-1; 0 if none
inline uint64_t root_of_unity(uint64_t m, uint64_t q) {
    if (m < 2 || (m & (m - 1)) || (q - 1) % m) return 0;
    uint64_t r = 0;
    for (uint64_t g = 2; g < q; ++g) {
        r = powmod(g, (q - 1) / m, q);
        if (powmod(r, m / 2, q) == q - 1) break;
    }
    uint64_t r2 = mulmod(r, r, q), x = r, mn = r;
    for (uint64_t k = 1; k < m; k += 2) {
        if (x < mn && x != 1) mn = x;
        x = mulmod(x, r2, q);
    }
    return mn;
}
// tw[brv(i)] = psi^i, twi[brv(i)] = psi^-i, i < N
inline void ntt_tables(uint32_t N, uint64_t Q, uint64_t psi, std::vector<uint64_t>& tw, std::vector<uint64_t>& twi) {
    tw.resize(N);
    twi.resize(N);
    uint32_t lg = ilog2(N);
    uint64_t x = 1, xi = 1, psii = invmod_prime(psi, Q);
    for (uint32_t i = 0; i < N; ++i) {
        uint32_t r = bitrev(i, lg);
        tw[r] = x;
        twi[r] = xi;
        x = mulmod(x, psi, Q);
        xi = mulmod(xi, psii, Q);
    }
}
inline uint32_t shoup32(uint64_t w, uint64_t Q) { return (uint32_t)(((u128_t)w << 32) / Q); }
inline uint64_t shoup64(uint64_t w, uint64_t Q) { return (uint64_t)(((u128_t)w << 64) / Q); }
// the reference's digit counts: ceil(log(x)/log(base)) evaluated in double (lwe-pke.cpp:355,
// rgsw-cryptoparameters.h:93-94, rgsw-cryptoparameters.cpp:39-40)
inline uint32_t digit_count(uint64_t x, uint64_t base) {
    return (uint32_t)std::ceil(std::log((double)x) / std::log((double)base));
}
}  // namespace lb
