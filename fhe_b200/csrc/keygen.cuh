// keygen.cuh — bootstrapping-key generation on the device: the NTT-dominated part of
// BinFHEContext::BTKeyGen (binfhecontext.cpp:285-307 -> BinFHEScheme::KeyGen, binfhe-base-scheme.cpp:45-72):
//   LWEEncryptionScheme::KeySwitchGen          lwe-pke.cpp:264-344
//   RingGSWAccumulatorCGGI::KeyGenAcc / KeyGenCGGI        rgsw-acc-cggi.cpp:39-96
//   RingGSWAccumulatorDM::KeyGenAcc / KeyGenDM            rgsw-acc-dm.cpp:39-114
//   RingGSWAccumulatorLMKCDEY::KeyGenAcc / KeyGenLMKCDEY / KeyGenAuto   rgsw-acc-lmkcdey.cpp:39-68, :163-223
// written straight into the packed device layouts the blind-rotation and key-switch kernels read.
//
// Randomness: the reference draws from a Blake2-based PRNG (blake2engine.cpp), so device-generated keys cannot be bit-equal
// to reference-generated ones; parity here is FUNCTIONAL and DISTRIBUTIONAL (tests: the CPU oracle evaluates gates correctly
// on device-generated keys, the device on its own keys, and the key noise has the reference's standard deviation 3.19).
// Generator: ChaCha20 (RFC 8439 block function) keyed by the 256-bit expansion of the caller's seed, one stream (nonce) per
// sampled object; uniform residues by Lemire rejection (exactly uniform), errors by inversion of a 64-bit cumulative table of
// the discrete Gaussian.  One deliberate difference: every RGSW row gets its OWN uniform mask polynomial; the reference's
// KeyGenCGGI / KeyGenDM / KeyGenLMKCDEY copy one mask to all rows of a ciphertext (`std::vector<NativePoly> tempA(digitsG2,
// NativePoly(dug, ...))`, rgsw-acc-cggi.cpp:83), which makes rows subtractable — not something to reproduce.
#pragma once
#include "br_wide.cuh"

struct ChaCha {
    u32 key[8];
};

__device__ __forceinline__ u32 kg_rotl(u32 v, int c) { return (v << c) | (v >> (32 - c)); }
#define KG_QR(a, b, c, d)                                   \
    a += b; d ^= a; d = kg_rotl(d, 16); c += d; b ^= c; b = kg_rotl(b, 12); \
    a += b; d ^= a; d = kg_rotl(d, 8);  c += d; b ^= c; b = kg_rotl(b, 7);

// one 64-byte ChaCha20 block -> 8 u64
__device__ __forceinline__ void chacha_block(const ChaCha& K, u64 stream, u64 counter, u64 (&out)[8]) {
    u32 s[16] = {0x61707865u, 0x3320646eu, 0x79622d32u, 0x6b206574u, K.key[0], K.key[1], K.key[2], K.key[3],
                 K.key[4], K.key[5], K.key[6], K.key[7], (u32)counter, (u32)(counter >> 32), (u32)stream, (u32)(stream >> 32)};
    u32 x[16];
#pragma unroll
    for (int i = 0; i < 16; ++i) x[i] = s[i];
#pragma unroll 1
    for (int r = 0; r < 10; ++r) {
        KG_QR(x[0], x[4], x[8], x[12]) KG_QR(x[1], x[5], x[9], x[13]) KG_QR(x[2], x[6], x[10], x[14]) KG_QR(x[3], x[7], x[11], x[15])
        KG_QR(x[0], x[5], x[10], x[15]) KG_QR(x[1], x[6], x[11], x[12]) KG_QR(x[2], x[7], x[8], x[13]) KG_QR(x[3], x[4], x[9], x[14])
    }
#pragma unroll
    for (int i = 0; i < 8; ++i) out[i] = (u64)(x[2 * i] + s[2 * i]) | ((u64)(x[2 * i + 1] + s[2 * i + 1]) << 32);
}

// sequential 64-bit draws from one stream
struct KgStream {
    const ChaCha* K;
    u64 stream, counter;
    u64 buf[8];
    int have;
    __device__ __forceinline__ KgStream(const ChaCha* k, u64 st) : K(k), stream(st), counter(0), have(0) {}
    __device__ __forceinline__ u64 next() {
        if (have == 0) { chacha_block(*K, stream, counter++, buf); have = 8; }
        return buf[--have];
    }
    // exactly uniform in [0, m): Lemire's multiply-and-reject
    __device__ __forceinline__ u64 uniform(u64 m) {
        for (;;) {
            const u64 r = next();
            const u64 hi = __umul64hi(r, m), lo = r * m;
            if (lo >= m) return hi;
            if (lo >= (0 - m) % m) return hi;
        }
    }
};

// discrete Gaussian (sigma = 3.19, DiscreteGaussianGeneratorImpl) by inversion: cdf[j] = floor(2^64 * P(X in first j+1 values))
// over the order 0, +1, -1, +2, -2, ...; the table is built on the host in long double
struct KgGauss {
    u64 cdf[96];
    u32 len, pad;
};
__device__ __forceinline__ long long kg_gauss(const KgGauss& G, u64 r) {
    u32 j = 0;
    while (j + 1 < G.len && r >= G.cdf[j]) ++j;
    const long long mag = (long long)((j + 1) >> 1);
    return (j & 1) ? mag : -mag;   // j = 0 -> 0, 1 -> +1, 2 -> -1, ...
}

struct KgParams {
    ChaCha K;
    KgGauss G;
    u64 Q, qKS;
    u32 N, n, logN, npad, baseKS, digitsKS, ks_elem, sigma_layout;   // sigma_layout: register-resident kernels' key order
    u32 rows, nd, pad0, pad1;
};

// ---- ring secret: uniform ternary coefficients (LWEEncryptionScheme::KeyGen(N, Q), lwe-pke.cpp:48-51) ---------------------
__global__ void kg_ternary_kernel(const __grid_constant__ KgParams P, long long* __restrict__ skN) {
    const u32 i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= P.N) return;
    KgStream st(&P.K, (1ull << 60) + i);
    skN[i] = (long long)st.uniform(3) - 1;
}

// ---- switching key (lwe-pke.cpp:264-344): one warp per row (i, j, k): a <- U(Z_qKS^n), b = e + <a, s> + skN_i j baseKS^k ------
template <typename KT>
__global__ void __launch_bounds__(128) kg_ksk_kernel(const __grid_constant__ KgParams P, const long long* __restrict__ sk,
                                                     const long long* __restrict__ skN, KT* __restrict__ kA, void* __restrict__ kB) {
    const u64 row = (u64)blockIdx.x * 4 + (threadIdx.x >> 5);
    const u32 lane = threadIdx.x & 31;
    const u64 rows = (u64)P.N * P.baseKS * P.digitsKS;
    if (row >= rows) return;
    const u32 k = (u32)(row % P.digitsKS), j = (u32)((row / P.digitsKS) % P.baseKS), i = (u32)(row / ((u64)P.digitsKS * P.baseKS));
    KgStream st(&P.K, (2ull << 60) + row * 32 + lane);
    const u64 q = P.qKS;
    unsigned __int128 acc = 0;
    KT* dst = kA + row * P.npad;
    for (u32 t = lane; t < P.npad; t += 32) {
        u64 a = 0;
        if (t < P.n) {
            a = st.uniform(q);
            const long long s = sk[t];
            const u64 sm = s >= 0 ? (u64)s % q : (q - (u64)(-s) % q) % q;
            acc += (unsigned __int128)a * sm;
        }
        dst[t] = (KT)a;
    }
    u64 part = (u64)(acc % q);
#pragma unroll
    for (int o = 16; o >= 1; o >>= 1) part = (part + __shfl_xor_sync(0xffffffffu, part, o)) % q;
    if (lane == 0) {
        KgStream se(&P.K, (3ull << 60) + row);
        const long long e = kg_gauss(P.G, se.next());
        u64 pw = 1;
        for (u32 d = 0; d < k; ++d) pw = pw * P.baseKS;   // baseKS^k < qKS
        const long long sn = skN[i];
        const u64 snm = sn >= 0 ? (u64)sn : q - (u64)(-sn);
        const u64 msg = (u64)(((unsigned __int128)snm * ((u64)j * pw % q)) % q);
        const u64 em = e >= 0 ? (u64)e % q : q - (u64)(-e) % q;
        const u64 b = (part + msg + em) % q;
        if (P.ks_elem == 8) reinterpret_cast<u64*>(kB)[row] = b;
        else reinterpret_cast<u32*>(kB)[row] = (u32)b;
    }
}

// storage position of EVALUATION index p of one polynomial: sigma order for the register-resident kernels, natural otherwise
__device__ __forceinline__ u32 kg_pos(const KgParams& P, u32 p) {
    if (!P.sigma_layout) return p;
    const u32 S = P.sigma_layout == 2 ? P.N >> 1 : P.N, ps = p & (S - 1);   // 2: the two-half layout of br_quad.cuh
    const u32 E = S >> 5, l = ps / E, r = ps % E;
    return (p - ps) + (r >> 2) * 128 + 4 * l + (r & 3);
}

// one RGSW / automorphism-key row per CTA.  desc[blk] = {kind, exponent mm in [0, 2N) or automorphism index k, first output
// polynomial (in units of N), rows}: kind 0 = RGSW encryption of 0, 1 = of X^mm (negative monomials are mm >= N), 2 = automorphism
// key for X -> X^k (rows = nd, no gadget row pairs).  Row r of a block:
//   RGSW:  (a + [r even] g_r X^mm,  e + a s + [r odd] g_r X^mm),  g_r = baseG^(r/2 + 1)     rgsw-acc-cggi.cpp:76-95
//   auto:  (a,  e + a s - baseG^(r+1) s(X^k))                                               rgsw-acc-lmkcdey.cpp:204-223
// a is drawn directly in EVALUATION format (uniform there iff uniform in coefficients), e in COEFFICIENT format and transformed.
template <typename T>
__global__ void __launch_bounds__(256) kg_rgsw_kernel(const __grid_constant__ KgParams P, const __grid_constant__ BrwParams<T> W,
                                                      const uint4* __restrict__ desc, const u64* __restrict__ gpow,
                                                      const T* __restrict__ skNTT, T* __restrict__ out) {
    extern __shared__ __align__(16) unsigned char kg_smem[];
    T* e = reinterpret_cast<T*>(kg_smem);
    const u32 N = P.N, PL = N + (N >> 4);
    const u32 blk = blockIdx.x / P.rows, r = blockIdx.x % P.rows;   // P.rows = rows of the largest block kind
    const uint4 d = desc[blk];
    if (r >= d.w) return;
    const u64 rowid = (u64)d.z / 2 + r;   // unique per output row
    {
        KgStream se(&P.K, (4ull << 60) + rowid * 256 + threadIdx.x);
        for (u32 j = threadIdx.x; j < N; j += blockDim.x) {
            const long long v = kg_gauss(P.G, se.next());
            e[wphys(j)] = v >= 0 ? (T)v : (T)(P.Q - (u64)(-v));
        }
    }
    __syncthreads();
    w_ntt_fwd(e, 1, PL, P.logN, W);
    KgStream sa(&P.K, (5ull << 60) + rowid * 256 + threadIdx.x);
    const u32 M = 2 * N;
    T* o0 = out + ((size_t)d.z + 2 * (size_t)r) * N;
    T* o1 = o0 + N;
    const u64 Q = P.Q;
    for (u32 p = threadIdx.x; p < N; p += blockDim.x) {
        const u64 a = sa.uniform(Q);
        const u64 s = skNTT[p];
        u64 c0 = a, c1 = (u64)(((unsigned __int128)a * s + e[wphys(p)]) % Q);
        const u32 brp = __brev(p) >> (32 - P.logN);
        if (d.x == 1) {
            const u64 g = gpow[(r >> 1) + 1];
            const u64 mono = W.pw[(d.y * (2 * brp + 1)) & (M - 1)];
            const u64 msg = (u64)(((unsigned __int128)g * mono) % Q);
            if (r & 1) c1 = (c1 + msg) % Q; else c0 = (c0 + msg) % Q;
        } else if (d.x == 2) {
            // s(X^k) in EVALUATION format: out[brv(j)] = in[brv(((2j+1) k mod 2N) >> 1)]  (poly-impl.h:345-356)
            const u32 src = __brev((u32)((((u64)(2 * brp + 1) * d.y) % M) >> 1)) >> (32 - P.logN);
            const u64 g = gpow[r + 1];
            const u64 sub = (u64)(((unsigned __int128)g * skNTT[src]) % Q);
            c1 = (c1 + Q - sub) % Q;
        }
        o0[kg_pos(P, p)] = (T)c0;
        o1[kg_pos(P, p)] = (T)c1;
    }
}

// skN (signed) -> canonical EVALUATION-format polynomial mod Q (natural order)
template <typename T>
__global__ void __launch_bounds__(256) kg_skntt_kernel(const __grid_constant__ KgParams P, const __grid_constant__ BrwParams<T> W,
                                                       const long long* __restrict__ skN, T* __restrict__ skNTT) {
    extern __shared__ __align__(16) unsigned char kg_smem[];
    T* e = reinterpret_cast<T*>(kg_smem);
    const u32 N = P.N, PL = N + (N >> 4);
    for (u32 j = threadIdx.x; j < N; j += blockDim.x) {
        const long long v = skN[j];
        e[wphys(j)] = v >= 0 ? (T)v : (T)(P.Q - (u64)(-v));
    }
    __syncthreads();
    w_ntt_fwd(e, 1, PL, P.logN, W);
    for (u32 p = threadIdx.x; p < N; p += blockDim.x) skNTT[p] = e[wphys(p)];
}
