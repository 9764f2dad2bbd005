/* oracle/oracle.c — TEST INFRASTRUCTURE ONLY. See oracle.h for status and rules.
 *
 * Plain-C restatement of the reference's CPU algorithms for the batched-bootstrapping hot path.
 * All reference operations return canonical residues, so plain (a*b)%q arithmetic on unsigned
 * __int128 reproduces ModMulFast/ModMulFastConst results bit for bit.
 */
#include "oracle.h"
#include <math.h>
#include <omp.h>
#include <stdlib.h>
#include <string.h>
#include <time.h>

typedef uint64_t u64;
typedef int64_t i64;
typedef uint32_t u32;
typedef unsigned __int128 u128;

/* operands are canonical residues: below 2^32 the product fits in 64 bits (same value, no 128-bit division) */
static inline u64 mulmod(u64 a, u64 b, u64 q) { return (q >> 32) == 0 ? a * b % q : (u64)((u128)a * b % q); }
static inline u64 addmod(u64 a, u64 b, u64 q) { u64 s = a + b; return s >= q ? s - q : s; }
static inline u64 submod(u64 a, u64 b, u64 q) { return a >= b ? a - b : a + q - b; }
static u64 powmod(u64 a, u64 e, u64 q) {
    u64 r = 1 % q;
    a %= q;
    while (e) {
        if (e & 1) r = mulmod(r, a, q);
        a = mulmod(a, a, q);
        e >>= 1;
    }
    return r;
}
static u64 invmod(u64 a, u64 q) { return powmod(a, q - 2, q); } /* q prime */

static u32 bitrev(u32 x, u32 bits) {
    u32 r = 0;
    for (u32 i = 0; i < bits; ++i) r |= ((x >> i) & 1u) << (bits - 1 - i);
    return r;
}
static u32 ilog2(u64 x) { u32 r = 0; while (x >>= 1) ++r; return r; }

/* deterministic Miller-Rabin for 64-bit (reference uses a randomized MillerRabinPrimalityTest; on primes
 * and composites both agree with overwhelming probability, and the results are pinned against it). */
static int is_prime(u64 n) {
    if (n < 2) return 0;
    static const u64 small[] = {2, 3, 5, 7, 11, 13, 17, 19, 23, 29, 31, 37};
    for (int i = 0; i < 12; ++i) {
        if (n == small[i]) return 1;
        if (n % small[i] == 0) return 0;
    }
    u64 d = n - 1; int s = 0;
    while (!(d & 1)) { d >>= 1; ++s; }
    for (int i = 0; i < 12; ++i) {
        u64 x = powmod(small[i], d, n);
        if (x == 1 || x == n - 1) continue;
        int comp = 1;
        for (int r = 1; r < s; ++r) {
            x = mulmod(x, x, n);
            if (x == n - 1) { comp = 0; break; }
        }
        if (comp) return 0;
    }
    return 1;
}

/* nbtheory-impl.h:350-373 */
u64 orc_last_prime(u32 bits, u64 m) {
    u64 q = (u64)1 << bits;
    u64 r = q % m;
    u64 qn = q + 1 - r;
    if (r < 2) qn -= m;
    while (!is_prime(qn)) qn -= m;
    return qn;
}
/* nbtheory-impl.h:386-393 */
u64 orc_previous_prime(u64 q, u64 m) {
    u64 qn = q - m;
    while (!is_prime(qn)) qn -= m;
    return qn;
}
/* nbtheory-impl.h:183-231: the *minimal* primitive m-th root of unity (m a power of two) */
u64 orc_root_of_unity(u64 m, u64 q) {
    if ((q - 1) % m) return 0;
    u64 r = 0;
    for (u64 g = 2; g < q; ++g) {
        r = powmod(g, (q - 1) / m, q);
        if (powmod(r, m / 2, q) == q - 1) break; /* order exactly m */
    }
    u64 r2 = mulmod(r, r, q), x = r, mn = r;
    for (u64 k = 1; k < m; k += 2) { /* all odd powers = all primitive roots */
        if (x < mn && x != 1) mn = x;
        x = mulmod(x, r2, q);
    }
    return mn;
}
/* ildcrtparams.h:100-118 */
int orc_dcrt_chain(u32 N, u32 L, u32 bits, u64* q, u64* psi) {
    u64 m = 2ull * N;
    u64 p = orc_last_prime(bits, m);
    for (u32 l = 0; l < L; ++l) {
        if (l) p = orc_previous_prime(p, m);
        q[l] = p;
        psi[l] = orc_root_of_unity(m, p);
    }
    return 0;
}

/* ---- NTT ---------------------------------------------------------------------------------------- */
/* transformnat-impl.h:777-831: Table[brv(i)] = psi^i, TableI[brv(i)] = psi^-i */
static void ntt_tables(u32 N, u64 Q, u64 psi, u64* tw, u64* twi) {
    u32 lg = ilog2(N);
    u64 x = 1, xi = 1, psii = invmod(psi, Q);
    for (u32 i = 0; i < N; ++i) {
        u32 r = bitrev(i, lg);
        tw[r] = x; twi[r] = xi;
        x = mulmod(x, psi, Q); xi = mulmod(xi, psii, Q);
    }
}
/* transformnat-impl.h:302-373 (Cooley-Tukey, natural in, bit-reversed out) */
static void ntt_fwd(u32 N, u64 Q, const u64* tw, u64* x) {
    for (u32 m = 1, t = N >> 1; m < N; m <<= 1, t >>= 1)
        for (u32 i = 0; i < m; ++i) {
            u64 w = tw[m + i];
            u32 j1 = 2 * i * t;
            for (u32 j = j1; j < j1 + t; ++j) {
                u64 u = mulmod(x[j + t], w, Q), lo = x[j];
                x[j] = addmod(lo, u, Q);
                x[j + t] = submod(lo, u, Q);
            }
        }
}
/* transformnat-impl.h:511-624 (Gentleman-Sande, bit-reversed in, natural out, times N^-1) */
static void ntt_inv(u32 N, u64 Q, const u64* twi, u64 ninv, u64* x) {
    for (u32 m = N >> 1, t = 1; m >= 1; m >>= 1, t <<= 1)
        for (u32 i = 0; i < m; ++i) {
            u64 w = twi[m + i];
            u32 j1 = 2 * i * t;
            for (u32 j = j1; j < j1 + t; ++j) {
                u64 lo = x[j], hi = x[j + t];
                x[j] = addmod(lo, hi, Q);
                x[j + t] = mulmod(submod(lo, hi, Q), w, Q);
            }
        }
    for (u32 i = 0; i < N; ++i) x[i] = mulmod(x[i], ninv, Q);
}
int orc_ntt(u64 Q, u64 psi, u32 N, u64* data, u64 count, int inverse, int threads) {
    u64* tw = malloc(sizeof(u64) * N * 2), *twi = tw + N;
    ntt_tables(N, Q, psi, tw, twi);
    u64 ninv = invmod(N % Q, Q);
#pragma omp parallel for num_threads(threads > 0 ? threads : 1)
    for (u64 i = 0; i < count; ++i) {
        if (inverse) ntt_inv(N, Q, twi, ninv, data + i * N);
        else ntt_fwd(N, Q, tw, data + i * N);
    }
    free(tw);
    return 0;
}

/* ---- context ------------------------------------------------------------------------------------ */
struct OrcCtx {
    u32 n, N, baseKS, digitsKS, baseG, digitsG, baseR, digitsR, numAutoKeys;
    int method;
    u64 q, Q, qKS, psi, ninv;
    u64 *tw, *twi;
    u64* mono;   /* GINX: [2N][N] EVAL, rgsw-cryptoparameters.cpp:96-113 */
    i64* logGen; /* LMKCDEY: [2N], rgsw-cryptoparameters.cpp:115-127 */
    u64* bsk;
    u64 *kskA, *kskB;
    u32 rows; /* digitsG2 */
};

OrcCtx* orc_ctx_create(u32 n, u32 N, u64 q, u64 Q, u64 qKS, u32 baseKS, u32 baseG, u32 baseR, u32 numAutoKeys,
                       int method) {
    OrcCtx* c = calloc(1, sizeof(OrcCtx));
    c->n = n; c->N = N; c->q = q; c->Q = Q; c->qKS = qKS; c->baseKS = baseKS; c->baseG = baseG; c->baseR = baseR;
    c->numAutoKeys = numAutoKeys; c->method = method;
    /* lwe-pke.cpp:355, rgsw-cryptoparameters.h:93-94, rgsw-cryptoparameters.cpp:39-40: same double expressions */
    c->digitsKS = (u32)ceil(log((double)qKS) / log((double)baseKS));
    c->digitsG = (u32)ceil(log((double)Q) / log((double)baseG));
    c->digitsR = (u32)ceil(log((double)q) / log((double)baseR));
    c->rows = (c->digitsG - 1) * 2;
    c->psi = orc_root_of_unity(2ull * N, Q);
    c->tw = malloc(sizeof(u64) * N); c->twi = malloc(sizeof(u64) * N);
    ntt_tables(N, Q, c->psi, c->tw, c->twi);
    c->ninv = invmod(N % Q, Q);
    if (method == ORC_GINX) {
        c->mono = calloc((size_t)2 * N * N, sizeof(u64));
        for (u32 i = 0; i < 2 * N; ++i) {
            u64* p = c->mono + (size_t)i * N;
            p[0] = Q - 1;
            if (i < N) p[i] = addmod(p[i], 1, Q);
            else p[i - N] = submod(p[i - N], 1, Q);
            ntt_fwd(N, Q, c->tw, p);
        }
    }
    if (method == ORC_LMKCDEY) {
        u32 M = 2 * N;
        c->logGen = calloc(M, sizeof(i64));
        u32 g = 1;
        c->logGen[M - g] = M;
        for (u32 i = 1; i < N / 2; ++i) {
            g = (g * 5) % M;
            c->logGen[g] = i;
            c->logGen[M - g] = -(i64)i;
        }
    }
    return c;
}
void orc_ctx_destroy(OrcCtx* c) {
    if (!c) return;
    free(c->tw); free(c->twi); free(c->mono); free(c->logGen); free(c->bsk); free(c->kskA); free(c->kskB);
    free(c);
}
int orc_ctx_params(OrcCtx* c, u64* o) {
    o[0] = c->n; o[1] = c->N; o[2] = c->q; o[3] = c->Q; o[4] = c->qKS; o[5] = c->baseKS; o[6] = c->baseG;
    o[7] = c->baseR; o[8] = c->digitsG; o[9] = c->numAutoKeys; o[10] = 0; o[11] = c->psi; o[12] = c->digitsKS;
    o[13] = c->method; o[14] = c->digitsR;
    return 0;
}
/* canonical flat BSK (EVALUATION format, reference index order within each polynomial):
 *  GINX    : [n][2 (ek[0][0][i], ek[0][1][i])][rows][2][N]
 *  AP      : [n][baseR][digitsR][rows][2][N]      (j = 0 slots unused, zero)
 *  LMKCDEY : [n][rows][2][N]  then  [numAutoKeys+1][digitsG-1][2][N]   (ek[0][0][i]; ek[0][1][0..w]) */
u64 orc_bsk_len(OrcCtx* c) {
    u64 rg = (u64)c->rows * 2 * c->N;
    if (c->method == ORC_GINX) return (u64)c->n * 2 * rg;
    if (c->method == ORC_AP) return (u64)c->n * c->baseR * c->digitsR * rg;
    return (u64)c->n * rg + (u64)(c->numAutoKeys + 1) * (c->digitsG - 1) * 2 * c->N;
}
u64 orc_ksk_len(OrcCtx* c) { return (u64)c->N * c->baseKS * c->digitsKS * c->n; }
int orc_load_bsk(OrcCtx* c, const u64* b) {
    free(c->bsk);
    u64 len = orc_bsk_len(c);
    c->bsk = malloc(len * sizeof(u64));
    memcpy(c->bsk, b, len * sizeof(u64));
    return 0;
}
int orc_load_ksk(OrcCtx* c, const u64* A, const u64* B) {
    free(c->kskA); free(c->kskB);
    u64 len = orc_ksk_len(c);
    c->kskA = malloc(len * sizeof(u64)); c->kskB = malloc(len / c->n * sizeof(u64));
    memcpy(c->kskA, A, len * sizeof(u64)); memcpy(c->kskB, B, len / c->n * sizeof(u64));
    return 0;
}

/* ---- own key generation (test infrastructure; not the reference PRNG) ------------------------------ */
typedef struct { u64 s; } Rng;
static u64 rng_next(Rng* r) {
    u64 z = (r->s += 0x9E3779B97F4A7C15ull);
    z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ull;
    z = (z ^ (z >> 27)) * 0x94D049BB133111EBull;
    return z ^ (z >> 31);
}
static u64 rng_uniform(Rng* r, u64 mod) { return (u64)(((u128)rng_next(r) * mod) >> 64); }
static i64 rng_gauss(Rng* r, double sigma) {
    double u1 = ((rng_next(r) >> 11) + 1.0) / 9007199254740993.0, u2 = (rng_next(r) >> 11) / 9007199254740992.0;
    return (i64)llround(sigma * sqrt(-2.0 * log(u1)) * cos(6.283185307179586 * u2));
}
static u64 to_mod(i64 v, u64 q) { return v >= 0 ? (u64)v % q : q - ((u64)(-v) % q == 0 ? q : (u64)(-v) % q); }

/* one RGSW row set "rows x 2 x N" encrypting X^mm * G (sign flips if mm >= N): KeyGenDM / KeyGenLMKCDEY /
 * KeyGenCGGI shape (rgsw-acc-dm.cpp:80-114, rgsw-acc-lmkcdey.cpp:163-200, rgsw-acc-cggi.cpp:71-96) */
static void gen_rgsw(OrcCtx* c, Rng* r, const u64* skNTT, int has_msg, u32 mm, int neg, u64* out) {
    u32 N = c->N; u64 Q = c->Q;
    u64* a = malloc(sizeof(u64) * N);
    for (u32 i = 0; i < c->rows; ++i) {
        u64* r0 = out + ((size_t)i * 2 + 0) * N, *r1 = out + ((size_t)i * 2 + 1) * N;
        for (u32 k = 0; k < N; ++k) { r0[k] = rng_uniform(r, Q); a[k] = r0[k]; }
        ntt_fwd(N, Q, c->tw, a);
        for (u32 k = 0; k < N; ++k) r1[k] = to_mod(rng_gauss(r, 3.19), Q);
        if (has_msg) {
            u64 gp = powmod(c->baseG, (i >> 1) + 1, Q);
            u64* tgt = (i & 1) ? r1 : r0;
            tgt[mm] = neg ? submod(tgt[mm], gp, Q) : addmod(tgt[mm], gp, Q);
        }
        ntt_fwd(N, Q, c->tw, r0);
        ntt_fwd(N, Q, c->tw, r1);
        for (u32 k = 0; k < N; ++k) r1[k] = addmod(r1[k], mulmod(a[k], skNTT[k], Q), Q);
    }
    free(a);
}
static void rgsw_of_exponent(OrcCtx* c, Rng* r, const u64* skNTT, i64 m, u64* out) {
    i64 q = (i64)c->q, N = c->N;
    i64 mm = (((m % q) + q) % q) * (2 * N / q);
    int neg = 0;
    if (mm >= N) { mm -= N; neg = 1; }
    gen_rgsw(c, r, skNTT, 1, (u32)mm, neg, out);
}
static void automorph_eval(u32 N, u32 k, const u64* in, u64* out);

int orc_keygen(OrcCtx* c, u64 seed, int key_dist, i64* sk, u64* bsk, u64* kskA, u64* kskB) {
    Rng r = {seed};
    u32 n = c->n, N = c->N; u64 Q = c->Q, qKS = c->qKS;
    for (u32 i = 0; i < n; ++i) sk[i] = key_dist ? rng_gauss(&r, 3.19) : (i64)rng_uniform(&r, 3) - 1;
    i64* skN = malloc(sizeof(i64) * N);
    for (u32 i = 0; i < N; ++i) skN[i] = (i64)rng_uniform(&r, 3) - 1;
    /* KSK: lwe-pke.cpp:264-344 */
    for (u32 i = 0; i < N; ++i)
        for (u32 j = 0; j < c->baseKS; ++j) {
            u64 dk = 1;
            for (u32 k = 0; k < c->digitsKS; ++k, dk *= c->baseKS) {
                size_t row = ((size_t)i * c->baseKS + j) * c->digitsKS + k;
                u64* a = kskA + row * n;
                u64 b = addmod(to_mod(rng_gauss(&r, 3.19), qKS), mulmod(to_mod(skN[i], qKS), (j * dk) % qKS, qKS), qKS);
                for (u32 t = 0; t < n; ++t) {
                    a[t] = rng_uniform(&r, qKS);
                    b = addmod(b, mulmod(a[t], to_mod(sk[t], qKS), qKS), qKS);
                }
                kskB[row] = b;
            }
        }
    u64* skNTT = malloc(sizeof(u64) * N);
    for (u32 i = 0; i < N; ++i) skNTT[i] = to_mod(skN[i], Q);
    ntt_fwd(N, Q, c->tw, skNTT);
    size_t rg = (size_t)c->rows * 2 * N;
    if (c->method == ORC_GINX) { /* rgsw-acc-cggi.cpp:39-57 */
        for (u32 i = 0; i < n; ++i) {
            gen_rgsw(c, &r, skNTT, sk[i] == 1, 0, 0, bsk + ((size_t)i * 2 + 0) * rg);
            gen_rgsw(c, &r, skNTT, sk[i] == -1, 0, 0, bsk + ((size_t)i * 2 + 1) * rg);
        }
    } else if (c->method == ORC_AP) { /* rgsw-acc-dm.cpp:39-60 */
        memset(bsk, 0, orc_bsk_len(c) * sizeof(u64));
        for (u32 i = 0; i < n; ++i)
            for (u32 j = 1; j < c->baseR; ++j) {
                i64 dr = 1;
                for (u32 k = 0; k < c->digitsR; ++k, dr *= c->baseR)
                    rgsw_of_exponent(c, &r, skNTT, sk[i] * (i64)j * dr,
                                     bsk + (((size_t)i * c->baseR + j) * c->digitsR + k) * rg);
            }
    } else { /* rgsw-acc-lmkcdey.cpp:39-68, :203-223 */
        for (u32 i = 0; i < n; ++i) rgsw_of_exponent(c, &r, skNTT, sk[i], bsk + (size_t)i * rg);
        u64* ak = bsk + (size_t)n * rg;
        u32 dg = c->digitsG - 1, M = 2 * N;
        u64* skA = malloc(sizeof(u64) * N);
        for (u32 t = 0; t <= c->numAutoKeys; ++t) {
            u32 k = t == 0 ? M - 5 : (u32)powmod(5, t, M);
            automorph_eval(N, k, skNTT, skA);
            for (u32 d = 0; d < dg; ++d) {
                u64* r0 = ak + (((size_t)t * dg + d) * 2 + 0) * N, *r1 = r0 + N;
                u64 gp = powmod(c->baseG, d + 1, Q);
                u64* e = malloc(sizeof(u64) * N);
                for (u32 x = 0; x < N; ++x) { r0[x] = rng_uniform(&r, Q); e[x] = to_mod(rng_gauss(&r, 3.19), Q); }
                ntt_fwd(N, Q, c->tw, e); /* reference samples e directly in EVAL; distribution detail only */
                for (u32 x = 0; x < N; ++x)
                    r1[x] = addmod(submod(e[x], mulmod(skA[x], gp, Q), Q), mulmod(r0[x], skNTT[x], Q), Q);
                free(e);
            }
        }
        free(skA);
    }
    free(skNTT); free(skN);
    return 0;
}
/* lwe-pke.cpp:104-130: b = <a,s> + e + m*(mod/p) */
int orc_encrypt(OrcCtx* c, const i64* sk, i64 m, u64 p, u64 mod, u64 seed, u64* a, u64* b) {
    Rng r = {seed};
    u64 bb = addmod(((u64)(((m % (i64)p) + (i64)p) % (i64)p)) * (mod / p) % mod, to_mod(rng_gauss(&r, 3.19), mod), mod);
    for (u32 i = 0; i < c->n; ++i) {
        a[i] = rng_uniform(&r, mod);
        bb = addmod(bb, mulmod(a[i], to_mod(sk[i], mod), mod), mod);
    }
    *b = bb;
    return 0;
}
/* lwe-pke.cpp:183-223 */
i64 orc_decrypt(OrcCtx* c, const i64* sk, const u64* a, u64 b, u32 len, u64 mod, u64 p) {
    (void)c;
    u64 inner = 0;
    for (u32 i = 0; i < len; ++i) inner = addmod(inner, mulmod(a[i] % mod, to_mod(sk[i], mod), mod), mod);
    u64 r = submod(b % mod, inner, mod);
    r = addmod(r, mod / (p * 2), mod);
    return (i64)(((u128)p * r) / mod);
}

/* ---- hot path ----------------------------------------------------------------------------------- */
/* rgsw-acc.cpp:54-91 (two-component) — out [rows][N], out[2j+c] from component c */
static void decompose2(OrcCtx* c, const u64* in0, const u64* in1, u64* out) {
    u32 N = c->N, rows = c->rows;
    u64 QHalf = c->Q >> 1; i64 Q = (i64)c->Q;
    i64 gBits = __builtin_ctz(c->baseG), sh = 64 - gBits;
    memset(out, 0, sizeof(u64) * rows * N);
    const u64* in[2] = {in0, in1};
    for (int comp = 0; comp < 2; ++comp)
        for (u32 k = 0; k < N; ++k) {
            u64 t = in[comp][k];
            i64 d = (i64)(t < QHalf ? t : t - (u64)Q);
            i64 r = (i64)((u64)d << sh) >> sh;
            d = (d - r) >> gBits;
            for (u32 dd = 0; dd < rows; dd += 2) {
                r = (i64)((u64)d << sh) >> sh;
                d = (d - r) >> gBits;
                if (r < 0) r += Q;
                out[(size_t)(dd + comp) * N + k] += (u64)r;
            }
        }
}
/* rgsw-acc.cpp:94-119 (single polynomial) — out [digitsG-1][N] */
static void decompose1(OrcCtx* c, const u64* in, u64* out) {
    u32 N = c->N, dg = c->digitsG - 1;
    u64 QHalf = c->Q >> 1; i64 Q = (i64)c->Q;
    i64 gBits = __builtin_ctz(c->baseG), sh = 64 - gBits;
    for (u32 k = 0; k < N; ++k) {
        u64 t = in[k];
        i64 d = (i64)(t < QHalf ? t : t - (u64)Q);
        i64 r = (i64)((u64)d << sh) >> sh;
        d = (d - r) >> gBits;
        for (u32 dd = 0; dd < dg; ++dd) {
            r = (i64)((u64)d << sh) >> sh;
            d = (d - r) >> gBits;
            if (r < 0) r += Q;
            out[(size_t)dd * N + k] = (u64)r;
        }
    }
}
int orc_decompose(OrcCtx* c, const u64* in, u64* out) { decompose2(c, in, in + c->N, out); return 0; }

/* INTT both components, decompose, NTT the digits: the common front half of AddToAcc* */
static void acc_to_dct(OrcCtx* c, const u64* acc, u64* dct, u64* tmp) {
    u32 N = c->N;
    memcpy(tmp, acc, sizeof(u64) * 2 * N);
    ntt_inv(N, c->Q, c->twi, c->ninv, tmp);
    ntt_inv(N, c->Q, c->twi, c->ninv, tmp + N);
    decompose2(c, tmp, tmp + N, dct);
    for (u32 d = 0; d < c->rows; ++d) ntt_fwd(N, c->Q, c->tw, dct + (size_t)d * N);
}
/* sum_d dct[d] * ek[d][col]  (ek: [rows][2][N]) */
static void extprod_col(OrcCtx* c, const u64* dct, const u64* ek, int col, u64* out) {
    u32 N = c->N; u64 Q = c->Q;
    for (u32 k = 0; k < N; ++k) {
        u64 s = 0;
        for (u32 d = 0; d < c->rows; ++d)
            s = addmod(s, mulmod(dct[(size_t)d * N + k], ek[((size_t)d * 2 + col) * N + k], Q), Q);
        out[k] = s;
    }
}
/* rgsw-acc-cggi.cpp:102-151 */
static void add_to_acc_cggi(OrcCtx* c, const u64* ek1, const u64* ek2, u32 a, u64* acc, u64* ws) {
    u32 N = c->N, M = 2 * N; u64 Q = c->Q;
    u64 *dct = ws, *tmp = ws + (size_t)c->rows * N, *t2 = tmp + 2 * N;
    acc_to_dct(c, acc, dct, tmp);
    u32 ip = a, in = (a == 0) ? 0 : M - a;
    const u64* mp = c->mono + (size_t)(ip == M ? 0 : ip) * N;
    const u64* mn = c->mono + (size_t)(in == M ? 0 : in) * N;
    for (int col = 0; col < 2; ++col) {
        extprod_col(c, dct, ek1, col, t2);
        for (u32 k = 0; k < N; ++k) acc[col * N + k] = addmod(acc[col * N + k], mulmod(t2[k], mp[k], Q), Q);
        extprod_col(c, dct, ek2, col, t2);
        for (u32 k = 0; k < N; ++k) acc[col * N + k] = addmod(acc[col * N + k], mulmod(t2[k], mn[k], Q), Q);
    }
}
/* rgsw-acc-dm.cpp:118-143 and rgsw-acc-lmkcdey.cpp:228-254: acc = dct x ek (replace) */
static void add_to_acc_replace(OrcCtx* c, const u64* ek, u64* acc, u64* ws) {
    u32 N = c->N;
    u64 *dct = ws, *tmp = ws + (size_t)c->rows * N;
    acc_to_dct(c, acc, dct, tmp);
    extprod_col(c, dct, ek, 0, acc);
    extprod_col(c, dct, ek, 1, acc + N);
}
/* poly-impl.h:345-356 (EVALUATION-format automorphism) */
static void automorph_eval(u32 N, u32 k, const u64* in, u64* out) {
    u32 lg = ilog2(N), mask = N - 1;
    for (u32 j = 0, jk = k; j < N; ++j, jk += 2 * k) out[bitrev(j, lg)] = in[bitrev((jk >> 1) & mask, lg)];
}
/* rgsw-acc-lmkcdey.cpp:257-287 */
static void automorphism_step(OrcCtx* c, u32 k, const u64* ak, u64* acc, u64* ws) {
    u32 N = c->N, dg = c->digitsG - 1; u64 Q = c->Q;
    u64 *cta = ws, *dcta = ws + N, *t = dcta + (size_t)dg * N;
    automorph_eval(N, k, acc + N, t);
    memcpy(acc + N, t, sizeof(u64) * N);
    automorph_eval(N, k, acc, cta);
    memset(acc, 0, sizeof(u64) * N);
    ntt_inv(N, Q, c->twi, c->ninv, cta);
    decompose1(c, cta, dcta);
    for (u32 d = 0; d < dg; ++d) ntt_fwd(N, Q, c->tw, dcta + (size_t)d * N);
    for (int col = 0; col < 2; ++col)
        for (u32 d = 0; d < dg; ++d)
            for (u32 x = 0; x < N; ++x)
                acc[col * N + x] =
                    addmod(acc[col * N + x], mulmod(dcta[(size_t)d * N + x], ak[((size_t)d * 2 + col) * N + x], Q), Q);
}

int orc_evalacc(OrcCtx* c, const u64* a, u64 amod, u64* acc) {
    u32 n = c->n, N = c->N, M = 2 * N;
    size_t rg = (size_t)c->rows * 2 * N;
    u64* ws = malloc(sizeof(u64) * ((size_t)c->rows * N + 8 * N));
    if (c->method == ORC_GINX) { /* rgsw-acc-cggi.cpp:59-68 */
        u64 MbyMod = M / amod;
        for (u32 i = 0; i < n; ++i) {
            u64 ai = (a[i] == 0 ? 0 : amod - a[i]) * MbyMod;
            add_to_acc_cggi(c, c->bsk + ((size_t)i * 2) * rg, c->bsk + ((size_t)i * 2 + 1) * rg, (u32)ai, acc, ws);
        }
    } else if (c->method == ORC_AP) { /* rgsw-acc-dm.cpp:62-77 (uses params q, u64 wrap as the reference) */
        for (u32 i = 0; i < n; ++i) {
            u64 aI = (0 < a[i]) ? (u64)0 + c->q - a[i] : 0;
            for (u32 k = 0; k < c->digitsR; ++k, aI /= c->baseR) {
                u32 a0 = (u32)(aI % c->baseR);
                if (a0) add_to_acc_replace(c, c->bsk + (((size_t)i * c->baseR + a0) * c->digitsR + k) * rg, acc, ws);
            }
        }
    } else { /* rgsw-acc-lmkcdey.cpp:70-158 */
        u32 Nh = N / 2, w = c->numAutoKeys, dg = c->digitsG - 1;
        const u64* ak = c->bsk + (size_t)n * rg;
        size_t akst = (size_t)dg * 2 * N;
        /* buckets keyed by logGen value in [-(Nh-1), Nh-1] or M; members in ascending i */
        i64* idx = malloc(sizeof(i64) * n);
        for (u32 i = 0; i < n; ++i) {
            u32 aodd = (u32)((a[i] == 0 ? 0 : M - a[i]) | 1u);
            idx[i] = c->logGen[aodd];
        }
        u64* t = malloc(sizeof(u64) * N);
        automorph_eval(N, M - 5, acc + N, t);
        memcpy(acc + N, t, sizeof(u64) * N);
        free(t);
        u32 nSkips = 0;
#define BUCKET(key)                                                                   \
    for (u32 j_ = 0; j_ < n; ++j_)                                                    \
        if (idx[j_] == (key)) add_to_acc_replace(c, c->bsk + (size_t)j_ * rg, acc, ws)
#define HAS(key, flag)                                                                \
    do { flag = 0; for (u32 j_ = 0; j_ < n; ++j_) if (idx[j_] == (key)) { flag = 1; break; } } while (0)
        for (int pass = 0; pass < 2; ++pass) {
            for (u32 i = Nh - 1; i > 0; --i) {
                i64 key = pass == 0 ? -(i64)i : (i64)i;
                int has; HAS(key, has);
                if (has) {
                    if (nSkips != 0) {
                        automorphism_step(c, (u32)powmod(5, nSkips, M), ak + nSkips * akst, acc, ws);
                        nSkips = 0;
                    }
                    BUCKET(key);
                }
                nSkips++;
                if (nSkips == w || i == 1) {
                    automorphism_step(c, (u32)powmod(5, nSkips, M), ak + nSkips * akst, acc, ws);
                    nSkips = 0;
                }
            }
            if (pass == 0) {
                BUCKET((i64)M);
                automorphism_step(c, M - 5, ak, acc, ws);
            } else {
                BUCKET(0);
            }
        }
#undef BUCKET
#undef HAS
        free(idx);
    }
    free(ws);
    return 0;
}

/* lwe-pke.cpp:41-46: floor(0.5 + v*q/Q) in double, then mod q */
static u64 round_qQ(u64 v, u64 q, u64 Q) {
    volatile double prod = (double)v * (double)q; /* keep the reference's evaluation order, no FMA contraction */
    volatile double quo = prod / (double)Q;
    return (u64)floor(0.5 + quo) % q;
}
int orc_modswitch(u64 qto, u64 qfrom, const u64* in, u64* out, u64 len) {
    for (u64 i = 0; i < len; ++i) out[i] = round_qQ(in[i], qto, qfrom);
    return 0;
}
/* lwe-pke.cpp:348-372 */
int orc_keyswitch(OrcCtx* c, const u64* a, u64 b, u64* oa, u64* ob) {
    u32 n = c->n, N = c->N; u64 q = c->qKS;
    memset(oa, 0, sizeof(u64) * n);
    for (u32 i = 0; i < N; ++i) {
        u64 at = a[i];
        for (u32 j = 0; j < c->digitsKS; ++j) {
            u64 a0 = at % c->baseKS;
            at /= c->baseKS;
            size_t row = ((size_t)i * c->baseKS + a0) * c->digitsKS + j;
            b = submod(b, c->kskB[row], q);
            const u64* r = c->kskA + row * n;
            for (u32 k = 0; k < n; ++k) oa[k] = submod(oa[k], r[k], q);
        }
    }
    *ob = b;
    return 0;
}

/* BootstrapGateCore / BootstrapFuncCore (binfhe-base-scheme.cpp:525-583, :589-626) with the test vector given
 * as a table over x = (b - j) mod ctmod, then Transpose + 2 INTT + extraction (:111-121, :634-639).
 * ext: N coefficients of a then b = acc1[0] (caller adds the gate offset). */
int orc_bootstrap_core(OrcCtx* c, const u64* a, u64 b, u64 ctmod, const u64* lut, u64* ext) {
    u32 N = c->N; u64 Q = c->Q;
    u64* acc = calloc(2 * (size_t)N, sizeof(u64));
    u32 factor = (u32)(2ull * N / ctmod);
    for (u64 j = 0; j < ctmod / 2; ++j) acc[N + j * factor] = lut[submod(b % ctmod, j % ctmod, ctmod)];
    ntt_fwd(N, Q, c->tw, acc + N);
    orc_evalacc(c, a, ctmod, acc);
    /* Transpose = automorphism by 2N-1 in EVALUATION format (poly-interface.h:440-447) */
    u64* t = malloc(sizeof(u64) * N);
    automorph_eval(N, 2 * N - 1, acc, t);
    ntt_inv(N, Q, c->twi, c->ninv, t);
    ntt_inv(N, Q, c->twi, c->ninv, acc + N);
    memcpy(ext, t, sizeof(u64) * N);
    ext[N] = acc[N];
    free(t); free(acc);
    return 0;
}
/* SwitchCTtoqn (lwe-pke.cpp:170-178) with final modulus fmod */
static void switch_to_qn(OrcCtx* c, const u64* ext, u64 fmod, u64* oa, u64* ob) {
    u32 N = c->N, n = c->n;
    u64* ms = malloc(sizeof(u64) * (N + 1 + n + 1));
    orc_modswitch(c->qKS, c->Q, ext, ms, N + 1);
    u64* ks = ms + N + 1;
    orc_keyswitch(c, ms, ms[N], ks, ks + n);
    orc_modswitch(fmod, c->qKS, ks, oa, n);
    orc_modswitch(fmod, c->qKS, ks + n, ob, 1);
    free(ms);
}
/* gate test-vector table: binfhe-base-scheme.cpp:535-567, gate constants rgsw-cryptoparameters.cpp:78-92 */
static void gate_lut(OrcCtx* c, int gate, u64 q, u64 p, u64* lut) {
    u64 gc;
    switch (gate) {
        case ORC_OR: gc = 5 * (q >> 3); break;
        case ORC_AND: gc = 7 * (q >> 3); break;
        case ORC_NOR: gc = 1 * (q >> 3); break;
        case ORC_NAND: gc = 3 * (q >> 3); break;
        case ORC_XOR: case ORC_XOR_FAST: gc = 6 * (q >> 3); break;
        case ORC_XNOR: case ORC_XNOR_FAST: gc = 2 * (q >> 3); break;
        case ORC_MAJORITY: gc = 7 * (q >> 3); break;
        case ORC_AND3: gc = 11 * (q / 12); break;
        case ORC_OR3: gc = 7 * (q / 12); break;
        case ORC_AND4: gc = 15 * (q >> 4); break;
        default: gc = 9 * (q >> 4); break; /* OR4 */
    }
    u64 q1 = gc, q2 = addmod(q1, q >> 1, q);
    int swap = q1 >= q2;
    u64 lb = swap ? q2 : q1, ub = swap ? q1 : q2;
    u64 Q2p = c->Q / (2 * p) + 1, Q2pNeg = c->Q - Q2p;
    u64 lv = swap ? Q2p : Q2pNeg, uv = swap ? Q2pNeg : Q2p;
    for (u64 x = 0; x < q; ++x) lut[x] = (x >= lb && x < ub) ? lv : uv;
}
static void gate_finish(OrcCtx* c, u64* ext, u64 p, int extended, u64* oa, u64* ob) {
    u32 N = c->N;
    ext[N] = addmod(ext[N], c->Q / (2 * p) + 1, c->Q);
    if (extended) { memcpy(oa, ext, sizeof(u64) * N); *ob = ext[N]; }
    else switch_to_qn(c, ext, c->q, oa, ob);
}
/* binfhe-base-scheme.cpp:76-126 (inputs already at (n, q)) */
static void bingate_one(OrcCtx* c, int gate, const u64* a1, u64 b1, const u64* a2, u64 b2, int extended, u64* oa,
                        u64* ob) {
    u32 n = c->n, N = c->N; u64 q = c->q;
    u64* a = malloc(sizeof(u64) * (n + q + N + 1));
    u64 *lut = a + n, *ext = lut + q;
    u64 b = addmod(b1, b2, q);
    for (u32 i = 0; i < n; ++i) a[i] = addmod(a1[i], a2[i], q);
    if (gate == ORC_XOR || gate == ORC_XNOR || gate == ORC_XOR_FAST || gate == ORC_XNOR_FAST) {
        for (u32 i = 0; i < n; ++i) a[i] = addmod(a[i], a[i], q);
        b = addmod(b, b, q);
    }
    gate_lut(c, gate, q, 4, lut);
    orc_bootstrap_core(c, a, b, q, lut, ext);
    gate_finish(c, ext, 4, extended, oa, ob);
    free(a);
}
int orc_evalbingate(OrcCtx* c, int gate, const u64* a1, const u64* b1, const u64* a2, const u64* b2, u64* oa, u64* ob,
                    u64 count, int extended, int threads) {
    u32 n = c->n, on = extended ? c->N : n;
#pragma omp parallel for num_threads(threads > 0 ? threads : 1) schedule(dynamic)
    for (u64 i = 0; i < count; ++i)
        bingate_one(c, gate, a1 + i * n, b1[i], a2 + i * n, b2[i], extended, oa + i * on, ob + i);
    return 0;
}
/* binfhe-base-scheme.cpp:223-237 */
static void eval_not(OrcCtx* c, const u64* a, u64 b, u64* oa, u64* ob) {
    u64 q = c->q;
    for (u32 i = 0; i < c->n; ++i) oa[i] = a[i] == 0 ? 0 : q - a[i];
    *ob = submod(q >> 2, b, q);
}
/* binfhe-base-scheme.cpp:129-187 */
int orc_evalbingate_multi(OrcCtx* c, int gate, u32 k, const u64* a, const u64* b, u64* oa, u64* ob, u64 count, u64 p,
                          int threads) {
    u32 n = c->n, N = c->N; u64 q = c->q;
    int rc = 0;
#pragma omp parallel for num_threads(threads > 0 ? threads : 1) schedule(dynamic)
    for (u64 i = 0; i < count; ++i) {
        const u64* ai = a + i * k * n; const u64* bi = b + i * k;
        if (gate == ORC_CMUX) {
            if (k != 3) { rc = -1; continue; }
            u64* t = malloc(sizeof(u64) * 3 * (n + 1));
            u64 *na = t, *x1 = t + n + 1, *x2 = x1 + n + 1;
            eval_not(c, ai + 2 * n, bi[2], na, na + n);
            bingate_one(c, ORC_NAND, ai, bi[0], na, na[n], 0, x1, x1 + n);
            bingate_one(c, ORC_NAND, ai + n, bi[1], ai + 2 * n, bi[2], 0, x2, x2 + n);
            bingate_one(c, ORC_NAND, x1, x1[n], x2, x2[n], 0, oa + i * n, ob + i);
            free(t);
        } else {
            u64* s = calloc(n + q + N + 1, sizeof(u64));
            u64 *lut = s + n, *ext = lut + q, bs = 0;
            for (u32 j = 0; j < k; ++j) {
                for (u32 t = 0; t < n; ++t) s[t] = addmod(s[t], ai[j * n + t], q);
                bs = addmod(bs, bi[j], q);
            }
            gate_lut(c, gate, q, p, lut);
            orc_bootstrap_core(c, s, bs, q, lut, ext);
            gate_finish(c, ext, p, 0, oa + i * n, ob + i);
            free(s);
        }
    }
    return rc;
}
/* binfhe-base-scheme.cpp:190-220 */
int orc_bootstrap(OrcCtx* c, const u64* a, const u64* b, u64* oa, u64* ob, u64 count, u64 p, int extended,
                  int threads) {
    u32 n = c->n, N = c->N, on = extended ? N : n; u64 q = c->q;
#pragma omp parallel for num_threads(threads > 0 ? threads : 1) schedule(dynamic)
    for (u64 i = 0; i < count; ++i) {
        u64* lut = malloc(sizeof(u64) * (q + N + 1));
        u64* ext = lut + q;
        /* BootstrapGateCore gets a COPY of ct (:200) and the copy constructor drops m_p (lwe-ciphertext.h:56):
         * the table is built with p = 4; the caller's p only enters the offset (:214) */
        gate_lut(c, ORC_AND, q, 4, lut);
        orc_bootstrap_core(c, a + i * n, addmod(b[i], q >> 2, q), q, lut, ext);
        gate_finish(c, ext, p, extended, oa + i * on, ob + i);
        free(lut);
    }
    return 0;
}
/* BootstrapFunc (binfhe-base-scheme.cpp:630-648) with f given as a table over x in [0, ctmod), values in Z_fmod */
static void bootstrap_func(OrcCtx* c, const u64* a, u64 b, u64 ctmod, const u64* ftab, u64 fmod, u64* oa, u64* ob) {
    u64* lut = malloc(sizeof(u64) * (ctmod + c->N + 1));
    u64* ext = lut + ctmod;
    for (u64 x = 0; x < ctmod; ++x) lut[x] = c->Q / fmod * ftab[x];
    orc_bootstrap_core(c, a, b, ctmod, lut, ext);
    switch_to_qn(c, ext, fmod, oa, ob);
    free(lut);
}
/* binfhe-base-scheme.h:245-260 */
static int check_input_function(const u64* lut, u64 len, u64 mod) {
    u64 mid = len / 2;
    if (lut[0] == mod - lut[mid]) {
        for (u64 i = 1; i < mid; ++i) if (lut[i] != mod - lut[mid + i]) return 2;
        return 0;
    }
    if (lut[0] == lut[mid]) {
        for (u64 i = 1; i < mid; ++i) if (lut[i] != lut[mid + i]) return 2;
        return 1;
    }
    return 2;
}
/* binfhe-base-scheme.cpp:241-332 */
static int evalfunc_one(OrcCtx* c, const u64* a, u64 b, const u64* LUT, u64 lutlen, u64 beta, u64 q, u64* oa, u64* ob) {
    u32 n = c->n;
    int prop = check_input_function(LUT, lutlen, q);
    if (prop == 0) {
        bootstrap_func(c, a, addmod(b, beta, q), q, LUT, q, oa, ob);
        return 0;
    }
    u64 dq = q << 1;
    u64 wm = prop == 2 ? dq : q; /* working modulus */
    if (prop == 2 && q > c->N) return -1;
    u64* f = malloc(sizeof(u64) * wm * 2);
    u64* f2 = f + wm;
    for (u64 x = 0; x < wm; ++x) f[x] = x < (wm >> 1) ? wm - (wm >> 2) : (wm >> 2); /* f0 */
    u64* a3 = malloc(sizeof(u64) * (n + 1));
    u64 b3;
    bootstrap_func(c, a, addmod(b, beta, wm), wm, f, wm, a3, &b3);
    /* ct3 = ct1 - ct3 (EvalSubEq2), + beta, - (q>>1 | q>>2) */
    for (u32 i = 0; i < n; ++i) a3[i] = submod(a[i], a3[i], wm);
    b3 = submod(b, b3, wm);
    b3 = addmod(b3, beta, wm);
    b3 = submod(b3, prop == 2 ? (q >> 1) : (q >> 2), wm);
    if (prop == 2) { /* LUT2 = LUT || LUT */
        for (u64 x = 0; x < wm; ++x)
            f2[x] = x < (wm >> 1) ? LUT[x % lutlen] : wm - LUT[(x - wm / 2) % lutlen];
    } else {
        for (u64 x = 0; x < wm; ++x) f2[x] = x < (wm >> 1) ? LUT[x] : wm - LUT[x - wm / 2];
    }
    bootstrap_func(c, a3, b3, wm, f2, wm, oa, ob);
    if (prop == 2) { /* ct4->SetModulus(q): lwe-ciphertext.h:116-120 reduces a and b mod q */
        for (u32 i = 0; i < n; ++i) oa[i] %= q;
        *ob %= q;
    }
    free(a3); free(f);
    return 0;
}
int orc_evalfunc(OrcCtx* c, const u64* a, const u64* b, const u64* lut, u64 lutlen, u64 beta, u64* oa, u64* ob,
                 u64 count, u64 ctmod, int threads) {
    u32 n = c->n;
    int rc = 0;
#pragma omp parallel for num_threads(threads > 0 ? threads : 1) schedule(dynamic)
    for (u64 i = 0; i < count; ++i)
        if (evalfunc_one(c, a + i * n, b[i], lut, lutlen, beta, ctmod, oa + i * n, ob + i)) rc = -1;
    return rc;
}
/* ---- large precision: EvalFloor / EvalSign / EvalDecomp (binfhe-base-scheme.cpp:335-521) ------------------------- */
u64 orc_ctx_baseG(OrcCtx* c) { return c->baseG; }
/* binfhe-base-scheme.cpp:335-377; ciphertext (a, b) mod `mod`, in place */
static void floor_one(OrcCtx* c, u64* a, u64* b, u64 mod, u64 beta, u64 qs) {
    u32 n = c->n;
    u64* w = malloc(sizeof(u64) * (2 * (n + 1) + qs));
    u64 *r = w + n + 1, *f = r + n + 1;
    *b = addmod(*b, beta % mod, mod);                              /* EvalAddConstEq(ct1, beta) */
    for (int pass = 0; pass < 2; ++pass) {
        for (u32 i = 0; i < n; ++i) w[i] = a[i] % qs;              /* ct1Modq: SetModulus(q) */
        u64 wb = *b % qs, rb;
        for (u64 x = 0; x < qs; ++x) {
            if (pass == 0) f[x] = x < (qs >> 1) ? mod - (qs >> 2) : (qs >> 2);
            else f[x] = x < (qs >> 2) ? mod - (qs >> 1) - x : (x < 3 * (qs >> 2) ? x : mod + (qs >> 1) - x);
        }
        bootstrap_func(c, w, wb, qs, f, mod, r, &rb);
        for (u32 i = 0; i < n; ++i) a[i] = submod(a[i], r[i], mod); /* EvalSubEq(ct1, ct2) */
        *b = submod(*b, rb, mod);
    }
    free(w);
}
int orc_evalfloor(OrcCtx* c, const u64* a, const u64* b, u64* oa, u64* ob, u64 count, u64 mod, u64 beta, u32 roundbits,
                  int threads) {
    u32 n = c->n;
    u64 qs = roundbits == 0 ? c->q : beta * ((u64)1 << (roundbits + 1));
#pragma omp parallel for num_threads(threads > 0 ? threads : 1) schedule(dynamic)
    for (u64 i = 0; i < count; ++i) {
        memcpy(oa + i * n, a + i * n, sizeof(u64) * n);
        ob[i] = b[i];
        floor_one(c, oa + i * n, ob + i, mod, beta, qs);
    }
    return 0;
}
static OrcCtx* find_base(OrcCtx** cs, int ncs, u64 base) {
    for (int i = 0; i < ncs; ++i) if (cs[i]->baseG == base) return cs[i];
    return NULL;
}
static u64 dynamic_base(u64 mod, int sign_variant) {
    u32 binLog = sign_variant ? (u32)ceil((double)(64 - __builtin_clzll(mod)) - 1) : (u32)ceil(log2((double)mod));
    if (binLog <= 17) return (u64)1 << 27;
    if (binLog <= 26) return (u64)1 << 18;
    return 0;
}
u32 orc_evaldecomp_count(u64 q, u64 mod, u64 beta) {
    u32 k = 1;
    while (mod > q && k < 64) { mod = mod / q * 2 * beta; ++k; }
    return k;
}
/* what 0: EvalSign (:380-455), 1: EvalSign(schemeSwitch), 2: EvalDecomp (:458-521; out [k][count][n], out_mod [k]).
 * cs[0] = the context of the current gadget base; ncs == 3 = the reference's dynamic mode. Returns outputs per input. */
int orc_evalsign_decomp(OrcCtx** cs, int ncs, const u64* a, const u64* b, u64* oa, u64* ob, u64* out_mod, u64 count,
                        u64 mod0, u64 beta, int what, int threads) {
    OrcCtx* c0 = cs[0];
    u32 n = c0->n;
    u64 q = c0->q;
    if (mod0 <= q) return -1;
    int rc = what == 2 ? (int)orc_evaldecomp_count(q, mod0, beta) : 1;
#pragma omp parallel for num_threads(threads > 0 ? threads : 1) schedule(dynamic)
    for (u64 i = 0; i < count; ++i) {
        u64* t = malloc(sizeof(u64) * (n + 1));
        memcpy(t, a + i * n, sizeof(u64) * n);
        u64 tb = b[i], mod = mod0;
        OrcCtx* cur = c0;
        u32 k = 0;
        while (mod > q) {
            if (what == 2) {
                for (u32 j = 0; j < n; ++j) oa[((u64)k * count + i) * n + j] = t[j] % q;
                ob[(u64)k * count + i] = tb % q;
                if (i == 0) out_mod[k] = q;
                ++k;
            }
            floor_one(cur, t, &tb, mod, beta, q);
            u64 mod_new = what == 2 ? mod / q * 2 * beta : (mod << 1) * beta / q;
            for (u32 j = 0; j < n; ++j) t[j] = round_qQ(t[j], mod_new, mod);   /* ModSwitch(mod, cttmp) */
            tb = round_qQ(tb, mod_new, mod);
            mod = mod_new;
            if (ncs == 3) {
                u64 base = dynamic_base(mod, what != 2);
                if (base != 0) { OrcCtx* nx = find_base(cs, ncs, base); if (nx) cur = nx; }
            }
        }
        if (what == 2) {
            memcpy(oa + ((u64)k * count + i) * n, t, sizeof(u64) * n);
            ob[(u64)k * count + i] = tb;
            if (i == 0) out_mod[k] = mod;
        } else {
            tb = addmod(tb, beta % mod, mod);
            u64* f = malloc(sizeof(u64) * mod);
            for (u64 x = 0; x < mod; ++x) {
                int low = x < mod / 2;
                f[x] = ((what == 0) == low) ? q / 4 : q - q / 4;
            }
            bootstrap_func(cur, t, tb, mod, f, q, oa + i * n, ob + i);
            if (what == 0) ob[i] = submod(ob[i], q >> 2, q);        /* EvalSubConstEq(cttmp, q >> 2) */
            free(f);
        }
        free(t);
    }
    return rc;
}

/* ---- RNS basis conversion: DCRTPolyImpl::ApproxSwitchCRTBasis / ApproxModUp / ApproxModDown (dcrtpoly-impl.h:910-1030) ----- */
static u64 prod_mod_except(const u64* m, u32 cnt, u32 skip, u64 mod) {
    u64 r = 1 % mod;
    for (u32 k = 0; k < cnt; ++k) if (k != skip) r = mulmod(r, m[k] % mod, mod);
    return r;
}
/* x [count][sizeQ][N] residues mod q_i (COEFFICIENT) -> out [count][sizeP][N]: out_j = sum_i [x_i (Q/q_i)^-1]_{q_i} (Q/q_i) mod p_j;
 * the reference sums the 128-bit products and reduces once (:921-936) — the same integer, reduced term by term here */
int orc_approx_switch_crt_basis(u32 N, u32 sizeQ, const u64* q, u32 sizeP, const u64* p, const u64* x, u64* out, u64 count) {
    u64* hinv = malloc(sizeof(u64) * sizeQ);
    u64* hat = malloc(sizeof(u64) * sizeQ * sizeP);
    for (u32 i = 0; i < sizeQ; ++i) {
        hinv[i] = invmod(prod_mod_except(q, sizeQ, i, q[i]), q[i]);
        for (u32 j = 0; j < sizeP; ++j) hat[i * sizeP + j] = prod_mod_except(q, sizeQ, i, p[j]);
    }
#pragma omp parallel for
    for (u64 c = 0; c < count; ++c)
        for (u32 k = 0; k < N; ++k)
            for (u32 j = 0; j < sizeP; ++j) {
                u64 s = 0;
                for (u32 i = 0; i < sizeQ; ++i) {
                    u64 xh = mulmod(x[(c * sizeQ + i) * N + k], hinv[i], q[i]);
                    s = addmod(s, mulmod(xh % p[j], hat[i * sizeP + j], p[j]), p[j]);
                }
                out[(c * sizeP + j) * N + k] = s;
            }
    free(hinv); free(hat);
    return 0;
}
/* ApproxModUp (:960-986): EVALUATION x over Q -> Q limbs unchanged, then the P limbs (INTT, basis switch, NTT) */
int orc_approx_mod_up(u32 N, u32 sizeQ, const u64* q, u32 sizeP, const u64* p, const u64* x, u64* out, u64 count) {
    u64* t = malloc(sizeof(u64) * count * sizeQ * N);
    u64* s = malloc(sizeof(u64) * count * sizeP * N);
    memcpy(t, x, sizeof(u64) * count * sizeQ * N);
    for (u64 c = 0; c < count; ++c)
        for (u32 i = 0; i < sizeQ; ++i) orc_ntt(q[i], orc_root_of_unity(2 * (u64)N, q[i]), N, t + (c * sizeQ + i) * N, 1, 1, 1);
    orc_approx_switch_crt_basis(N, sizeQ, q, sizeP, p, t, s, count);
    for (u64 c = 0; c < count; ++c) {
        for (u32 j = 0; j < sizeP; ++j) orc_ntt(p[j], orc_root_of_unity(2 * (u64)N, p[j]), N, s + (c * sizeP + j) * N, 1, 0, 1);
        memcpy(out + c * (sizeQ + sizeP) * N, x + c * sizeQ * N, sizeof(u64) * sizeQ * N);
        memcpy(out + (c * (sizeQ + sizeP) + sizeQ) * N, s + c * sizeP * N, sizeof(u64) * sizeP * N);
    }
    free(t); free(s);
    return 0;
}
/* ApproxModDown (:988-1030): EVALUATION x over Q u P -> (x_i - Switch_{P->Q}(x_P)_i) P^-1 mod q_i.  ptm = 0: CKKS; ptm > 0
 * (BGV, plaintext modulus t): the P part is multiplied by t^-1 mod p_j before the basis switch (:1003-1005, tInvModp =
 * t.ModInverse(p_j), bgvrns-cryptoparameters.cpp:78-83) and the switched part by t mod q_i before its NTT (:1021-1023) */
int orc_approx_mod_down_t(u32 N, u32 sizeQ, const u64* q, u32 sizeP, const u64* p, const u64* x, u64* out, u64 count, u64 ptm) {
    u64* t = malloc(sizeof(u64) * count * sizeP * N);
    u64* s = malloc(sizeof(u64) * count * sizeQ * N);
    for (u64 c = 0; c < count; ++c)
        for (u32 j = 0; j < sizeP; ++j) {
            u64* tj = t + (c * sizeP + j) * N;
            memcpy(tj, x + (c * (sizeQ + sizeP) + sizeQ + j) * N, sizeof(u64) * N);
            orc_ntt(p[j], orc_root_of_unity(2 * (u64)N, p[j]), N, tj, 1, 1, 1);
            if (ptm > 0) {
                const u64 tinv = invmod(ptm % p[j], p[j]);
                for (u32 k = 0; k < N; ++k) tj[k] = mulmod(tj[k], tinv, p[j]);
            }
        }
    orc_approx_switch_crt_basis(N, sizeP, p, sizeQ, q, t, s, count);
    for (u64 c = 0; c < count; ++c)
        for (u32 i = 0; i < sizeQ; ++i) {
            u64* si = s + (c * sizeQ + i) * N;
            if (ptm > 0)
                for (u32 k = 0; k < N; ++k) si[k] = mulmod(si[k], ptm % q[i], q[i]);
            orc_ntt(q[i], orc_root_of_unity(2 * (u64)N, q[i]), N, si, 1, 0, 1);
            u64 pinv = 1;
            for (u32 j = 0; j < sizeP; ++j) pinv = mulmod(pinv, p[j] % q[i], q[i]);
            pinv = invmod(pinv, q[i]);
            for (u32 k = 0; k < N; ++k)
                out[(c * sizeQ + i) * N + k] = mulmod(submod(x[(c * (sizeQ + sizeP) + i) * N + k], si[k], q[i]), pinv, q[i]);
        }
    free(t); free(s);
    return 0;
}
/* KeySwitchHYBRID::EvalFastKeySwitchCoreExt (src/pke/lib/keyswitch/keyswitch-hybrid.cpp:438-481): over the basis Ql u P (the
 * first sizeQl moduli of Q, then P), out_c[i] = sum_j digits_j[i] * evk_c[j][idx(i)] mod m_i with c = 0 the key's b vector and
 * c = 1 its a vector, idx(i) = i for a Q limb and sizeQ + (i - sizeQl) for a P limb (the key lives on the full Q u P).
 * digits [count][nd][sizeQl + sizeP][N], evk_b / evk_a [nd][sizeQ + sizeP][N], out [count][2][sizeQl + sizeP][N] */
int orc_keyswitch_inner(u32 N, u32 sizeQ, const u64* q, u32 sizeP, const u64* p, u32 sizeQl, u32 nd, const u64* digits,
                        const u64* evk_b, const u64* evk_a, u64* out, u64 count) {
    const u32 L = sizeQl + sizeP, LK = sizeQ + sizeP;
    for (u64 c = 0; c < count; ++c)
        for (u32 i = 0; i < L; ++i) {
            const u64 m = i < sizeQl ? q[i] : p[i - sizeQl];
            const u32 idx = i < sizeQl ? i : sizeQ + (i - sizeQl);
            for (u32 k = 0; k < N; ++k) {
                u64 s0 = 0, s1 = 0;
                for (u32 j = 0; j < nd; ++j) {
                    const u64 d = digits[((c * nd + j) * L + i) * N + k];
                    s0 = addmod(s0, mulmod(d, evk_b[((u64)j * LK + idx) * N + k], m), m);
                    s1 = addmod(s1, mulmod(d, evk_a[((u64)j * LK + idx) * N + k], m), m);
                }
                out[((c * 2 + 0) * L + i) * N + k] = s0;
                out[((c * 2 + 1) * L + i) * N + k] = s1;
            }
        }
    return 0;
}
/* Hybrid key switching of one polynomial (KeySwitchHYBRID::EvalKeySwitchPrecomputeCore, keyswitch-hybrid.cpp:328-412, then
 * EvalFastKeySwitchCore, :414-436).  c [count][sizeQl][N] EVALUATION over Ql = q[0 .. sizeQl); the digits are the parts of
 * `alpha` consecutive limbs, each raised to Ql u P by an approximate basis switch from its own limbs to all the others (:377-403);
 * inner product with the key's b / a vectors [numPart][sizeQ + sizeP][N]; both results brought down by P (ptm as in
 * orc_approx_mod_down_t).  out [count][2][sizeQl][N]. */
int orc_hybrid_keyswitch(u32 N, u32 sizeQ, const u64* q, u32 sizeP, const u64* p, u32 alpha, u32 sizeQl, const u64* c, const u64* evk_b,
                         const u64* evk_a, u64* out, u64 count, u64 ptm) {
    const u32 L = sizeQl + sizeP;
    u32 nd = (sizeQl + alpha - 1) / alpha, ndmax = (sizeQ + alpha - 1) / alpha;
    if (nd > ndmax) nd = ndmax;
    u64* digits = malloc(sizeof(u64) * count * nd * L * N);
    u64* tilda = malloc(sizeof(u64) * count * 2 * L * N);
    u64 *src = malloc(sizeof(u64) * (sizeQ + 1)), *dst = malloc(sizeof(u64) * (sizeQ + sizeP + 1));
    for (u32 part = 0; part < nd; ++part) {
        const u32 start = alpha * part, sp = (part == nd - 1) ? sizeQl - start : alpha, end = start + sp, sc = L - sp;
        for (u32 i = 0; i < sp; ++i) src[i] = q[start + i];
        for (u32 i = 0, k = 0; i < L; ++i)
            if (i < start || i >= end) dst[k++] = i < sizeQl ? q[i] : p[i - sizeQl];
        u64* t = malloc(sizeof(u64) * count * sp * N);
        u64* s = malloc(sizeof(u64) * count * sc * N);
        for (u64 ct = 0; ct < count; ++ct)
            for (u32 i = 0; i < sp; ++i) {
                memcpy(t + (ct * sp + i) * N, c + (ct * sizeQl + start + i) * N, sizeof(u64) * N);
                orc_ntt(src[i], orc_root_of_unity(2 * (u64)N, src[i]), N, t + (ct * sp + i) * N, 1, 1, 1);
            }
        orc_approx_switch_crt_basis(N, sp, src, sc, dst, t, s, count);
        for (u64 ct = 0; ct < count; ++ct) {
            u64* dg = digits + (ct * nd + part) * L * N;
            for (u32 i = 0, k = 0; i < L; ++i) {
                if (i >= start && i < end) {
                    memcpy(dg + (u64)i * N, c + (ct * sizeQl + i) * N, sizeof(u64) * N);
                } else {
                    u64* sk = s + (ct * sc + k) * N;
                    orc_ntt(dst[k], orc_root_of_unity(2 * (u64)N, dst[k]), N, sk, 1, 0, 1);
                    memcpy(dg + (u64)i * N, sk, sizeof(u64) * N);
                    ++k;
                }
            }
        }
        free(t); free(s);
    }
    orc_keyswitch_inner(N, sizeQ, q, sizeP, p, sizeQl, nd, digits, evk_b, evk_a, tilda, count);
    orc_approx_mod_down_t(N, sizeQl, q, sizeP, p, tilda, out, 2 * count, ptm);
    free(digits); free(tilda); free(src); free(dst);
    return 0;
}
int orc_approx_mod_down(u32 N, u32 sizeQ, const u64* q, u32 sizeP, const u64* p, const u64* x, u64* out, u64 count) {
    return orc_approx_mod_down_t(N, sizeQ, q, sizeP, p, x, out, count, 0);
}

double orc_time_bingate(OrcCtx* c, int gate, const u64* a1, const u64* b1, const u64* a2, const u64* b2, u64* oa,
                        u64* ob, u64 count, int threads) {
    struct timespec t0, t1;
    clock_gettime(CLOCK_MONOTONIC, &t0);
    orc_evalbingate(c, gate, a1, b1, a2, b2, oa, ob, count, 0, threads);
    clock_gettime(CLOCK_MONOTONIC, &t1);
    return (t1.tv_sec - t0.tv_sec) + 1e-9 * (t1.tv_nsec - t0.tv_nsec);
}
int orc_max_threads(void) { return omp_get_max_threads(); }
