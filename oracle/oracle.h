/* oracle/oracle.h — TEST INFRASTRUCTURE ONLY (CPU restatement of the reference's hot path).
 *
 * Plain-C restatement of luxcpp/fhe's batched-bootstrapping path and bit-reverse NTT, used as the
 * parity checker for the CUDA library.  NOT part of the product: only tests/, __graft_entry__.smoke()
 * and bench.py's cpu_baseline / --impl reference legs may load liboracle.so.
 *
 * Parity status: PINNED by execution — tests/test_oracle_vs_reference.py compares every function here
 * against the real reference (oracle/_ref/libref_shim.so, built from /root/reference by oracle/Makefile)
 * on the reference's own keys, and tests/golden/ holds small vectors produced by the reference
 * (tests/golden/make_golden.py) so the pin also holds where /root/reference is absent.
 *
 * Each function cites the reference file:line (relative to /root/reference) it restates.
 */
#ifndef LUX_ORACLE_H
#define LUX_ORACLE_H
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

enum { ORC_AP = 1, ORC_GINX = 2, ORC_LMKCDEY = 3 }; /* BINFHE_METHOD, binfhe-constants.h:112-117 */
/* BINGATE, binfhe-constants.h:123 */
enum { ORC_OR, ORC_AND, ORC_NOR, ORC_NAND, ORC_XOR, ORC_XNOR, ORC_MAJORITY, ORC_AND3, ORC_OR3, ORC_AND4, ORC_OR4,
       ORC_XOR_FAST, ORC_XNOR_FAST, ORC_CMUX };

typedef struct OrcCtx OrcCtx;

/* number theory: nbtheory-impl.h:183-231 (RootOfUnity), :350-373 (LastPrime), :386-393 (PreviousPrime) */
uint64_t orc_last_prime(uint32_t bits, uint64_t m);
uint64_t orc_previous_prime(uint64_t q, uint64_t m);
uint64_t orc_root_of_unity(uint64_t m, uint64_t q);
int orc_dcrt_chain(uint32_t N, uint32_t L, uint32_t bits, uint64_t* q, uint64_t* psi); /* ildcrtparams.h:100-118 */

/* NTT: transformnat-impl.h:302-373 (forward), :511-624 (inverse), tables :777-831. data [count][N] in place */
int orc_ntt(uint64_t Q, uint64_t psi, uint32_t N, uint64_t* data, uint64_t count, int inverse, int threads);

/* context */
OrcCtx* orc_ctx_create(uint32_t n, uint32_t N, uint64_t q, uint64_t Q, uint64_t qKS, uint32_t baseKS, uint32_t baseG,
                       uint32_t baseR, uint32_t numAutoKeys, int method);
void orc_ctx_destroy(OrcCtx*);
/* out: n N q Q qKS baseKS baseG baseR digitsG numAutoKeys 0 psi digitsKS method digitsR */
int orc_ctx_params(OrcCtx*, uint64_t* out);
uint64_t orc_bsk_len(OrcCtx*); /* number of u64 in the canonical flat BSK */
uint64_t orc_ksk_len(OrcCtx*); /* number of u64 in KSK A (B has len/n) */
int orc_load_bsk(OrcCtx*, const uint64_t* bsk);
int orc_load_ksk(OrcCtx*, const uint64_t* A, const uint64_t* B);
/* own key generation (deterministic from seed; NOT the reference's PRNG — the bootstrap path draws no
 * randomness, so parity does not depend on it). sk [n] signed. key_dist 0 ternary, 1 gaussian */
int orc_keygen(OrcCtx*, uint64_t seed, int key_dist, int64_t* sk, uint64_t* bsk, uint64_t* kskA, uint64_t* kskB);
int orc_encrypt(OrcCtx*, const int64_t* sk, int64_t m, uint64_t p, uint64_t mod, uint64_t seed, uint64_t* a,
                uint64_t* b);
int64_t orc_decrypt(OrcCtx*, const int64_t* sk, const uint64_t* a, uint64_t b, uint32_t len, uint64_t mod,
                    uint64_t p);

/* hot path */
int orc_decompose(OrcCtx*, const uint64_t* in2N, uint64_t* out);                 /* rgsw-acc.cpp:54-91 */
int orc_evalacc(OrcCtx*, const uint64_t* a, uint64_t amod, uint64_t* acc2N);    /* rgsw-acc-{cggi,dm,lmkcdey}.cpp */
int orc_modswitch(uint64_t qto, uint64_t qfrom, const uint64_t* in, uint64_t* out, uint64_t len); /* lwe-pke.cpp:41-46,254-261 */
int orc_keyswitch(OrcCtx*, const uint64_t* a, uint64_t b, uint64_t* oa, uint64_t* ob); /* lwe-pke.cpp:348-372 */
/* generic bootstrap: test vector m[j*factor] = lut[(b - j) mod ctmod]; out (N+1): a' then b (no offset added) */
int orc_bootstrap_core(OrcCtx*, const uint64_t* a, uint64_t b, uint64_t ctmod, const uint64_t* lut, uint64_t* ext);
/* batch entry points; threads = OpenMP threads over independent ciphertexts */
int orc_evalbingate(OrcCtx*, int gate, const uint64_t* a1, const uint64_t* b1, const uint64_t* a2,
                    const uint64_t* b2, uint64_t* oa, uint64_t* ob, uint64_t count, int extended, int threads);
int orc_evalbingate_multi(OrcCtx*, int gate, uint32_t k, const uint64_t* a, const uint64_t* b, uint64_t* oa,
                          uint64_t* ob, uint64_t count, uint64_t p, int threads);
int orc_bootstrap(OrcCtx*, const uint64_t* a, const uint64_t* b, uint64_t* oa, uint64_t* ob, uint64_t count,
                  uint64_t p, int extended, int threads);
int orc_evalfunc(OrcCtx*, const uint64_t* a, const uint64_t* b, const uint64_t* lut, uint64_t lutlen, uint64_t beta,
                 uint64_t* oa, uint64_t* ob, uint64_t count, uint64_t ctmod, int threads);
/* large precision (binfhe-base-scheme.cpp:335-521): ciphertexts (n, mod); one context per gadget base, cs[0] current */
uint64_t orc_ctx_baseG(OrcCtx*);
int orc_evalfloor(OrcCtx*, const uint64_t* a, const uint64_t* b, uint64_t* oa, uint64_t* ob, uint64_t count, uint64_t mod,
                  uint64_t beta, uint32_t roundbits, int threads);
uint32_t orc_evaldecomp_count(uint64_t q, uint64_t mod, uint64_t beta);
int orc_evalsign_decomp(OrcCtx** cs, int ncs, const uint64_t* a, const uint64_t* b, uint64_t* oa, uint64_t* ob,
                        uint64_t* out_mod, uint64_t count, uint64_t mod, uint64_t beta, int what, int threads);
/* RNS basis conversion (dcrtpoly-impl.h:910-1030); data [count][limbs][N] */
int orc_approx_switch_crt_basis(uint32_t N, uint32_t sizeQ, const uint64_t* q, uint32_t sizeP, const uint64_t* p,
                                const uint64_t* x, uint64_t* out, uint64_t count);
int orc_approx_mod_up(uint32_t N, uint32_t sizeQ, const uint64_t* q, uint32_t sizeP, const uint64_t* p, const uint64_t* x,
                      uint64_t* out, uint64_t count);
int orc_approx_mod_down(uint32_t N, uint32_t sizeQ, const uint64_t* q, uint32_t sizeP, const uint64_t* p, const uint64_t* x,
                        uint64_t* out, uint64_t count);
int orc_keyswitch_inner(uint32_t N, uint32_t sizeQ, const uint64_t* q, uint32_t sizeP, const uint64_t* p, uint32_t sizeQl, uint32_t nd,
                        const uint64_t* digits, const uint64_t* evk_b, const uint64_t* evk_a, uint64_t* out, uint64_t count);
int orc_hybrid_keyswitch(uint32_t N, uint32_t sizeQ, const uint64_t* q, uint32_t sizeP, const uint64_t* p, uint32_t alpha, uint32_t sizeQl,
                         const uint64_t* c, const uint64_t* evk_b, const uint64_t* evk_a, uint64_t* out, uint64_t count, uint64_t ptm);
int orc_approx_mod_down_t(uint32_t N, uint32_t sizeQ, const uint64_t* q, uint32_t sizeP, const uint64_t* p, const uint64_t* x,
                          uint64_t* out, uint64_t count, uint64_t ptm);   /* ptm > 0: the BGV branch */
double orc_time_bingate(OrcCtx*, int gate, const uint64_t* a1, const uint64_t* b1, const uint64_t* a2,
                        const uint64_t* b2, uint64_t* oa, uint64_t* ob, uint64_t count, int threads);
int orc_max_threads(void);

#ifdef __cplusplus
}
#endif
#endif
