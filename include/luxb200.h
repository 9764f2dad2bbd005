/* luxb200.h — C ABI of libluxb200.so: B200-native (sm_100a) batched TFHE/FHEW bootstrapping + bit-reverse NTT.
 *
 * Drop-in boundary for luxcpp/fhe's batched hot path.  Every entry point replaces a reference interface
 * (file:line relative to the reference tree):
 *
 *   luxb200_evalbingate_batch       lux::fhe::EvalBinGateBatch            src/binfhe/include/batch/binfhe-batch.h:100-111
 *                                   (= OMP loop over BinFHEContext::EvalBinGate, src/binfhe/lib/batch/batch.cpp:176-210;
 *                                   BinFHEScheme::EvalBinGate src/binfhe/lib/binfhe-base-scheme.cpp:76-126)
 *   luxb200_evalbingate_multi_batch BinFHEScheme::EvalBinGate(vector)     src/binfhe/lib/binfhe-base-scheme.cpp:129-187
 *                                   and lux::fhe::EvalCMUXBatch            binfhe-batch.h:113-125
 *   luxb200_bootstrap_batch         lux::fhe::BootstrapBatch               binfhe-batch.h:48-58 (spec: BinFHEScheme::Bootstrap,
 *                                                                          binfhe-base-scheme.cpp:190-220)
 *   luxb200_evalfunc_batch          lux::fhe::EvalFuncBatch                binfhe-batch.h:60-75 (BinFHEScheme::EvalFunc,
 *                                                                          binfhe-base-scheme.cpp:241-332)
 *   luxb200_blind_rotate_batch      backend::Backend::BlindRotateBatch     src/binfhe/include/backend/backend.h:177-183
 *                                   (RingGSWAccumulator::EvalAcc, src/binfhe/include/rgsw-acc.h:73-76)
 *   luxb200_keyswitch_batch         backend::Backend::KeySwitchBatch       backend.h:196-202 (LWEEncryptionScheme::KeySwitch,
 *                                                                          src/binfhe/lib/lwe-pke.cpp:348-372)
 *   luxb200_modswitch_batch         backend::Backend::ModSwitchBatch       backend.h:204-211 (lwe-pke.cpp:254-261)
 *   luxb200_ntt_batch               PolyImpl::SwitchFormat                 src/core/include/lattice/hal/default/poly-impl.h:420-440
 *                                   (NumberTheoreticTransformNat, src/core/include/math/hal/intnat/transformnat-impl.h:302-373,511-624)
 *   luxb200_dcrt_ntt_batch          DCRTPolyImpl::SwitchFormat             src/core/include/lattice/hal/default/dcrtpoly-impl.h:1953-1959
 *   luxb200_circuit_batch           lux::fhe::BatchDAG::Execute            binfhe-batch.h:255-279 (src/binfhe/lib/batch/batch.cpp:429-482,
 *                                                                          sequential there): gate DAG, level-scheduled on the device
 *   luxb200_load_bsk / _ksk         backend::Backend::PackBootstrappingKey backend.h:223-233 (flat layouts below)
 *   luxb200_ctx_create              BinFHEContext::GenerateBinFHEContext   src/binfhe/lib/binfhecontext.cpp:109-187 (parameters only)
 *
 * Conventions (same as the reference's C ABI, include/lux/fhe/c_api.h:47-70): every call returns 0 on
 * success or a negative LUXB200_ERR_* code; no exception crosses the boundary; luxb200_last_error() gives
 * the message of the calling thread's last failure.  Host buffers are caller-owned; ciphertext data uses
 * the reference's NativeVector element type (uint64_t, one residue per element).  There is NO CPU fallback:
 * without a CUDA device every compute call fails with LUXB200_ERR_NO_DEVICE.
 * Threading: a context (one device, one stream, one set of work buffers) serialises its entry points on an internal
 * mutex, so host threads may share it (the reference's std::async batch futures do); different contexts — e.g. one per
 * GPU — run concurrently.  An NTT plan is not locked: one thread per plan.
 *
 * Canonical flat key layouts (EVALUATION format, the reference's own coefficient order):
 *   BSK  GINX    [n][2][rows][2][N]              ([i][0] = (*ek)[0][0][i], [i][1] = (*ek)[0][1][i]; rgsw-acc-cggi.cpp:39-57)
 *        AP      [n][baseR][digitsR][rows][2][N] ((*ek)[i][j][k]; j = 0 unused; rgsw-acc-dm.cpp:39-60)
 *        LMKCDEY [n][rows][2][N] then [numAutoKeys+1][digitsG-1][2][N]  ((*ek)[0][0][i]; (*ek)[0][1][t]; rgsw-acc-lmkcdey.cpp:39-68)
 *        rows = 2*(digitsG-1)
 *   KSK  A [N][baseKS][digitsKS][n], B [N][baseKS][digitsKS]   (LWESwitchingKeyImpl, lwe-keyswitchkey.h:49-58)
 */
#ifndef LUXB200_H
#define LUXB200_H
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define LUXB200_OK 0
#define LUXB200_ERR_INVALID_ARG (-1)
#define LUXB200_ERR_NO_DEVICE (-2)
#define LUXB200_ERR_CUDA (-3)
#define LUXB200_ERR_UNSUPPORTED (-4)
#define LUXB200_ERR_NO_KEYS (-5)
#define LUXB200_ERR_ALLOC (-6)

/* BINFHE_METHOD (binfhe-constants.h:112-117) and BINGATE (:123) values */
enum { LUXB200_AP = 1, LUXB200_GINX = 2, LUXB200_LMKCDEY = 3 };
enum { LUXB200_OR, LUXB200_AND, LUXB200_NOR, LUXB200_NAND, LUXB200_XOR, LUXB200_XNOR, LUXB200_MAJORITY,
       LUXB200_AND3, LUXB200_OR3, LUXB200_AND4, LUXB200_OR4, LUXB200_XOR_FAST, LUXB200_XNOR_FAST, LUXB200_CMUX };

typedef struct luxb200_ctx luxb200_ctx;

typedef struct {
    uint32_t n, N;            /* LWE dimension, ring dimension */
    uint64_t q, Q, qKS;       /* LWE modulus, RLWE (NTT) modulus, key-switch modulus */
    uint32_t baseKS, baseG, baseR, numAutoKeys;
    uint32_t method;          /* LUXB200_AP / GINX / LMKCDEY */
    uint32_t reserved;
    uint64_t psi;             /* primitive 2N-th root of unity mod Q; 0 = compute the minimal one (nbtheory-impl.h:183-231) */
} luxb200_params;

const char* luxb200_last_error(void);
int luxb200_version(void);
int luxb200_device_count(void);

/* ---- number theory used for table generation (host) ------------------------------------------------ */
uint64_t luxb200_last_prime(uint32_t bits, uint64_t m);      /* LastPrime, nbtheory-impl.h:350-373 */
uint64_t luxb200_previous_prime(uint64_t q, uint64_t m);     /* PreviousPrime, nbtheory-impl.h:386-393 */
uint64_t luxb200_root_of_unity(uint64_t m, uint64_t q);      /* RootOfUnity, nbtheory-impl.h:183-231 */

/* ---- context ----------------------------------------------------------------------------------------- */
int luxb200_ctx_create(const luxb200_params* p, int device, luxb200_ctx** out);
void luxb200_ctx_destroy(luxb200_ctx* ctx);
/* out[16]: n N q Q qKS baseKS baseG baseR digitsG numAutoKeys 0 psi digitsKS method digitsR device */
int luxb200_ctx_info(luxb200_ctx* ctx, uint64_t* out);
uint64_t luxb200_bsk_len(luxb200_ctx* ctx);   /* elements of the canonical flat BSK */
uint64_t luxb200_ksk_len(luxb200_ctx* ctx);   /* elements of KSK A (B has ksk_len / n) */
/* keys in the canonical layout; on_device != 0: the pointers are device pointers on ctx's device
 * (e.g. the receive buffer of an NCCL broadcast) */
int luxb200_load_bsk(luxb200_ctx* ctx, const uint64_t* bsk, int on_device);
int luxb200_load_ksk(luxb200_ctx* ctx, const uint64_t* A, const uint64_t* B, int on_device);
/* BinFHEContext::BTKeyGen (binfhecontext.cpp:285-307; KeySwitchGen lwe-pke.cpp:264-344, KeyGenAcc rgsw-acc-{cggi,dm,
 * lmkcdey}.cpp) on the device, straight into the packed key layouts.  sk [n] signed LWE secret, seed = caller entropy
 * (ChaCha20 streams); skN_out [N] optional (the ring secret).  Keys are NOT bit-equal to reference-generated ones (other
 * PRNG); they have the same structure and noise distribution and interoperate: luxb200_export_keys returns them in the
 * canonical layouts, which the reference's evaluator accepts through BTKeyLoad. */
int luxb200_btkeygen(luxb200_ctx* ctx, const int64_t* sk, uint64_t seed, int64_t* skN_out);
int luxb200_export_keys(luxb200_ctx* ctx, uint64_t* bsk, uint64_t* kskA, uint64_t* kskB);
int luxb200_synchronize(luxb200_ctx* ctx);
/* One-time key exchange inside one process (the only exchange step of the path): ctxs[0] holds the keys,
 * ctxs[1..n-1] (same parameters, other GPUs of the box) receive the packed device images peer-to-peer over
 * NVLink/NVSwitch in ceil(log2 n) doubling rounds.  Multi-process jobs broadcast the canonical arrays with NCCL and load
 * them with on_device = 1 (fhe_b200/parallel.py). */
int luxb200_broadcast_keys(luxb200_ctx* const* ctxs, int n);
/* page-locked host staging for ciphertext batches (so the batch pipeline's copies overlap its kernels) */
int luxb200_host_alloc(uint64_t bytes, void** out);
int luxb200_host_free(void* ptr);

/* ---- batched operators, HOST buffers (H2D / D2H inside) ---------------------------------------------- */
/* a1,a2 [count][n], b1,b2 [count] mod q; out oa [count][n] (or [count][N] mod Q when extended), ob [count] */
int luxb200_evalbingate_batch(luxb200_ctx* ctx, int gate, const uint64_t* a1, const uint64_t* b1, const uint64_t* a2,
                              const uint64_t* b2, uint64_t* oa, uint64_t* ob, uint64_t count, int extended);
/* the same with the per-ciphertext semantics of binfhe-base-scheme.cpp:76-126 made explicit: p = plaintext modulus the
 * test vector is built for (BootstrapGateCore, :535-567; the reference itself always ends up with 4 here because it
 * bootstraps a copy of ct1 and the copy constructor drops m_p); in_ext bit 0 / 1: ct1 / ct2 arrive extended (a [count][N]
 * mod Q) and are run through SwitchCTtoqn first, as :92-93 does */
int luxb200_evalbingate_batch_ex(luxb200_ctx* ctx, int gate, const uint64_t* a1, const uint64_t* b1, const uint64_t* a2,
                                 const uint64_t* b2, uint64_t* oa, uint64_t* ob, uint64_t count, int extended, uint64_t p,
                                 int in_ext);
/* LWEEncryptionScheme::SwitchCTtoqn (lwe-pke.cpp:170-178): a [count][N] mod Q, b [count] -> oa [count][n], ob [count] mod q */
int luxb200_switch_to_qn_batch(luxb200_ctx* ctx, const uint64_t* a, const uint64_t* b, uint64_t* oa, uint64_t* ob,
                               uint64_t count);
/* a [count][k][n], b [count][k]; p = plaintext modulus of the inputs (GetptModulus) */
int luxb200_evalbingate_multi_batch(luxb200_ctx* ctx, int gate, uint32_t k, const uint64_t* a, const uint64_t* b,
                                    uint64_t* oa, uint64_t* ob, uint64_t count, uint64_t p);
int luxb200_bootstrap_batch(luxb200_ctx* ctx, const uint64_t* a, const uint64_t* b, uint64_t* oa, uint64_t* ob,
                            uint64_t count, uint64_t p, int extended);
/* lut [lutlen] over Z_ctmod (GenerateLUTviaFunction, binfhecontext.cpp:372-387); beta = BinFHEContext::GetBeta() */
int luxb200_evalfunc_batch(luxb200_ctx* ctx, const uint64_t* a, const uint64_t* b, const uint64_t* lut,
                           uint64_t lutlen, uint64_t beta, uint64_t* oa, uint64_t* ob, uint64_t count, uint64_t ctmod);
/* ---- large-precision floor / sign / digit decomposition (binfhe-base-scheme.cpp:335-521) ----------------------------
 * Ciphertexts a [count][n] mod `mod` (the large plaintext-space modulus, a power of two up to 2^29, binfhecontext.cpp:60),
 * beta = BinFHEContext::GetBeta().  One context per gadget base: ctxs[0] is the current base (RingGSWParams->GetBaseG()),
 * the others are the rest of the reference's key map (BTKeyGen with timeOptimization fills baseG = 2^14, 2^18, 2^27,
 * binfhecontext.cpp:292-298); n_ctx == 3 switches bases with the shrinking modulus exactly as :415-434 does. */
int luxb200_evalfloor_batch(luxb200_ctx* ctx, const uint64_t* a, const uint64_t* b, uint64_t* oa, uint64_t* ob, uint64_t count,
                            uint64_t mod, uint64_t beta, uint32_t roundbits);
int luxb200_evalsign_batch(luxb200_ctx* const* ctxs, int n_ctx, const uint64_t* a, const uint64_t* b, uint64_t* oa,
                           uint64_t* ob, uint64_t count, uint64_t mod, uint64_t beta, int scheme_switch);
/* out_a [n_out][count][n], out_b [n_out][count], out_mod [n_out] (modulus of each returned digit ciphertext);
 * n_out = luxb200_evaldecomp_count(q, mod, beta) <= max_out */
int luxb200_evaldecomp_batch(luxb200_ctx* const* ctxs, int n_ctx, const uint64_t* a, const uint64_t* b, uint64_t* out_a,
                             uint64_t* out_b, uint64_t* out_mod, uint32_t max_out, uint32_t* n_out, uint64_t count,
                             uint64_t mod, uint64_t beta);
uint32_t luxb200_evaldecomp_count(uint64_t q, uint64_t mod, uint64_t beta);
/* dst borrows src's device copy of the switching key (contexts that differ in baseG only) */
int luxb200_share_ksk(luxb200_ctx* dst, luxb200_ctx* src);
/* a [count][n] mod amod; acc [count][2][N] EVALUATION format, in place */
int luxb200_blind_rotate_batch(luxb200_ctx* ctx, const uint64_t* a, uint64_t amod, uint64_t* acc, uint64_t count);
/* backend::Backend::ExternalProductBatch (backend.h:185-194): RGSW x RLWE -> RLWE as one AddToAccDM step computes it
 * (rgsw-acc-dm.cpp:118-143).  rgsw [count][rows][2][N] (RingGSWEvalKeyImpl::GetElements order), rlwe / out [count][2][N];
 * EVALUATION format, residues mod Q */
int luxb200_external_product_batch(luxb200_ctx* ctx, const uint64_t* rgsw, const uint64_t* rlwe, uint64_t* out,
                                   uint64_t count);
/* a [count][N] mod qKS, b [count]  ->  oa [count][n], ob [count] mod qKS */
int luxb200_keyswitch_batch(luxb200_ctx* ctx, const uint64_t* a, const uint64_t* b, uint64_t* oa, uint64_t* ob,
                            uint64_t count);
/* elementwise RoundqQ (lwe-pke.cpp:41-46) over len values */
int luxb200_modswitch_batch(luxb200_ctx* ctx, uint64_t qto, uint64_t qfrom, const uint64_t* in, uint64_t* out,
                            uint64_t len);

/* ---- raw device memory: backend::Backend::Allocate / Free / CopyToDevice / CopyToHost / DeviceMemory
 * (backend.h:88-113).  kind: 1 host->device, 2 device->host, 3 device->device */
int luxb200_dev_alloc(int device, uint64_t bytes, void** out);
int luxb200_dev_free(int device, void* ptr);
int luxb200_dev_copy(int device, void* dst, const void* src, uint64_t bytes, int kind);
int luxb200_dev_mem_info(int device, uint64_t* free_bytes, uint64_t* total_bytes);

/* ---- NTT (context-free plan objects) ----------------------------------------------------------------- */
typedef struct luxb200_ntt_plan luxb200_ntt_plan;
/* L moduli q[l] with 2N-th roots psi[l] (psi may be NULL: minimal roots are computed). N = 2^k, 256..131072 */
int luxb200_ntt_plan_create(uint32_t N, uint32_t L, const uint64_t* q, const uint64_t* psi, int device,
                            luxb200_ntt_plan** out);
void luxb200_ntt_plan_destroy(luxb200_ntt_plan* plan);
/* data [count][L][N] uint64 host, in place; inverse = 0: COEFFICIENT->EVALUATION, 1: EVALUATION->COEFFICIENT */
int luxb200_ntt_batch(luxb200_ntt_plan* plan, uint64_t* data, uint64_t count, int inverse);
/* device-resident variants: data on the plan's device. elem_bytes 4 (only when every q < 2^30) or 8 */
int luxb200_ntt_batch_dev(luxb200_ntt_plan* plan, void* data_dev, uint64_t count, int inverse, int elem_bytes);
int luxb200_ntt_plan_sync(luxb200_ntt_plan* plan);
float luxb200_ntt_last_ms(luxb200_ntt_plan* plan);   /* CUDA-event time of the last *_dev call's kernels */
#define luxb200_dcrt_ntt_batch luxb200_ntt_batch     /* DCRTPoly::SwitchFormat = the L > 1 case */

/* ---- RNS basis conversion around the DCRT NTT (CKKS / BGV hybrid key switching) ---------------------------------------
 *   luxb200_approx_switch_crt_basis  DCRTPolyImpl::ApproxSwitchCRTBasis  src/core/include/lattice/hal/default/dcrtpoly-impl.h:910-957
 *   luxb200_approx_mod_switch what=0 DCRTPolyImpl::ApproxModUp           :960-986  (KeySwitchHYBRID::EvalKeySwitchPrecomputeCore,
 *                                                                         src/pke/lib/keyswitch/keyswitch-hybrid.cpp:328-412)
 *                             what=1 DCRTPolyImpl::ApproxModDown (t = 0) :988-1030 (EvalFastKeySwitchCore, :414-436)
 * A plan holds two bases Q = {q_i} and P = {p_j} (pairwise distinct primes below 2^62) and the tables derived from them
 * ((Q/q_i)^-1 mod q_i, (Q/q_i) mod p_j, P^-1 mod q_i, and the same with the roles swapped).  Data [count][limbs][N] uint64,
 * one canonical residue per element. */
typedef struct luxb200_rns_plan luxb200_rns_plan;
int luxb200_rns_plan_create(uint32_t N, uint32_t sizeQ, const uint64_t* q, uint32_t sizeP, const uint64_t* p, int device,
                            luxb200_rns_plan** out);
void luxb200_rns_plan_destroy(luxb200_rns_plan* plan);
/* COEFFICIENT-format limbs; direction 0: x over Q -> out over P, 1: x over P -> out over Q.  Host buffers. */
int luxb200_approx_switch_crt_basis(luxb200_rns_plan* plan, const uint64_t* x, uint64_t* out, uint64_t count, int direction);
/* EVALUATION format in and out; what = 0: x [count][sizeQ][N] -> out [count][sizeQ+sizeP][N] (Q limbs, then P limbs);
 * what = 1: x [count][sizeQ+sizeP][N] -> out [count][sizeQ][N].  Host buffers; the _dev variants take device pointers. */
int luxb200_approx_mod_switch(luxb200_rns_plan* plan, const uint64_t* x, uint64_t* out, uint64_t count, int what);
int luxb200_approx_mod_up_dev(luxb200_rns_plan* plan, const void* x_dev, void* out_dev, uint64_t count);
int luxb200_approx_mod_down_dev(luxb200_rns_plan* plan, const void* x_dev, void* out_dev, uint64_t count);
/* The BGV branch of DCRTPolyImpl::ApproxModDown (t > 0, dcrtpoly-impl.h:1003-1005 and :1021-1023 with tInvModp / tModqPrecon of
 * src/pke/lib/scheme/bgvrns/bgvrns-cryptoparameters.cpp:78-95): the P part is multiplied by t^-1 mod p_j before the basis switch
 * and the switched part by t mod q_i before its NTT.  t = 0 is the CKKS form above; `what` as in luxb200_approx_mod_switch
 * (t is ignored for what = 0). */
int luxb200_approx_mod_switch_bgv(luxb200_rns_plan* plan, const uint64_t* x, uint64_t* out, uint64_t count, int what, uint64_t t);
int luxb200_approx_mod_down_bgv_dev(luxb200_rns_plan* plan, const void* x_dev, void* out_dev, uint64_t count, uint64_t t);
/* KeySwitchHYBRID::EvalFastKeySwitchCoreExt (src/pke/lib/keyswitch/keyswitch-hybrid.cpp:438-481), the inner product of hybrid key
 * switching between ApproxModUp and ApproxModDown: over the basis Ql u P (the first sizeQl moduli of the plan's Q, then P),
 * out_c[i] = sum_j digits_j[i] * evk_c[j][idx(i)] mod m_i, c = 0 the key's b vector, c = 1 its a vector, idx(i) = i for a Q limb and
 * sizeQ + (i - sizeQl) for a P limb.  EVALUATION format; digits [count][nd][sizeQl + sizeP][N], evk_b / evk_a
 * [nd][sizeQ + sizeP][N], out [count][2][sizeQl + sizeP][N].  Host buffers, or device pointers for the _dev form. */
int luxb200_rns_keyswitch_inner(luxb200_rns_plan* plan, const uint64_t* digits, const uint64_t* evk_b, const uint64_t* evk_a,
                                uint64_t* out, uint64_t count, uint32_t nd, uint32_t sizeQl);
int luxb200_rns_keyswitch_inner_dev(luxb200_rns_plan* plan, const void* digits_dev, const void* evk_b_dev, const void* evk_a_dev,
                                    void* out_dev, uint64_t count, uint32_t nd, uint32_t sizeQl);
/* Hybrid key switching of one polynomial as one call: KeySwitchHYBRID::EvalKeySwitchPrecomputeCore (keyswitch-hybrid.cpp:328-412:
 * digits = parts of `alpha` consecutive limbs of Ql = q[0 .. sizeQl), each raised to Ql u P) and EvalFastKeySwitchCore (:414-436:
 * inner product with the key, ApproxModDown of both results; t > 0 = BGV).  c [count][sizeQl][N] EVALUATION, evk_b / evk_a
 * [luxb200_hks_num_digits][sizeQ + sizeP][N], out [count][2][sizeQl][N].  Host buffers. */
typedef struct luxb200_hks_plan luxb200_hks_plan;
int luxb200_hks_plan_create(uint32_t N, uint32_t sizeQ, const uint64_t* q, uint32_t sizeP, const uint64_t* p, uint32_t alpha,
                            uint32_t sizeQl, int device, luxb200_hks_plan** out);
void luxb200_hks_plan_destroy(luxb200_hks_plan* plan);
uint32_t luxb200_hks_num_digits(luxb200_hks_plan* plan);
int luxb200_hks_keyswitch(luxb200_hks_plan* plan, const uint64_t* c, const uint64_t* evk_b, const uint64_t* evk_a, uint64_t* out,
                          uint64_t count, uint64_t t);
/* the same on device pointers (plan's device; the key stays resident between calls); returns when the result is complete */
int luxb200_hks_keyswitch_dev(luxb200_hks_plan* plan, const void* c_dev, const void* evk_b_dev, const void* evk_a_dev, void* out_dev,
                              uint64_t count, uint64_t t);
/* Allocation-free variant of luxb200_approx_switch_crt_basis.  It has no counterpart in the reference and was added
 * for callers that cannot call cudaMalloc (code under CUDA graph capture, tight serving loops,
 * harnesses that forbid it).  Same arguments, same result; the device staging of the basis-switch kernel is a static device
 * array of LUXB200_STATIC_SCRATCH_BYTES that every GPU loading the library holds until exit, instead of the plan's growable
 * buffers, and the call makes no cudaMalloc / cudaFree.  count * (sizeQ + sizeP) * N * 8 bytes (input staging rounded up to
 * 256 bytes) must fit in the array: anything larger is refused with LUXB200_ERR_INVALID_ARG — split the batch or use the
 * allocating entry point.  All calls on one device share the array (it keeps the last call's contents; nothing in it needs to
 * start at zero), so calls that overlap on one GPU race: serialise them. */
#define LUXB200_STATIC_SCRATCH_BYTES (16ull << 20)
int luxb200_approx_switch_crt_basis_static(luxb200_rns_plan* plan, const uint64_t* x, uint64_t* out, uint64_t count,
                                           int direction);

/* ---- device-resident batched gates (inputs already in HBM as uint32) ---------------------------------- */
/* a1,a2 [count][n] u32, b1,b2 [count] u32 on ctx's device; out oa [count][n] u32, ob [count] u32. Asynchronous on the
 * context stream; call luxb200_synchronize before reading. */
int luxb200_evalbingate_batch_dev(luxb200_ctx* ctx, int gate, const uint32_t* a1, const uint32_t* b1,
                                  const uint32_t* a2, const uint32_t* b2, uint32_t* oa, uint32_t* ob, uint64_t count);
/* CUDA-event times (ms) of the kernels of the last gate/bootstrap call on this context:
 * out[0] prep, out[1] blind rotation, out[2] key switch, out[3] total; out[4] = number of kernel launches */
int luxb200_last_timing(luxb200_ctx* ctx, float* out);
/* cumulative number of kernels this library launched on the context (for bench.py's gpu_launches) */
uint64_t luxb200_launch_count(luxb200_ctx* ctx);

/* ---- gate circuits: lux::fhe::BatchDAG (binfhe-batch.h:255-279; Execute, batch.cpp:429-482) --------------------- */
/* One circuit applied to `count` independent input sets.  Wires 0 .. n_inputs-1 are the inputs, wire n_inputs + g is the
 * output of gates[g]; a gate may only read wires with a smaller id (the order BatchDAG::Add* produces).
 * op: a 2-input BINGATE code (LUXB200_OR .. LUXB200_XNOR, *_FAST) = BinFHEScheme::EvalBinGate on (in0, in1);
 *     LUXB200_OP_NOT = EvalNOT(in0) (binfhe-base-scheme.cpp:223-237, no bootstrap);
 *     LUXB200_OP_BOOTSTRAP = BinFHEScheme::Bootstrap(in0) with plaintext modulus 4 (binfhe-base-scheme.cpp:190-220).
 * Gates are grouped into dependency levels; one level = one gather kernel + one blind-rotation launch + one key-switch
 * launch over every (gate, input set) pair of the level, mixed gate kinds included.  Every intermediate ciphertext stays in
 * HBM; only the inputs go up and the wires listed in `outputs` come back.
 * in_a [n_inputs][count][n], in_b [n_inputs][count] mod q; out_a [n_outputs][count][n], out_b [n_outputs][count]. */
enum { LUXB200_OP_NOT = 100, LUXB200_OP_BOOTSTRAP = 101, LUXB200_OP_EVALFUNC = 102 };
typedef struct {
    uint32_t op;         /* BINGATE code or LUXB200_OP_* */
    uint32_t in0, in1;   /* wire ids; in1 ignored by the 1-input ops */
} luxb200_gate;
int luxb200_circuit_batch(luxb200_ctx* ctx, uint32_t n_inputs, const luxb200_gate* gates, uint32_t n_gates,
                          const uint32_t* outputs, uint32_t n_outputs, const uint64_t* in_a, const uint64_t* in_b,
                          uint64_t* out_a, uint64_t* out_b, uint64_t count);
/* The same with BatchOp::OP_EVAL_FUNC nodes (BatchDAG::AddEvalFunc, batch.cpp:455-460): op = LUXB200_OP_EVALFUNC,
 * in0 = the input wire, in1 = index of the node's look-up table in `luts` [n_luts][q] (values over Z_q as
 * GenerateLUTviaFunction writes them, binfhecontext.cpp:372-387); beta = BinFHEContext::GetBeta().  Per node exactly
 * BinFHEScheme::EvalFunc (binfhe-base-scheme.cpp:241-332): a negacyclic table is one bootstrap, a periodic or arbitrary one
 * two chained bootstraps (two levels); an arbitrary table with q > N is refused with the reference's message. */
int luxb200_circuit_batch_ex(luxb200_ctx* ctx, uint32_t n_inputs, const luxb200_gate* gates, uint32_t n_gates,
                             const uint32_t* outputs, uint32_t n_outputs, const uint64_t* luts, uint32_t n_luts,
                             uint64_t beta, const uint64_t* in_a, const uint64_t* in_b, uint64_t* out_a, uint64_t* out_b,
                             uint64_t count);
/* out[0] = dependency levels, out[1] = bootstrapped gates, out[2] = kernel launches of the last circuit call */
int luxb200_circuit_last_stats(luxb200_ctx* ctx, uint64_t* out);

/* ---- integer-pipe roofline probe: measured IMAD-class issue rate of this GPU (ops/s) ------------------- */
int luxb200_intpipe_probe(int device, double* imad_per_s, double* mixed_per_s);
/* 64-bit Shoup butterflies per second on register-resident data: the integer-pipe roofline of the u64 NTT
 * (replaces nothing in the reference; measurement only) */
int luxb200_intpipe_probe64(int device, double* bfly64_per_s);

#ifdef __cplusplus
}
#endif
#endif
