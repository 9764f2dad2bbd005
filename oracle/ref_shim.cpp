// oracle/ref_shim.cpp — TEST INFRASTRUCTURE ONLY.
//
// A flat C ABI over the *unmodified* reference library (oracle/_ref/libfheref.so, compiled by
// oracle/Makefile from the sources under /root/reference).  Nothing here is shipped or linked into
// libluxb200.so; only tests/, __graft_entry__.smoke() and bench.py's reference / cpu_baseline legs
// load it.  Every function drives the reference through its own public API:
//   BinFHEContext            src/binfhe/include/binfhecontext.h:130-447
//   BinFHEScheme / EvalAcc   src/binfhe/include/binfhe-base-scheme.h, rgsw-acc.h:60-99
//   LWEEncryptionScheme      src/binfhe/include/lwe-pke.h
//   ChineseRemainderTransformFTT / NativePoly   src/core/include/math/hal/intnat/transformnat.h
//   ILDCRTParams             src/core/include/lattice/ildcrtparams.h:100-118
#include "binfhecontext.h"
#include "lattice/lat-hal.h"
#include "math/math-hal.h"
#include "math/nbtheory.h"

#include <omp.h>
#include <chrono>
#include <cstdint>
#include <cstring>
#include <memory>
#include <vector>

using namespace lux::fhe;

struct RefCtx {
    BinFHEContext cc;
    LWEPrivateKey sk;
    bool haveKeys = false;
};

static LWECiphertext make_ct(const uint64_t* a, uint64_t b, uint32_t n, uint64_t mod, uint64_t p = 4) {
    NativeVector v(n, NativeInteger(mod));
    for (uint32_t i = 0; i < n; ++i)
        v[i] = NativeInteger(a[i]);
    auto ct = std::make_shared<LWECiphertextImpl>(std::move(v), NativeInteger(b));
    ct->SetptModulus(p);
    return ct;
}

static void put_ct(ConstLWECiphertext& ct, uint64_t* a, uint64_t* b) {
    uint32_t n = ct->GetLength();
    for (uint32_t i = 0; i < n; ++i)
        a[i] = ct->GetA(i).ConvertToInt();
    *b = ct->GetB().ConvertToInt();
}

extern "C" {

void* ref_ctx_create(int paramset, int method) {
    try {
        auto* h = new RefCtx();
        h->cc.GenerateBinFHEContext(static_cast<BINFHE_PARAMSET>(paramset), static_cast<BINFHE_METHOD>(method));
        return h;
    }
    catch (...) {
        return nullptr;
    }
}

// custom parameter set: BinFHEContext::GenerateBinFHEContext(const BinFHEContextParams&, method), binfhecontext.cpp:176-187
void* ref_ctx_create_custom(uint32_t bits, uint32_t cyclOrder, uint32_t n, uint32_t q, uint32_t modKS, uint32_t baseKS,
                            uint32_t baseG, uint32_t baseR, uint32_t numAutoKeys, int keyDist, int method) {
    try {
        auto* h = new RefCtx();
        BinFHEContextParams p{bits, cyclOrder, n, q, modKS, 3.19, baseKS, baseG, baseR, numAutoKeys,
                              static_cast<SecretKeyDist>(keyDist)};
        h->cc.GenerateBinFHEContext(p, static_cast<BINFHE_METHOD>(method));
        return h;
    }
    catch (...) {
        return nullptr;
    }
}

void ref_ctx_destroy(void* hv) {
    delete static_cast<RefCtx*>(hv);
}

// out[0..15]: n N q Q qKS baseKS baseG baseR digitsG numAutoKeys keyDist psi(2N-th root) digitsKS method
int ref_ctx_params(void* hv, uint64_t* out) {
    auto* h   = static_cast<RefCtx*>(hv);
    auto lwe  = h->cc.GetParams()->GetLWEParams();
    auto rgsw = h->cc.GetParams()->GetRingGSWParams();
    out[0]    = lwe->Getn();
    out[1]    = lwe->GetN();
    out[2]    = lwe->Getq().ConvertToInt();
    out[3]    = lwe->GetQ().ConvertToInt();
    out[4]    = lwe->GetqKS().ConvertToInt();
    out[5]    = lwe->GetBaseKS();
    out[6]    = rgsw->GetBaseG();
    out[7]    = rgsw->GetBaseR();
    out[8]    = rgsw->GetDigitsG();
    out[9]    = rgsw->GetNumAutoKeys();
    out[10]   = static_cast<uint64_t>(lwe->GetKeyDist());
    out[11]   = rgsw->GetPolyParams()->GetRootOfUnity().ConvertToInt();
    out[12]   = static_cast<uint64_t>(
        std::ceil(log(lwe->GetqKS().ConvertToDouble()) / log(static_cast<double>(lwe->GetBaseKS()))));
    out[13] = static_cast<uint64_t>(rgsw->GetMethod());
    return 0;
}

int ref_keygen(void* hv) {
    auto* h = static_cast<RefCtx*>(hv);
    try {
        h->sk = h->cc.KeyGen();
        h->cc.BTKeyGen(h->sk);
        h->haveKeys = true;
        return 0;
    }
    catch (...) {
        return -1;
    }
}

// secret key as signed values in (-mod/2, mod/2]
int ref_get_sk(void* hv, int64_t* out) {
    auto* h        = static_cast<RefCtx*>(hv);
    const auto& sv = h->sk->GetElement();
    uint64_t mod   = sv.GetModulus().ConvertToInt();
    for (uint32_t i = 0; i < sv.GetLength(); ++i) {
        uint64_t v = sv[i].ConvertToInt();
        out[i]     = v > mod / 2 ? static_cast<int64_t>(v) - static_cast<int64_t>(mod) : static_cast<int64_t>(v);
    }
    return 0;
}

// BSK dims: d0 d1 d2 of RingGSWACCKeyImpl, then rows (of entry [0][0][0]) and N
int ref_bsk_dims(void* hv, uint64_t* out) {
    auto* h        = static_cast<RefCtx*>(hv);
    const auto& ek = h->cc.GetRefreshKey()->GetElements();
    out[0]         = ek.size();
    out[1]         = ek[0].size();
    out[2]         = ek[0][0].size();
    out[3]         = 0;
    out[4]         = 0;
    for (auto& x : ek)
        for (auto& y : x)
            for (auto& z : y)
                if (z) {
                    out[3] = z->GetElements().size();
                    out[4] = z->GetElements()[0][0].GetLength();
                    return 0;
                }
    return 0;
}

// Dump entry [i0][i1][i2] (rows x 2 x N, EVALUATION format, reference order) ; returns rows, 0 if null
int ref_bsk_entry(void* hv, uint32_t i0, uint32_t i1, uint32_t i2, uint64_t* out) {
    auto* h        = static_cast<RefCtx*>(hv);
    const auto& ek = h->cc.GetRefreshKey()->GetElements();
    const auto& e  = ek[i0][i1][i2];
    if (!e)
        return 0;
    const auto& el = e->GetElements();
    size_t pos     = 0;
    for (auto& row : el)
        for (auto& poly : row) {
            uint32_t N = poly.GetLength();
            for (uint32_t k = 0; k < N; ++k)
                out[pos++] = poly[k].ConvertToInt();
        }
    return static_cast<int>(el.size());
}

// KSK: A [N][baseKS][digits][n], B [N][baseKS][digits]
int ref_ksk_dump(void* hv, uint64_t* A, uint64_t* B) {
    auto* h        = static_cast<RefCtx*>(hv);
    const auto& kA = h->cc.GetSwitchKey()->GetElementsA();
    const auto& kB = h->cc.GetSwitchKey()->GetElementsB();
    size_t pa = 0, pb = 0;
    for (size_t i = 0; i < kA.size(); ++i)
        for (size_t j = 0; j < kA[i].size(); ++j)
            for (size_t k = 0; k < kA[i][j].size(); ++k) {
                const auto& v = kA[i][j][k];
                for (uint32_t t = 0; t < v.GetLength(); ++t)
                    A[pa++] = v[t].ConvertToInt();
                B[pb++] = kB[i][j][k].ConvertToInt();
            }
    return 0;
}

int ref_encrypt(void* hv, int64_t m, uint64_t p, uint64_t mod, uint64_t* a, uint64_t* b) {
    auto* h = static_cast<RefCtx*>(hv);
    auto ct = h->cc.Encrypt(h->sk, m, SMALL_DIM, p, NativeInteger(mod));
    put_ct(ct, a, b);
    return 0;
}

int64_t ref_decrypt(void* hv, const uint64_t* a, uint64_t b, uint32_t len, uint64_t mod, uint64_t p) {
    auto* h = static_cast<RefCtx*>(hv);
    auto ct = make_ct(a, b, len, mod, p);
    LWEPlaintext r;
    h->cc.Decrypt(h->sk, ct, &r, p);
    return r;
}

// count independent 2-input gates; inputs [count][n] / [count]; out_len = n (or N if extended)
int ref_evalbingate(void* hv, int gate, const uint64_t* a1, const uint64_t* b1, const uint64_t* a2,
                    const uint64_t* b2, uint64_t* oa, uint64_t* ob, uint64_t count, int extended, int threads) {
    auto* h      = static_cast<RefCtx*>(hv);
    auto lwe     = h->cc.GetParams()->GetLWEParams();
    uint32_t n   = lwe->Getn();
    uint64_t q   = lwe->Getq().ConvertToInt();
    uint32_t on  = extended ? lwe->GetN() : n;
    int failures = 0;
#pragma omp parallel for num_threads(threads > 0 ? threads : 1) schedule(dynamic)
    for (uint64_t i = 0; i < count; ++i) {
        try {
            auto c1 = make_ct(a1 + i * n, b1[i], n, q);
            auto c2 = make_ct(a2 + i * n, b2[i], n, q);
            auto r  = h->cc.EvalBinGate(static_cast<BINGATE>(gate), c1, c2, extended != 0);
            put_ct(r, oa + i * on, ob + i);
        }
        catch (...) {
#pragma omp atomic
            failures++;
        }
    }
    return failures;
}

// multi-input gates (AND3/OR3/MAJORITY/AND4/OR4/CMUX): inputs [count][k][n]
int ref_evalbingate_multi(void* hv, int gate, uint32_t k, const uint64_t* a, const uint64_t* b, uint64_t* oa,
                          uint64_t* ob, uint64_t count, uint64_t p, int threads) {
    auto* h      = static_cast<RefCtx*>(hv);
    auto lwe     = h->cc.GetParams()->GetLWEParams();
    uint32_t n   = lwe->Getn();
    uint64_t q   = lwe->Getq().ConvertToInt();
    int failures = 0;
#pragma omp parallel for num_threads(threads > 0 ? threads : 1) schedule(dynamic)
    for (uint64_t i = 0; i < count; ++i) {
        try {
            std::vector<LWECiphertext> v;
            for (uint32_t j = 0; j < k; ++j)
                v.push_back(make_ct(a + (i * k + j) * n, b[i * k + j], n, q, p));
            auto r = h->cc.EvalBinGate(static_cast<BINGATE>(gate), v);
            put_ct(r, oa + i * n, ob + i);
        }
        catch (...) {
#pragma omp atomic
            failures++;
        }
    }
    return failures;
}

int ref_bootstrap(void* hv, const uint64_t* a, const uint64_t* b, uint64_t* oa, uint64_t* ob, uint64_t count,
                  uint64_t p, int extended, int threads) {
    auto* h     = static_cast<RefCtx*>(hv);
    auto lwe    = h->cc.GetParams()->GetLWEParams();
    uint32_t n  = lwe->Getn();
    uint64_t q  = lwe->Getq().ConvertToInt();
    uint32_t on = extended ? lwe->GetN() : n;
#pragma omp parallel for num_threads(threads > 0 ? threads : 1) schedule(dynamic)
    for (uint64_t i = 0; i < count; ++i) {
        auto c = make_ct(a + i * n, b[i], n, q, p);
        auto r = h->cc.Bootstrap(c, extended != 0);
        put_ct(r, oa + i * on, ob + i);
    }
    return 0;
}

// EvalFunc with an explicit LUT of length q (ciphertext modulus). Returns -1 if the reference throws.
int ref_evalfunc(void* hv, const uint64_t* a, const uint64_t* b, const uint64_t* lut, uint64_t lutlen, uint64_t* oa,
                 uint64_t* ob, uint64_t count, uint64_t ctmod, int threads) {
    auto* h    = static_cast<RefCtx*>(hv);
    auto lwe   = h->cc.GetParams()->GetLWEParams();
    uint32_t n = lwe->Getn();
    std::vector<NativeInteger> L(lutlen);
    for (uint64_t i = 0; i < lutlen; ++i)
        L[i] = NativeInteger(lut[i]);
    int failures = 0;
#pragma omp parallel for num_threads(threads > 0 ? threads : 1) schedule(dynamic)
    for (uint64_t i = 0; i < count; ++i) {
        try {
            auto c = make_ct(a + i * n, b[i], n, ctmod);
            auto r = h->cc.EvalFunc(c, L);
            put_ct(r, oa + i * n, ob + i);
        }
        catch (...) {
#pragma omp atomic
            failures++;
        }
    }
    return failures ? -1 : 0;
}

uint64_t ref_get_beta(void* hv) {
    return static_cast<RefCtx*>(hv)->cc.GetBeta().ConvertToInt();
}

// Blind rotation only (RingGSWAccumulator::EvalAcc through the scheme bound in the context):
// acc [2][N] in EVALUATION format in/out; a [n] mod amod.
int ref_evalacc(void* hv, const uint64_t* a, uint64_t amod, uint64_t* acc) {
    auto* h     = static_cast<RefCtx*>(hv);
    auto params = h->cc.GetParams();
    auto rgsw   = params->GetRingGSWParams();
    auto pp     = rgsw->GetPolyParams();
    uint32_t N  = rgsw->GetN();
    uint32_t n  = params->GetLWEParams()->Getn();
    std::vector<NativePoly> res(2);
    for (int c = 0; c < 2; ++c) {
        res[c] = NativePoly(pp, Format::EVALUATION, true);
        for (uint32_t k = 0; k < N; ++k)
            res[c][k] = NativeInteger(acc[c * N + k]);
    }
    auto accct = std::make_shared<RLWECiphertextImpl>(std::move(res));
    NativeVector av(n, NativeInteger(amod));
    for (uint32_t i = 0; i < n; ++i)
        av[i] = NativeInteger(a[i]);
    std::shared_ptr<RingGSWAccumulator> sch;
    switch (rgsw->GetMethod()) {
        case AP:
            sch = std::make_shared<RingGSWAccumulatorDM>();
            break;
        case GINX:
            sch = std::make_shared<RingGSWAccumulatorCGGI>();
            break;
        default:
            sch = std::make_shared<RingGSWAccumulatorLMKCDEY>();
    }
    sch->EvalAcc(rgsw, h->cc.GetRefreshKey(), accct, av);
    for (int c = 0; c < 2; ++c)
        for (uint32_t k = 0; k < N; ++k)
            acc[c * N + k] = accct->GetElements()[c][k].ConvertToInt();
    return 0;
}

// SignedDigitDecompose of a 2-component ciphertext (COEFFICIENT values) -> out [digitsG2][N]
int ref_decompose(void* hv, const uint64_t* in, uint64_t* out) {
    auto* h     = static_cast<RefCtx*>(hv);
    auto rgsw   = h->cc.GetParams()->GetRingGSWParams();
    auto pp     = rgsw->GetPolyParams();
    uint32_t N  = rgsw->GetN();
    uint32_t d2 = (rgsw->GetDigitsG() - 1) << 1;
    std::vector<NativePoly> ct(2, NativePoly(pp, Format::COEFFICIENT, true));
    for (int c = 0; c < 2; ++c)
        for (uint32_t k = 0; k < N; ++k)
            ct[c][k] = NativeInteger(in[c * N + k]);
    std::vector<NativePoly> dct(d2, NativePoly(pp, Format::COEFFICIENT, true));
    RingGSWAccumulatorCGGI().SignedDigitDecompose(rgsw, ct, dct);
    for (uint32_t d = 0; d < d2; ++d)
        for (uint32_t k = 0; k < N; ++k)
            out[d * N + k] = dct[d][k].ConvertToInt();
    return 0;
}

// LWE ModSwitch / KeySwitch on flat data (lwe-pke.cpp:254-261, :348-372)
int ref_modswitch(void* hv, uint64_t qto, uint64_t qfrom, const uint64_t* a, const uint64_t* b, uint64_t* oa,
                  uint64_t* ob, uint32_t len, uint64_t count) {
    auto* h = static_cast<RefCtx*>(hv);
    for (uint64_t i = 0; i < count; ++i) {
        auto c = make_ct(a + i * len, b[i], len, qfrom);
        auto r = h->cc.GetLWEScheme()->ModSwitch(NativeInteger(qto), c);
        put_ct(r, oa + i * len, ob + i);
    }
    return 0;
}

int ref_keyswitch(void* hv, const uint64_t* a, const uint64_t* b, uint64_t* oa, uint64_t* ob, uint64_t count) {
    auto* h    = static_cast<RefCtx*>(hv);
    auto lwe   = h->cc.GetParams()->GetLWEParams();
    uint32_t N = lwe->GetN(), n = lwe->Getn();
    for (uint64_t i = 0; i < count; ++i) {
        auto c = make_ct(a + i * N, b[i], N, lwe->GetqKS().ConvertToInt());
        auto r = h->cc.GetLWEScheme()->KeySwitch(lwe, h->cc.GetSwitchKey(), c);
        put_ct(r, oa + i * n, ob + i);
    }
    return 0;
}

// ---- NTT (NativePoly::SwitchFormat path) -------------------------------------------------------
uint64_t ref_root_of_unity(uint64_t m, uint64_t Q) {
    return RootOfUnity<NativeInteger>(static_cast<usint>(m), NativeInteger(Q)).ConvertToInt();
}

uint64_t ref_last_prime(uint32_t bits, uint64_t m) {
    return LastPrime<NativeInteger>(bits, m).ConvertToInt();
}

uint64_t ref_previous_prime(uint64_t q, uint64_t m) {
    return PreviousPrime<NativeInteger>(NativeInteger(q), m).ConvertToInt();
}

// data [count][N], in place. inverse=0: ForwardTransformToBitReverseInPlace, 1: InverseTransformFromBitReverseInPlace
int ref_ntt(uint64_t Q, uint64_t psi, uint32_t N, uint64_t* data, uint64_t count, int inverse, int threads) {
    NativeInteger mod(Q), root(psi);
    // fill the process-global tables once, single-threaded (transformnat-impl.h:777-831)
    {
        NativeVector v(N, mod);
        ChineseRemainderTransformFTT<NativeVector>().ForwardTransformToBitReverseInPlace(root, 2 * N, &v);
        ChineseRemainderTransformFTT<NativeVector>().InverseTransformFromBitReverseInPlace(root, 2 * N, &v);
    }
#pragma omp parallel for num_threads(threads > 0 ? threads : 1)
    for (uint64_t i = 0; i < count; ++i) {
        NativeVector v(N, mod);
        uint64_t* p = data + i * N;
        for (uint32_t k = 0; k < N; ++k)
            v[k] = NativeInteger(p[k]);
        if (inverse)
            ChineseRemainderTransformFTT<NativeVector>().InverseTransformFromBitReverseInPlace(root, 2 * N, &v);
        else
            ChineseRemainderTransformFTT<NativeVector>().ForwardTransformToBitReverseInPlace(root, 2 * N, &v);
        for (uint32_t k = 0; k < N; ++k)
            p[k] = v[k].ConvertToInt();
    }
    return 0;
}

// Moduli chain of ILDCRTParams(order=2N, depth=L, bits) (ildcrtparams.h:100-118): q[L], psi[L]
int ref_dcrt_chain(uint32_t N, uint32_t L, uint32_t bits, uint64_t* q, uint64_t* psi) {
    ILDCRTParams<BigInteger> p(2 * N, L, bits);
    for (uint32_t l = 0; l < L; ++l) {
        q[l]   = p.GetParams()[l]->GetModulus().ConvertToInt();
        psi[l] = p.GetParams()[l]->GetRootOfUnity().ConvertToInt();
    }
    return 0;
}

// DCRTPoly::SwitchFormat on flat data [count][L][N] (dcrtpoly-impl.h:1953-1959), in place.
// fmt_in: 0 = COEFFICIENT (forward NTT), 1 = EVALUATION (inverse NTT)
int ref_dcrt_switchformat(uint32_t N, uint32_t L, uint32_t bits, uint64_t* data, uint64_t count, int fmt_in) {
    auto params = std::make_shared<ILDCRTParams<BigInteger>>(2 * N, L, bits);
    for (uint64_t i = 0; i < count; ++i) {
        DCRTPoly poly(params, fmt_in ? Format::EVALUATION : Format::COEFFICIENT, true);
        for (uint32_t l = 0; l < L; ++l) {
            NativePoly limb = poly.GetElementAtIndex(l);
            uint64_t* p     = data + (i * L + l) * N;
            for (uint32_t k = 0; k < N; ++k)
                limb[k] = NativeInteger(p[k]);
            poly.SetElementAtIndex(l, std::move(limb));
        }
        poly.SwitchFormat();
        for (uint32_t l = 0; l < L; ++l) {
            const NativePoly& limb = poly.GetElementAtIndex(l);
            uint64_t* p            = data + (i * L + l) * N;
            for (uint32_t k = 0; k < N; ++k)
                p[k] = limb[k].ConvertToInt();
        }
    }
    return 0;
}


// ---- large-precision contexts (binfhecontext.cpp:53-104) and EvalFloor / EvalSign / EvalDecomp -----------------------
void* ref_ctx_create_largeprec(int paramset, int arbFunc, uint32_t logQ, uint32_t N, int timeOpt) {
    try {
        auto* h = new RefCtx();
        h->cc.GenerateBinFHEContext(static_cast<BINFHE_PARAMSET>(paramset), arbFunc != 0, logQ, N, GINX, timeOpt != 0);
        return h;
    }
    catch (...) {
        return nullptr;
    }
}
// gadget bases of the key map after BTKeyGen (binfhecontext.cpp:292-306); returns the count
int ref_btkey_bases(void* hv, uint32_t* out, int max) {
    auto* h = static_cast<RefCtx*>(hv);
    int k   = 0;
    auto map = h->cc.GetBTKeyMap();   // a copy held for the loop (the accessor returns a temporary)
    for (auto& kv : *map)
        if (k < max)
            out[k++] = kv.first;
    return k;
}
static RingGSWBTKey key_of_base(RefCtx* h, uint32_t baseG) {
    return h->cc.GetBTKeyMap()->at(baseG);   // by value: shared_ptr members keep the key objects alive
}
int ref_bsk_entry_base(void* hv, uint32_t baseG, uint32_t i0, uint32_t i1, uint32_t i2, uint64_t* out) {
    auto* h        = static_cast<RefCtx*>(hv);
    const RingGSWBTKey key = key_of_base(h, baseG);
    const auto& ek = key.BSkey->GetElements();
    const auto& e  = ek[i0][i1][i2];
    if (!e)
        return 0;
    const auto& el = e->GetElements();
    size_t pos     = 0;
    for (auto& row : el)
        for (auto& poly : row) {
            uint32_t N = poly.GetLength();
            for (uint32_t k = 0; k < N; ++k)
                out[pos++] = poly[k].ConvertToInt();
        }
    return static_cast<int>(el.size());
}
int ref_ksk_dump_base(void* hv, uint32_t baseG, uint64_t* A, uint64_t* B) {
    auto* h        = static_cast<RefCtx*>(hv);
    const RingGSWBTKey key = key_of_base(h, baseG);
    const auto& ks = key.KSkey;
    const auto& kA = ks->GetElementsA();
    const auto& kB = ks->GetElementsB();
    size_t pa = 0, pb = 0;
    for (size_t i = 0; i < kA.size(); ++i)
        for (size_t j = 0; j < kA[i].size(); ++j)
            for (size_t k = 0; k < kA[i][j].size(); ++k) {
                const auto& v = kA[i][j][k];
                for (uint32_t t = 0; t < v.GetLength(); ++t)
                    A[pa++] = v[t].ConvertToInt();
                B[pb++] = kB[i][j][k].ConvertToInt();
            }
    return 0;
}
uint32_t ref_current_baseG(void* hv) {
    return static_cast<RefCtx*>(hv)->cc.GetParams()->GetRingGSWParams()->GetBaseG();
}
int ref_evalfloor(void* hv, const uint64_t* a, const uint64_t* b, uint64_t* oa, uint64_t* ob, uint64_t count, uint64_t mod,
                  uint32_t roundbits, int threads) {
    auto* h    = static_cast<RefCtx*>(hv);
    uint32_t n = h->cc.GetParams()->GetLWEParams()->Getn();
    int rc     = 0;
#pragma omp parallel for num_threads(threads > 0 ? threads : 1) schedule(dynamic)
    for (uint64_t i = 0; i < count; ++i) {
        try {
            auto r = h->cc.EvalFloor(make_ct(a + i * n, b[i], n, mod), roundbits);
            put_ct(r, oa + i * n, ob + i);
        }
        catch (...) {
            rc = -1;
        }
    }
    return rc;
}
// EvalSign mutates the shared RingGSW parameters (Change_BaseG): one ciphertext at a time
int ref_evalsign(void* hv, const uint64_t* a, const uint64_t* b, uint64_t* oa, uint64_t* ob, uint64_t count, uint64_t mod,
                 int schemeSwitch) {
    auto* h    = static_cast<RefCtx*>(hv);
    uint32_t n = h->cc.GetParams()->GetLWEParams()->Getn();
    try {
        for (uint64_t i = 0; i < count; ++i) {
            auto r = h->cc.EvalSign(make_ct(a + i * n, b[i], n, mod), schemeSwitch != 0);
            put_ct(r, oa + i * n, ob + i);
        }
    }
    catch (...) {
        return -1;
    }
    return 0;
}
// out_a [k][count][n], out_b [k][count], out_mod [k]; returns the number of ciphertexts per input (or -1)
int ref_evaldecomp(void* hv, const uint64_t* a, const uint64_t* b, uint64_t* out_a, uint64_t* out_b, uint64_t* out_mod,
                   uint32_t max_out, uint64_t count, uint64_t mod) {
    auto* h    = static_cast<RefCtx*>(hv);
    uint32_t n = h->cc.GetParams()->GetLWEParams()->Getn();
    int k      = -1;
    try {
        for (uint64_t i = 0; i < count; ++i) {
            auto v = h->cc.EvalDecomp(make_ct(a + i * n, b[i], n, mod));
            if (v.size() > max_out)
                return -1;
            k = static_cast<int>(v.size());
            for (size_t d = 0; d < v.size(); ++d) {
                put_ct(v[d], out_a + (d * count + i) * n, out_b + d * count + i);
                out_mod[d] = v[d]->GetModulus().ConvertToInt();
            }
        }
    }
    catch (...) {
        return -1;
    }
    return k;
}


// ---- RNS basis conversion (CKKS/BGV hybrid key switching): the reference's own DCRTPoly methods --------------------------
// ApproxSwitchCRTBasis / ApproxModUp / ApproxModDown (dcrtpoly-impl.h:910-1030) with the tables CryptoParametersRNS::
// PrecomputeCRTTables (src/pke/lib/schemerns/rns-cryptoparameters.cpp) would hand them, computed here from their textbook
// definitions with BigInteger: QHat_i = Q / q_i, QHatInvModq_i = QHat_i^-1 mod q_i, QHatModp[i][j] = QHat_i mod p_j,
// modpBarrettMu_j = floor(2^128 / p_j), PInvModq_i = P^-1 mod q_i.
namespace {
struct RnsTables {
    std::vector<NativeInteger> hatInv, hatInvPrecon;
    std::vector<std::vector<NativeInteger>> hatMod;
    std::vector<DoubleNativeInt> mu;
};
RnsTables rns_tables(const std::vector<NativeInteger>& from, const std::vector<NativeInteger>& to) {
    RnsTables T;
    BigInteger Q(1);
    for (auto& m : from) Q *= BigInteger(m.ConvertToInt<uint64_t>());
    for (size_t i = 0; i < from.size(); ++i) {
        BigInteger qi(from[i].ConvertToInt<uint64_t>());
        BigInteger hat = Q / qi;
        NativeInteger inv = NativeInteger(hat.Mod(qi).ConvertToInt<uint64_t>()).ModInverse(from[i]);
        T.hatInv.push_back(inv);
        T.hatInvPrecon.push_back(inv.PrepModMulConst(from[i]));
        std::vector<NativeInteger> row;
        for (auto& pj : to) row.push_back(NativeInteger(hat.Mod(BigInteger(pj.ConvertToInt<uint64_t>())).ConvertToInt<uint64_t>()));
        T.hatMod.push_back(row);
    }
    for (auto& pj : to) {
        const DoubleNativeInt all = ~DoubleNativeInt(0);   // floor(2^128 / p) = floor((2^128 - 1) / p) for p not a power of two
        T.mu.push_back(all / DoubleNativeInt(pj.ConvertToInt<uint64_t>()));
    }
    return T;
}
}  // namespace

// x [count][limbs_in][N] -> out [count][limbs_out][N].  what 0: ApproxSwitchCRTBasis Q -> P on COEFFICIENT data (sizeQ in,
// sizeP out); 1: ApproxModUp, EVALUATION (sizeQ in, sizeQ + sizeP out); 2: ApproxModDown with t = 0, EVALUATION (sizeQ + sizeP
// in, sizeQ out)
static int ref_rns_op_t(uint32_t N, uint32_t sizeQ, const uint64_t* q, uint32_t sizeP, const uint64_t* p, const uint64_t* x, uint64_t* out,
                        uint64_t count, int what, uint64_t ptm);
int ref_rns_op(uint32_t N, uint32_t sizeQ, const uint64_t* q, uint32_t sizeP, const uint64_t* p, const uint64_t* x, uint64_t* out,
               uint64_t count, int what) {
    return ref_rns_op_t(N, sizeQ, q, sizeP, p, x, out, count, what, 0);
}
// ApproxModDown with the BGV tables of bgvrns-cryptoparameters.cpp:78-95 (tInvModp = t^-1 mod p_j, tModqPrecon) and t = ptm
int ref_rns_moddown_t(uint32_t N, uint32_t sizeQ, const uint64_t* q, uint32_t sizeP, const uint64_t* p, const uint64_t* x, uint64_t* out,
                      uint64_t count, uint64_t ptm) {
    return ref_rns_op_t(N, sizeQ, q, sizeP, p, x, out, count, 2, ptm);
}
static int ref_rns_op_t(uint32_t N, uint32_t sizeQ, const uint64_t* q, uint32_t sizeP, const uint64_t* p, const uint64_t* x, uint64_t* out,
                        uint64_t count, int what, uint64_t ptm) {
    try {
        std::vector<NativeInteger> mq, mp, rq, rp, mqp, rqp;
        for (uint32_t i = 0; i < sizeQ; ++i) { mq.emplace_back(q[i]); rq.push_back(RootOfUnity<NativeInteger>(2 * N, mq.back())); }
        for (uint32_t j = 0; j < sizeP; ++j) { mp.emplace_back(p[j]); rp.push_back(RootOfUnity<NativeInteger>(2 * N, mp.back())); }
        mqp = mq; mqp.insert(mqp.end(), mp.begin(), mp.end());
        rqp = rq; rqp.insert(rqp.end(), rp.begin(), rp.end());
        auto paramsQ  = std::make_shared<ILDCRTParams<BigInteger>>(2 * N, mq, rq);
        auto paramsP  = std::make_shared<ILDCRTParams<BigInteger>>(2 * N, mp, rp);
        auto paramsQP = std::make_shared<ILDCRTParams<BigInteger>>(2 * N, mqp, rqp);
        const RnsTables up = rns_tables(mq, mp), down = rns_tables(mp, mq);
        std::vector<NativeInteger> pInv, pInvPrecon;
        BigInteger P(1);
        for (auto& m : mp) P *= BigInteger(m.ConvertToInt<uint64_t>());
        for (auto& qi : mq) {
            NativeInteger inv = NativeInteger(P.Mod(BigInteger(qi.ConvertToInt<uint64_t>())).ConvertToInt<uint64_t>()).ModInverse(qi);
            pInv.push_back(inv);
            pInvPrecon.push_back(inv.PrepModMulConst(qi));
        }
        const uint32_t lin = what == 2 ? sizeQ + sizeP : sizeQ, lout = what == 0 ? sizeP : (what == 1 ? sizeQ + sizeP : sizeQ);
        for (uint64_t c = 0; c < count; ++c) {
            DCRTPoly in(what == 2 ? paramsQP : paramsQ, what == 0 ? Format::COEFFICIENT : Format::EVALUATION, true);
            for (uint32_t l = 0; l < lin; ++l) {
                NativePoly& limb = in.GetAllElements()[l];
                for (uint32_t k = 0; k < N; ++k) limb[k] = NativeInteger(x[(c * lin + l) * N + k]);
            }
            DCRTPoly res;
            if (what == 0) res = in.ApproxSwitchCRTBasis(paramsQ, paramsP, up.hatInv, up.hatInvPrecon, up.hatMod, up.mu);
            else if (what == 1) { in.ApproxModUp(paramsQ, paramsP, paramsQP, up.hatInv, up.hatInvPrecon, up.hatMod, up.mu); res = in; }
            else if (ptm == 0) res = in.ApproxModDown(paramsQ, paramsP, pInv, pInvPrecon, down.hatInv, down.hatInvPrecon, down.hatMod, down.mu, {}, {},
                                                      NativeInteger(0), {});
            else {
                const NativeInteger t(ptm);
                std::vector<NativeInteger> tInvModp, tInvModpPrecon, tModqPrecon;
                for (auto& pj : mp) { tInvModp.push_back(t.ModInverse(pj)); tInvModpPrecon.push_back(tInvModp.back().PrepModMulConst(pj)); }
                for (auto& qi : mq) tModqPrecon.push_back(t.Mod(qi).PrepModMulConst(qi));
                res = in.ApproxModDown(paramsQ, paramsP, pInv, pInvPrecon, down.hatInv, down.hatInvPrecon, down.hatMod, down.mu, tInvModp,
                                       tInvModpPrecon, t, tModqPrecon);
            }
            if (res.GetNumOfElements() != lout) return -2;
            for (uint32_t l = 0; l < lout; ++l) {
                const NativePoly& limb = res.GetAllElements()[l];
                for (uint32_t k = 0; k < N; ++k) out[(c * lout + l) * N + k] = limb[k].ConvertToInt<uint64_t>();
            }
        }
        return 0;
    }
    catch (...) {
        return -1;
    }
}

// The loop body of KeySwitchHYBRID::EvalFastKeySwitchCoreExt (keyswitch-hybrid.cpp:457-477) on the reference's own DCRTPoly /
// NativePoly objects: cTilda_c[i] += cj[i] * (b|a)_j[idx(i)].  (The function itself needs an EvalKey with crypto parameters; its
// arithmetic is exactly these element operations.)  Layouts as orc_keyswitch_inner.
int ref_keyswitch_inner(uint32_t N, uint32_t sizeQ, const uint64_t* q, uint32_t sizeP, const uint64_t* p, uint32_t sizeQl, uint32_t nd,
                        const uint64_t* digits, const uint64_t* evk_b, const uint64_t* evk_a, uint64_t* out, uint64_t count) {
    try {
        std::vector<NativeInteger> mqlp, rqlp, mqp, rqp;
        for (uint32_t i = 0; i < sizeQl; ++i) { mqlp.emplace_back(q[i]); rqlp.push_back(RootOfUnity<NativeInteger>(2 * N, mqlp.back())); }
        for (uint32_t j = 0; j < sizeP; ++j) { mqlp.emplace_back(p[j]); rqlp.push_back(RootOfUnity<NativeInteger>(2 * N, mqlp.back())); }
        for (uint32_t i = 0; i < sizeQ; ++i) { mqp.emplace_back(q[i]); rqp.push_back(RootOfUnity<NativeInteger>(2 * N, mqp.back())); }
        for (uint32_t j = 0; j < sizeP; ++j) { mqp.emplace_back(p[j]); rqp.push_back(RootOfUnity<NativeInteger>(2 * N, mqp.back())); }
        auto paramsQlP = std::make_shared<ILDCRTParams<BigInteger>>(2 * N, mqlp, rqlp);
        auto paramsQP  = std::make_shared<ILDCRTParams<BigInteger>>(2 * N, mqp, rqp);
        const uint32_t L = sizeQl + sizeP, LK = sizeQ + sizeP;
        auto fill = [&](DCRTPoly& d, const uint64_t* src, uint32_t limbs) {
            for (uint32_t l = 0; l < limbs; ++l) {
                NativePoly& limb = d.GetAllElements()[l];
                for (uint32_t k = 0; k < N; ++k) limb[k] = NativeInteger(src[(size_t)l * N + k]);
            }
        };
        std::vector<DCRTPoly> bv, av;
        for (uint32_t j = 0; j < nd; ++j) {
            bv.emplace_back(paramsQP, Format::EVALUATION, true); fill(bv.back(), evk_b + (size_t)j * LK * N, LK);
            av.emplace_back(paramsQP, Format::EVALUATION, true); fill(av.back(), evk_a + (size_t)j * LK * N, LK);
        }
        for (uint64_t c = 0; c < count; ++c) {
            DCRTPoly cTilda0(paramsQlP, Format::EVALUATION, true), cTilda1(paramsQlP, Format::EVALUATION, true);
            for (uint32_t j = 0; j < nd; ++j) {
                DCRTPoly cj(paramsQlP, Format::EVALUATION, true);
                fill(cj, digits + ((size_t)c * nd + j) * L * N, L);
                for (uint32_t i = 0; i < sizeQl; ++i) {
                    cTilda0.SetElementAtIndex(i, cTilda0.GetElementAtIndex(i) + cj.GetElementAtIndex(i) * bv[j].GetElementAtIndex(i));
                    cTilda1.SetElementAtIndex(i, cTilda1.GetElementAtIndex(i) + cj.GetElementAtIndex(i) * av[j].GetElementAtIndex(i));
                }
                for (uint32_t i = sizeQl, idx = sizeQ; i < L; ++i, ++idx) {
                    cTilda0.SetElementAtIndex(i, cTilda0.GetElementAtIndex(i) + cj.GetElementAtIndex(i) * bv[j].GetElementAtIndex(idx));
                    cTilda1.SetElementAtIndex(i, cTilda1.GetElementAtIndex(i) + cj.GetElementAtIndex(i) * av[j].GetElementAtIndex(idx));
                }
            }
            for (uint32_t l = 0; l < L; ++l)
                for (uint32_t k = 0; k < N; ++k) {
                    out[(((size_t)c * 2 + 0) * L + l) * N + k] = cTilda0.GetAllElements()[l][k].ConvertToInt<uint64_t>();
                    out[(((size_t)c * 2 + 1) * L + l) * N + k] = cTilda1.GetAllElements()[l][k].ConvertToInt<uint64_t>();
                }
        }
        return 0;
    }
    catch (...) {
        return -1;
    }
}

// Hybrid key switching of one polynomial, composed from the reference's own DCRTPoly primitives in the order of
// KeySwitchHYBRID::EvalKeySwitchPrecomputeCore (keyswitch-hybrid.cpp:328-412) and EvalFastKeySwitchCore (:414-436): Clone /
// SetFormat(COEFFICIENT) / ApproxSwitchCRTBasis / SetFormat(EVALUATION) per part, the element loops of :394-403 and :457-477,
// ApproxModDown twice.  (The member functions themselves need a CryptoContext of src/pke, which oracle/_ref does not build;
// the tables are those of CryptoParametersRNS::PrecomputeCRTTables for each part.)  Layouts as orc_hybrid_keyswitch.
int ref_hybrid_keyswitch(uint32_t N, uint32_t sizeQ, const uint64_t* q, uint32_t sizeP, const uint64_t* p, uint32_t alpha, uint32_t sizeQl,
                         const uint64_t* c, const uint64_t* evk_b, const uint64_t* evk_a, uint64_t* out, uint64_t count, uint64_t ptm) {
    try {
        auto mk = [&](const std::vector<NativeInteger>& m) {
            std::vector<NativeInteger> r;
            for (auto& x : m) r.push_back(RootOfUnity<NativeInteger>(2 * N, x));
            return std::make_shared<ILDCRTParams<BigInteger>>(2 * N, m, r);
        };
        std::vector<NativeInteger> mql, mp, mqlp, mqp;
        for (uint32_t i = 0; i < sizeQl; ++i) mql.emplace_back(q[i]);
        for (uint32_t j = 0; j < sizeP; ++j) mp.emplace_back(p[j]);
        mqlp = mql; mqlp.insert(mqlp.end(), mp.begin(), mp.end());
        for (uint32_t i = 0; i < sizeQ; ++i) mqp.emplace_back(q[i]);
        mqp.insert(mqp.end(), mp.begin(), mp.end());
        auto paramsQl = mk(mql), paramsP = mk(mp), paramsQlP = mk(mqlp), paramsQP = mk(mqp);
        const uint32_t L = sizeQl + sizeP, LK = sizeQ + sizeP;
        uint32_t nd = (sizeQl + alpha - 1) / alpha;
        if (nd > (sizeQ + alpha - 1) / alpha) nd = (sizeQ + alpha - 1) / alpha;
        auto fill = [&](DCRTPoly& d, const uint64_t* src, uint32_t limbs) {
            for (uint32_t l = 0; l < limbs; ++l) {
                NativePoly& limb = d.GetAllElements()[l];
                for (uint32_t k = 0; k < N; ++k) limb[k] = NativeInteger(src[(size_t)l * N + k]);
            }
        };
        std::vector<DCRTPoly> bv, av;
        for (uint32_t j = 0; j < nd; ++j) {
            bv.emplace_back(paramsQP, Format::EVALUATION, true); fill(bv.back(), evk_b + (size_t)j * LK * N, LK);
            av.emplace_back(paramsQP, Format::EVALUATION, true); fill(av.back(), evk_a + (size_t)j * LK * N, LK);
        }
        const RnsTables down = rns_tables(mp, mql);
        std::vector<NativeInteger> pInv, pInvPrecon, tInvModp, tInvModpPrecon, tModqPrecon;
        BigInteger Pbig(1);
        for (auto& m : mp) Pbig *= BigInteger(m.ConvertToInt<uint64_t>());
        for (auto& qi : mql) {
            NativeInteger inv = NativeInteger(Pbig.Mod(BigInteger(qi.ConvertToInt<uint64_t>())).ConvertToInt<uint64_t>()).ModInverse(qi);
            pInv.push_back(inv); pInvPrecon.push_back(inv.PrepModMulConst(qi));
        }
        const NativeInteger t(ptm);
        if (ptm) {
            for (auto& pj : mp) { tInvModp.push_back(t.ModInverse(pj)); tInvModpPrecon.push_back(tInvModp.back().PrepModMulConst(pj)); }
            for (auto& qi : mql) tModqPrecon.push_back(t.Mod(qi).PrepModMulConst(qi));
        }
        for (uint64_t ct = 0; ct < count; ++ct) {
            DCRTPoly cpoly(paramsQl, Format::EVALUATION, true);
            fill(cpoly, c + (size_t)ct * sizeQl * N, sizeQl);
            std::vector<DCRTPoly> partsCtExt(nd);
            for (uint32_t part = 0; part < nd; ++part) {
                const uint32_t start = alpha * part, sp = (part == nd - 1) ? sizeQl - start : alpha, end = start + sp;
                std::vector<NativeInteger> msrc, mdst;
                for (uint32_t i = 0; i < sp; ++i) msrc.push_back(mql[start + i]);
                for (uint32_t i = 0; i < L; ++i)
                    if (i < start || i >= end) mdst.push_back(mqlp[i]);
                auto paramsPart = mk(msrc), paramsCompl = mk(mdst);
                DCRTPoly partCt(paramsPart, Format::EVALUATION, true);
                for (uint32_t i = 0, idx = start; i < sp; ++i, ++idx) partCt.SetElementAtIndex(i, cpoly.GetElementAtIndex(idx));
                auto partCtClone = partCt.Clone();
                partCtClone.SetFormat(Format::COEFFICIENT);
                const RnsTables T = rns_tables(msrc, mdst);
                DCRTPoly compl_ = partCtClone.ApproxSwitchCRTBasis(paramsPart, paramsCompl, T.hatInv, T.hatInvPrecon, T.hatMod, T.mu);
                compl_.SetFormat(Format::EVALUATION);
                partsCtExt[part] = DCRTPoly(paramsQlP, Format::EVALUATION, true);
                for (uint32_t i = 0; i < start; ++i) partsCtExt[part].SetElementAtIndex(i, compl_.GetElementAtIndex(i));
                for (uint32_t i = start, idx = 0; i < end; ++i, ++idx) partsCtExt[part].SetElementAtIndex(i, partCt.GetElementAtIndex(idx));
                for (uint32_t i = end; i < L; ++i) partsCtExt[part].SetElementAtIndex(i, compl_.GetElementAtIndex(i - sp));
            }
            DCRTPoly cTilda0(paramsQlP, Format::EVALUATION, true), cTilda1(paramsQlP, Format::EVALUATION, true);
            for (uint32_t j = 0; j < nd; ++j) {
                const DCRTPoly& cj = partsCtExt[j];
                for (uint32_t i = 0; i < sizeQl; ++i) {
                    cTilda0.SetElementAtIndex(i, cTilda0.GetElementAtIndex(i) + cj.GetElementAtIndex(i) * bv[j].GetElementAtIndex(i));
                    cTilda1.SetElementAtIndex(i, cTilda1.GetElementAtIndex(i) + cj.GetElementAtIndex(i) * av[j].GetElementAtIndex(i));
                }
                for (uint32_t i = sizeQl, idx = sizeQ; i < L; ++i, ++idx) {
                    cTilda0.SetElementAtIndex(i, cTilda0.GetElementAtIndex(i) + cj.GetElementAtIndex(i) * bv[j].GetElementAtIndex(idx));
                    cTilda1.SetElementAtIndex(i, cTilda1.GetElementAtIndex(i) + cj.GetElementAtIndex(i) * av[j].GetElementAtIndex(idx));
                }
            }
            DCRTPoly* both[2] = {&cTilda0, &cTilda1};
            for (int w = 0; w < 2; ++w) {
                DCRTPoly r = both[w]->ApproxModDown(paramsQl, paramsP, pInv, pInvPrecon, down.hatInv, down.hatInvPrecon, down.hatMod, down.mu,
                                                    tInvModp, tInvModpPrecon, t, tModqPrecon);
                for (uint32_t l = 0; l < sizeQl; ++l)
                    for (uint32_t k = 0; k < N; ++k)
                        out[(((size_t)ct * 2 + w) * sizeQl + l) * N + k] = r.GetAllElements()[l][k].ConvertToInt<uint64_t>();
            }
        }
        return 0;
    }
    catch (...) {
        return -1;
    }
}

// ---- CPU baseline timers (bench.py --impl reference / cpu_baseline) --------------------------------
// `count` EvalBinGate(gate) on count ciphertext pairs, outer OMP loop over independent gates — the shape
// of EvalBinGateBatch (src/binfhe/lib/batch/batch.cpp:197-200). Returns seconds.
double ref_time_bingate(void* hv, int gate, const uint64_t* a1, const uint64_t* b1, const uint64_t* a2,
                        const uint64_t* b2, uint64_t* oa, uint64_t* ob, uint64_t count, int threads) {
    omp_set_max_active_levels(1);
    auto t0 = std::chrono::steady_clock::now();
    ref_evalbingate(hv, gate, a1, b1, a2, b2, oa, ob, count, 0, threads);
    auto t1 = std::chrono::steady_clock::now();
    return std::chrono::duration<double>(t1 - t0).count();
}

double ref_time_ntt(uint64_t Q, uint64_t psi, uint32_t N, uint64_t* data, uint64_t count, int inverse,
                    int threads) {
    // warm tables
    ref_ntt(Q, psi, N, data, 1, 0, 1);
    ref_ntt(Q, psi, N, data, 1, 1, 1);
    NativeInteger mod(Q), root(psi);
    std::vector<NativeVector> vs;
    vs.reserve(count);
    for (uint64_t i = 0; i < count; ++i) {
        NativeVector v(N, mod);
        for (uint32_t k = 0; k < N; ++k)
            v[k] = NativeInteger(data[i * N + k]);
        vs.push_back(std::move(v));
    }
    auto t0 = std::chrono::steady_clock::now();
#pragma omp parallel for num_threads(threads > 0 ? threads : 1)
    for (uint64_t i = 0; i < count; ++i) {
        if (inverse)
            ChineseRemainderTransformFTT<NativeVector>().InverseTransformFromBitReverseInPlace(root, 2 * N, &vs[i]);
        else
            ChineseRemainderTransformFTT<NativeVector>().ForwardTransformToBitReverseInPlace(root, 2 * N, &vs[i]);
    }
    auto t1 = std::chrono::steady_clock::now();
    for (uint64_t i = 0; i < count; ++i)
        for (uint32_t k = 0; k < N; ++k)
            data[i * N + k] = vs[i][k].ConvertToInt();
    return std::chrono::duration<double>(t1 - t0).count();
}

int ref_max_threads() {
    return omp_get_max_threads();
}

}  // extern "C"
