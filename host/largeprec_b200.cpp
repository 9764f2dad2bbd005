// largeprec_b200.cpp — see largeprec_b200.h.
#include "largeprec_b200.h"

#include <luxb200.h>

#include <cstring>
#include <map>
#include <mutex>
#include <stdexcept>

#include "backend_b200.h"

namespace lux::fhe {

namespace {

inline uint64_t u64(const NativeInteger& v) { return v.ConvertToInt<uint64_t>(); }

struct BaseCtx {
    luxb200_ctx* h = nullptr;
    RingGSWACCKey bsk;      // the key objects resident on the device (held: their addresses are the cache identity)
    LWESwitchingKey ksk;
};

struct KeyMapB200 {
    std::mutex mu;
    luxb200_params p{};     // parameters with baseG = 0
    int device = 0;
    std::map<uint32_t, BaseCtx> ctx;   // by gadget base

    void release() {
        for (auto& kv : ctx)
            if (kv.second.h) luxb200_ctx_destroy(kv.second.h);
        ctx.clear();
    }
    ~KeyMapB200() { release(); }
};

KeyMapB200& keymap() {
    static KeyMapB200 km;
    return km;
}

void check(int rc, const char* what) {
    if (rc != LUXB200_OK) throw std::runtime_error(std::string(what) + ": " + luxb200_last_error());
}

luxb200_params params_of(BinFHEContext& cc, uint32_t baseG) {
    auto lwe = cc.GetParams()->GetLWEParams();
    auto rg  = cc.GetParams()->GetRingGSWParams();
    luxb200_params p{};
    p.n = lwe->Getn(); p.N = lwe->GetN();
    p.q = u64(lwe->Getq()); p.Q = u64(lwe->GetQ()); p.qKS = u64(lwe->GetqKS());
    p.baseKS = u64(lwe->GetBaseKS()); p.baseG = baseG; p.baseR = rg->GetBaseR();
    p.numAutoKeys = rg->GetNumAutoKeys(); p.method = static_cast<uint32_t>(rg->GetMethod());
    p.psi = u64(rg->GetPolyParams()->GetRootOfUnity());
    return p;
}

// GINX refresh key -> canonical flat layout [n][2][rows][2][N] (include/luxb200.h)
void flatten_ginx(const RingGSWACCKeyImpl& ek, uint32_t n, uint32_t N, std::vector<uint64_t>& out) {
    size_t pos = 0;
    for (uint32_t i = 0; i < n; ++i)
        for (uint32_t s = 0; s < 2; ++s) {
            const auto& k = ek[0][s][i];
            if (!k) throw std::invalid_argument("refresh key has an empty entry");
            for (const auto& row : k->GetElements())
                for (const auto& poly : row) {
                    if (pos + N > out.size()) throw std::logic_error("refresh key does not match the parameter set");
                    std::memcpy(&out[pos], &poly.GetValues()[0], size_t(N) * 8);
                    pos += N;
                }
        }
    if (pos != out.size()) throw std::logic_error("refresh key does not match the parameter set");
}

void flatten_ksk(const LWESwitchingKeyImpl& ks, std::vector<uint64_t>& A, std::vector<uint64_t>& B) {
    const auto& kA = ks.GetElementsA();
    const auto& kB = ks.GetElementsB();
    size_t pa = 0, pb = 0;
    for (size_t i = 0; i < kA.size(); ++i)
        for (size_t j = 0; j < kA[i].size(); ++j)
            for (size_t k = 0; k < kA[i][j].size(); ++k) {
                const NativeVector& v = kA[i][j][k];
                if (pa + v.GetLength() > A.size() || pb >= B.size()) throw std::logic_error("switching key does not match the parameter set");
                std::memcpy(&A[pa], &v[0], size_t(v.GetLength()) * 8);
                pa += v.GetLength();
                B[pb++] = u64(kB[i][j][k]);
            }
}

// device contexts for every base of cc's key map, the current base first
std::vector<luxb200_ctx*> attach(BinFHEContext& cc) {
    KeyMapB200& km = keymap();
    if (cc.GetParams()->GetRingGSWParams()->GetMethod() != GINX) throw std::invalid_argument("CGGI is the only supported method");
    auto* be = dynamic_cast<backend::BackendB200*>(backend::BackendRegistry::Instance().Get(backend::BackendType::CUDA));
    const int device = be ? be->device() : 0;
    const luxb200_params base0 = params_of(cc, 0);
    if (std::memcmp(&base0, &km.p, sizeof(base0)) != 0 || device != km.device) {
        km.release();
        km.p = base0;
        km.device = device;
    }
    const auto map = cc.GetBTKeyMap();
    if (!map || map->empty())
        throw std::invalid_argument("Bootstrapping keys have not been generated. Please call BTKeyGen before calling bootstrapping.");
    const uint32_t cur = cc.GetParams()->GetRingGSWParams()->GetBaseG();
    if (map->find(cur) == map->end()) throw std::invalid_argument("ERROR: No key [" + std::to_string(cur) + "] found in the map");
    std::vector<uint32_t> bases{cur};
    for (const auto& kv : *map)
        if (kv.first != cur) bases.push_back(kv.first);
    std::vector<luxb200_ctx*> out;
    for (uint32_t bg : bases) {
        const RingGSWBTKey& key = map->at(bg);
        BaseCtx& bc = km.ctx[bg];
        if (!bc.h) {
            const luxb200_params p = params_of(cc, bg);
            check(luxb200_ctx_create(&p, device, &bc.h), "luxb200_ctx_create");
        }
        if (bc.bsk != key.BSkey) {
            std::vector<uint64_t> flat(luxb200_bsk_len(bc.h));
            flatten_ginx(*key.BSkey, base0.n, base0.N, flat);
            check(luxb200_load_bsk(bc.h, flat.data(), 0), "luxb200_load_bsk");
            bc.bsk = key.BSkey;
        }
        if (bc.ksk != key.KSkey) {   // every base has its own switching key (KeyGen draws one per call, binfhe-base-scheme.cpp:45-72)
            std::vector<uint64_t> A(luxb200_ksk_len(bc.h)), B(A.size() / base0.n);
            flatten_ksk(*key.KSkey, A, B);
            check(luxb200_load_ksk(bc.h, A.data(), B.data(), 0), "luxb200_load_ksk");
            bc.ksk = key.KSkey;
        }
        out.push_back(bc.h);
    }
    return out;
}

// all ciphertexts of a batch: dimension n and one (large) modulus
uint64_t flatten(const std::vector<LWECiphertext>& cts, uint32_t n, std::vector<uint64_t>& a, std::vector<uint64_t>& b) {
    if (!cts[0]) throw std::invalid_argument("Ciphertext is empty");
    const uint64_t mod = u64(cts[0]->GetModulus());
    a.resize(cts.size() * size_t(n));
    b.resize(cts.size());
    for (size_t i = 0; i < cts.size(); ++i) {
        if (!cts[i]) throw std::invalid_argument("Ciphertext is empty");
        if (cts[i]->GetLength() != n || u64(cts[i]->GetModulus()) != mod)
            throw std::invalid_argument("one batch = one ciphertext modulus and dimension n (split the batch)");
        std::memcpy(&a[i * size_t(n)], &cts[i]->GetA()[0], size_t(n) * 8);
        b[i] = u64(cts[i]->GetB());
    }
    return mod;
}

void build(const uint64_t* a, const uint64_t* b, size_t count, uint32_t n, uint64_t mod, LWECiphertext* out) {
    for (size_t i = 0; i < count; ++i) {
        NativeVector v(n, NativeInteger(mod));
        std::memcpy(&v[0], a + i * size_t(n), size_t(n) * 8);
        out[i] = std::make_shared<LWECiphertextImpl>(std::move(v), NativeInteger(b[i]));
    }
}

template <typename F>
BatchResult guarded(size_t count, F&& body) {
    BatchResult result{true, 0, 0, ""};
    if (count == 0) return result;
    try {
        std::lock_guard<std::mutex> lk(keymap().mu);
        body();
        result.processed = count;
    } catch (const std::exception& e) {
        result.success = false;
        result.failed  = count;
        result.error   = e.what();
    }
    return result;
}

}  // namespace

void ReleaseLargePrecisionB200() {
    std::lock_guard<std::mutex> lk(keymap().mu);
    keymap().release();
    keymap().p = luxb200_params{};
}

BatchResult EvalFloorBatch(BinFHEContext& cc, const std::vector<LWECiphertext>& ct_in, std::vector<LWECiphertext>& ct_out,
                           uint32_t roundbits, uint32_t flags) {
    return guarded(ct_in.size(), [&] {
        ct_out.resize(ct_in.size());
        if (flags & BATCH_CPU_FORCE) {
            for (size_t i = 0; i < ct_in.size(); ++i) ct_out[i] = cc.EvalFloor(ct_in[i], roundbits);
            return;
        }
        auto hs = attach(cc);
        const uint32_t n = cc.GetParams()->GetLWEParams()->Getn();
        std::vector<uint64_t> a, b, oa(ct_in.size() * size_t(n)), ob(ct_in.size());
        const uint64_t mod = flatten(ct_in, n, a, b);
        check(luxb200_evalfloor_batch(hs[0], a.data(), b.data(), oa.data(), ob.data(), ct_in.size(), mod, u64(cc.GetBeta()), roundbits),
              "EvalFloorBatch");
        build(oa.data(), ob.data(), ct_in.size(), n, mod, ct_out.data());
    });
}

BatchResult EvalSignBatch(BinFHEContext& cc, const std::vector<LWECiphertext>& ct_in, std::vector<LWECiphertext>& ct_out,
                          bool schemeSwitch, uint32_t flags) {
    return guarded(ct_in.size(), [&] {
        ct_out.resize(ct_in.size());
        if (flags & BATCH_CPU_FORCE) {   // the reference call mutates shared parameters (Change_BaseG): one at a time
            for (size_t i = 0; i < ct_in.size(); ++i) ct_out[i] = cc.EvalSign(ct_in[i], schemeSwitch);
            return;
        }
        auto hs = attach(cc);
        const uint32_t n = cc.GetParams()->GetLWEParams()->Getn();
        std::vector<uint64_t> a, b, oa(ct_in.size() * size_t(n)), ob(ct_in.size());
        const uint64_t mod = flatten(ct_in, n, a, b);
        check(luxb200_evalsign_batch(hs.data(), static_cast<int>(hs.size()), a.data(), b.data(), oa.data(), ob.data(), ct_in.size(), mod,
                                     u64(cc.GetBeta()), schemeSwitch ? 1 : 0), "EvalSignBatch");
        build(oa.data(), ob.data(), ct_in.size(), n, u64(cc.GetParams()->GetLWEParams()->Getq()), ct_out.data());
    });
}

BatchResult EvalDecompBatch(BinFHEContext& cc, const std::vector<LWECiphertext>& ct_in,
                            std::vector<std::vector<LWECiphertext>>& ct_out, uint32_t flags) {
    return guarded(ct_in.size(), [&] {
        ct_out.assign(ct_in.size(), {});
        if (flags & BATCH_CPU_FORCE) {
            for (size_t i = 0; i < ct_in.size(); ++i) ct_out[i] = cc.EvalDecomp(ct_in[i]);
            return;
        }
        auto hs = attach(cc);
        const uint32_t n = cc.GetParams()->GetLWEParams()->Getn();
        std::vector<uint64_t> a, b;
        const uint64_t mod = flatten(ct_in, n, a, b);
        const uint64_t q = u64(cc.GetParams()->GetLWEParams()->Getq());
        if (mod <= q)
            throw std::invalid_argument("ERROR: EvalDecomp is only for large precision. For small precision, please use bootstrapping directly");
        const uint32_t k = luxb200_evaldecomp_count(q, mod, u64(cc.GetBeta()));
        std::vector<uint64_t> oa(size_t(k) * ct_in.size() * n), ob(size_t(k) * ct_in.size()), om(k);
        uint32_t got = 0;
        check(luxb200_evaldecomp_batch(hs.data(), static_cast<int>(hs.size()), a.data(), b.data(), oa.data(), ob.data(), om.data(), k, &got,
                                       ct_in.size(), mod, u64(cc.GetBeta())), "EvalDecompBatch");
        for (size_t i = 0; i < ct_in.size(); ++i) {
            ct_out[i].resize(got);
            for (uint32_t d = 0; d < got; ++d)
                build(&oa[(size_t(d) * ct_in.size() + i) * n], &ob[size_t(d) * ct_in.size() + i], 1, n, om[d], &ct_out[i][d]);
        }
    });
}

}  // namespace lux::fhe
