// backend_b200.cpp — see backend_b200.h.  Pack -> one luxb200_* call -> unpack; no arithmetic here.
#include "backend_b200.h"
#include "packed_b200.h"

#include <atomic>
#include <chrono>
#include <cstring>
#include <map>
#include <stdexcept>
#include <thread>

namespace lux::fhe {
namespace backend {

namespace {

inline uint64_t u64(const NativeInteger& v) { return v.ConvertToInt<uint64_t>(); }

static_assert(sizeof(NativeInteger) == sizeof(uint64_t), "NativeVector storage is one uint64 residue per element");

// [count][len] a-parts and [count] b-parts of a batch of LWE ciphertexts, written to caller-provided (page-locked) arrays.
// NativeVector storage is contiguous uint64 residues, so a row is one memcpy; rows are packed by all host threads.
// Every element is checked: non-null, dimension `len`, modulus `mod` (0 = do not check).
void flatten_cts_to(const LWECiphertext* cts, size_t count, uint32_t len, uint64_t mod, uint64_t* a, uint64_t* b) {
    std::atomic<long> bad{-1};
#pragma omp parallel for schedule(static)
    for (size_t i = 0; i < count; ++i) {
        const LWECiphertext& ct = cts[i];
        if (!ct || ct->GetLength() != len || (mod && u64(ct->GetModulus()) != mod)) {
            bad.store(static_cast<long>(i));
            continue;
        }
        const NativeVector& v = ct->GetA();
        std::memcpy(a + i * size_t(len), &v[0], size_t(len) * 8);
        b[i] = u64(ct->GetB());
    }
    if (bad.load() >= 0)
        throw std::invalid_argument("BackendB200: ciphertext " + std::to_string(bad.load()) +
                                    " is empty or has the wrong dimension / modulus for this operation");
}
void flatten_cts(const std::vector<LWECiphertext>& cts, uint32_t len, std::vector<uint64_t>& a, std::vector<uint64_t>& b) {
    a.resize(cts.size() * size_t(len));
    b.resize(cts.size());
    flatten_cts_to(cts.data(), cts.size(), len, 0, a.data(), b.data());
}

void build_cts_from(const uint64_t* a, const uint64_t* b, size_t count, uint32_t len, const NativeInteger& mod,
                    const NativeInteger& pmod, LWECiphertext* out) {
#pragma omp parallel for schedule(static)
    for (size_t i = 0; i < count; ++i) {
        NativeVector v(len, mod);
        std::memcpy(&v[0], a + i * size_t(len), size_t(len) * 8);
        out[i] = std::make_shared<LWECiphertextImpl>(std::move(v), NativeInteger(b[i]));
        out[i]->SetptModulus(pmod);
    }
}
void build_cts(const std::vector<uint64_t>& a, const std::vector<uint64_t>& b, uint32_t len, const NativeInteger& mod,
               const NativeInteger& pmod, std::vector<LWECiphertext>& out) {
    out.resize(b.size());
    build_cts_from(a.data(), b.data(), b.size(), len, mod, pmod, out.data());
}

double now_ms() {
    return std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now().time_since_epoch()).count();
}

void flatten_rlwe(const RLWECiphertext& ct, uint32_t N, uint64_t* out) {
    const auto& el = ct->GetElements();
    if (el.size() != 2) throw std::invalid_argument("BackendB200: RLWE ciphertext must have 2 components");
    for (int c = 0; c < 2; ++c) {
        if (el[c].GetFormat() != Format::EVALUATION || el[c].GetLength() != N)
            throw std::invalid_argument("BackendB200: RLWE components must be EVALUATION-format polynomials of length N");
        for (uint32_t k = 0; k < N; ++k) out[c * N + k] = u64(el[c][k]);
    }
}

}  // namespace

BackendB200::BackendB200(int device) : devices_{device} {}
BackendB200::BackendB200(std::vector<int> devices) : devices_(std::move(devices)) {
    if (devices_.empty())
        for (int d = 0; d < luxb200_device_count(); ++d) devices_.push_back(d);
    if (devices_.empty()) devices_.push_back(0);
}

void BackendB200::destroy_contexts() {
    for (luxb200_ctx* h : hs_)
        if (h) luxb200_ctx_destroy(h);
    hs_.clear();
    h_ = nullptr;
}

BackendB200::~BackendB200() {
    destroy_contexts();
    for (Pinned& b : st_) b.release();
}

uint64_t* BackendB200::Pinned::ensure(size_t elems) {
    if (elems <= cap) return p;
    release();
    void* q = nullptr;
    if (luxb200_host_alloc(std::max<size_t>(elems, 1) * 8, &q) != LUXB200_OK)
        throw std::runtime_error(std::string("BackendB200: pinned staging allocation failed: ") + luxb200_last_error());
    p = static_cast<uint64_t*>(q);
    cap = elems;
    return p;
}
void BackendB200::Pinned::release() {
    if (p) luxb200_host_free(p);
    p = nullptr;
    cap = 0;
}

void BackendB200::check(int rc, const char* what) const {
    if (rc != LUXB200_OK) throw std::runtime_error(std::string("BackendB200::") + what + ": " + luxb200_last_error());
}

bool BackendB200::IsAvailable() const {
    const int cnt = luxb200_device_count();
    for (int d : devices_)
        if (d < 0 || d >= cnt) return false;
    return true;
}

size_t BackendB200::DeviceMemory() const {
    uint64_t f = 0, t = 0;
    return luxb200_dev_mem_info(devices_[0], &f, &t) == LUXB200_OK ? size_t(t) : 0;
}

static luxb200_params params_of(BinFHEContext& cc) {
    auto lwe = cc.GetParams()->GetLWEParams();
    auto rg  = cc.GetParams()->GetRingGSWParams();
    luxb200_params p{};
    p.n = lwe->Getn(); p.N = lwe->GetN();
    p.q = u64(lwe->Getq()); p.Q = u64(lwe->GetQ()); p.qKS = u64(lwe->GetqKS());
    p.baseKS = u64(lwe->GetBaseKS()); p.baseG = rg->GetBaseG(); p.baseR = rg->GetBaseR();
    p.numAutoKeys = rg->GetNumAutoKeys(); p.method = static_cast<uint32_t>(rg->GetMethod());
    p.psi = u64(rg->GetPolyParams()->GetRootOfUnity());
    return p;
}

void BackendB200::Bind(BinFHEContext& cc) {
    std::lock_guard<std::recursive_mutex> lk(mu_);
    auto rg = cc.GetParams()->GetRingGSWParams();
    const luxb200_params p = params_of(cc);
    destroy_contexts();
    hs_.assign(devices_.size(), nullptr);
    for (size_t d = 0; d < devices_.size(); ++d) check(luxb200_ctx_create(&p, devices_[d], &hs_[d]), "Bind");
    h_ = hs_[0];
    p_ = p;
    digitsG_ = rg->GetDigitsG(); digitsR_ = static_cast<uint32_t>(rg->GetDigitsR().size());
    rows_ = 2 * (digitsG_ - 1);
    poly_params_ = rg->GetPolyParams();
    bound_bsk_.reset(); bound_ksk_.reset();
    packed_bsk_ = packed_ksk_ = false;
    if (cc.GetRefreshKey() && cc.GetSwitchKey()) load_keys(cc.GetRefreshKey(), cc.GetSwitchKey());
}

void BackendB200::Attach(BinFHEContext& cc) {
    std::lock_guard<std::recursive_mutex> lk(mu_);
    if (!cc.GetParams()) throw std::invalid_argument("BinFHECryptoParams is empty");
    const luxb200_params p = params_of(cc);
    if (!h_ || std::memcmp(&p, &p_, sizeof(p)) != 0) {
        Bind(cc);   // another parameter set (a second BinFHEContext): new device contexts, keys reloaded below
        return;
    }
    // a context that holds keys must see exactly those on the device; one without keys keeps LoadPacked keys
    if (cc.GetRefreshKey()) ensure_bsk(cc.GetRefreshKey());
    if (cc.GetSwitchKey()) ensure_ksk(cc.GetSwitchKey());
}

// RingGSWACCKeyImpl -> canonical flat layout of include/luxb200.h (the orders of rgsw-acc-cggi.cpp:39-57,
// rgsw-acc-dm.cpp:39-60, rgsw-acc-lmkcdey.cpp:39-68)
void BackendB200::flatten_bsk(const RingGSWACCKeyImpl& ek, std::vector<uint64_t>& out) const {
    out.assign(luxb200_bsk_len(h_), 0);
    const uint32_t N = p_.N, n = p_.n;
    // slot list first, then all host threads copy polynomials (a 1 GB key for STD128_AP)
    std::vector<std::pair<const RingGSWEvalKeyImpl*, size_t>> slots;
    size_t pos = 0;
    auto put = [&](const RingGSWEvalKey& k) {
        if (!k) throw std::invalid_argument("BackendB200: refresh key has an empty entry where one is required");
        slots.emplace_back(k.get(), pos);
        for (const auto& row : k->GetElements()) pos += row.size() * size_t(N);
    };
    switch (p_.method) {
        case LUXB200_GINX:
            for (uint32_t i = 0; i < n; ++i) { put(ek[0][0][i]); put(ek[0][1][i]); }
            break;
        case LUXB200_AP:
            for (uint32_t i = 0; i < n; ++i)
                for (uint32_t j = 0; j < p_.baseR; ++j)
                    for (uint32_t k = 0; k < digitsR_; ++k) {
                        if (j == 0) pos += size_t(rows_) * 2 * N;   // digit 0 is never used (rgsw-acc-dm.cpp:47-58)
                        else put(ek[i][j][k]);
                    }
            break;
        default:
            for (uint32_t i = 0; i < n; ++i) put(ek[0][0][i]);
            for (uint32_t t = 0; t <= p_.numAutoKeys; ++t) put(ek[0][1][t]);
    }
    if (pos != out.size()) throw std::logic_error("BackendB200: refresh key does not match the parameter set");
#pragma omp parallel for schedule(dynamic, 16)
    for (size_t s = 0; s < slots.size(); ++s) {
        size_t o = slots[s].second;
        for (const auto& row : slots[s].first->GetElements())
            for (const auto& poly : row) {
                std::memcpy(&out[o], &poly.GetValues()[0], size_t(N) * 8);
                o += N;
            }
    }
}

// every device holds the keys of device 0 (peer-to-peer over NVLink, luxb200_broadcast_keys)
void BackendB200::share_keys() {
    if (hs_.size() > 1) check(luxb200_broadcast_keys(hs_.data(), static_cast<int>(hs_.size())), "broadcast_keys");
}

void BackendB200::ensure_bsk(const RingGSWACCKey& ek) {
    if (!ek) throw std::invalid_argument("Bootstrapping keys have not been generated. Please call BTKeyGen before calling bootstrapping.");
    std::lock_guard<std::recursive_mutex> lk(mu_);
    if (!packed_bsk_ && ek == bound_bsk_) return;   // the held shared_ptr pins the address: identity cannot be recycled
    std::vector<uint64_t> flat;
    flatten_bsk(*ek, flat);
    check(luxb200_load_bsk(h_, flat.data(), 0), "load_bsk");
    bound_bsk_ = ek;
    packed_bsk_ = false;
    share_keys();
}

void BackendB200::ensure_ksk(const LWESwitchingKey& ks) {
    if (!ks) throw std::invalid_argument("BackendB200: switching key is empty");
    std::lock_guard<std::recursive_mutex> lk(mu_);
    if (!packed_ksk_ && ks == bound_ksk_) return;
    const auto& kA = ks->GetElementsA();
    const auto& kB = ks->GetElementsB();
    std::vector<uint64_t> A(luxb200_ksk_len(h_)), B(A.size() / p_.n);
    size_t pa = 0, pb = 0;
    for (size_t i = 0; i < kA.size(); ++i)
        for (size_t j = 0; j < kA[i].size(); ++j)
            for (size_t k = 0; k < kA[i][j].size(); ++k) {
                const NativeVector& v = kA[i][j][k];
                if (pa + v.GetLength() > A.size() || pb >= B.size())
                    throw std::logic_error("BackendB200: switching key does not match the parameter set");
                std::memcpy(&A[pa], &v[0], size_t(v.GetLength()) * 8);
                pa += v.GetLength();
                B[pb++] = u64(kB[i][j][k]);
            }
    check(luxb200_load_ksk(h_, A.data(), B.data(), 0), "load_ksk");
    bound_ksk_ = ks;
    packed_ksk_ = false;
    share_keys();
}

void BackendB200::load_keys(const RingGSWACCKey& ek, const LWESwitchingKey& ks) {
    ensure_bsk(ek);
    ensure_ksk(ks);
}

void BackendB200::run_split(size_t count, const std::function<void(size_t, size_t, size_t)>& body) {
    const size_t G = std::min(hs_.size(), std::max<size_t>(count, 1));
    if (G <= 1) { body(0, 0, count); return; }
    std::vector<std::thread> th;
    std::vector<std::exception_ptr> err(G);
    const size_t base = count / G, rem = count % G;
    for (size_t d = 0; d < G; ++d) {
        const size_t lo = d * base + std::min(d, rem), hi = lo + base + (d < rem ? 1 : 0);
        th.emplace_back([&, d, lo, hi] {
            try { if (hi > lo) body(d, lo, hi); } catch (...) { err[d] = std::current_exception(); }
        });
    }
    for (auto& t : th) t.join();
    for (auto& e : err)
        if (e) std::rethrow_exception(e);
}

// ---- memory (backend.h:88-113) ----------------------------------------------------------------------------
DeviceBuffer BackendB200::Allocate(size_t bytes) {
    DeviceBuffer b;
    void* p = nullptr;
    check(luxb200_dev_alloc(devices_[0], bytes, &p), "Allocate");
    b.ptr = p; b.size = bytes; b.device = BackendType::CUDA;
    return b;
}
void BackendB200::Free(DeviceBuffer& buffer) {
    if (buffer.ptr) check(luxb200_dev_free(devices_[0], buffer.ptr), "Free");
    buffer.ptr = nullptr; buffer.size = 0;
}
void BackendB200::CopyToDevice(const void* host_ptr, DeviceBuffer& device_buffer, size_t bytes) {
    if (bytes > device_buffer.size) throw std::invalid_argument("BackendB200::CopyToDevice: buffer too small");
    check(luxb200_dev_copy(devices_[0], device_buffer.ptr, host_ptr, bytes, 1), "CopyToDevice");
}
void BackendB200::CopyToHost(const DeviceBuffer& device_buffer, void* host_ptr, size_t bytes) {
    if (bytes > device_buffer.size) throw std::invalid_argument("BackendB200::CopyToHost: buffer too small");
    check(luxb200_dev_copy(devices_[0], host_ptr, device_buffer.ptr, bytes, 2), "CopyToHost");
}
void BackendB200::Synchronize() {
    for (luxb200_ctx* h : hs_) check(luxb200_synchronize(h), "Synchronize");
}

// ---- blind rotation: RingGSWAccumulator::EvalAcc (rgsw-acc.h:73-76) on EVALUATION-format accumulators -----
void BackendB200::BlindRotateBatch(const std::shared_ptr<RingGSWCryptoParams>&, const std::vector<LWECiphertext>& cts,
                                   const RingGSWACCKey& ek, std::vector<RLWECiphertext>& accs) {
    if (cts.size() != accs.size()) throw std::invalid_argument("BackendB200::BlindRotateBatch: size mismatch");
    if (cts.empty()) return;
    ensure_bsk(ek);
    const uint32_t n = p_.n, N = p_.N;
    std::vector<uint64_t> a, b, acc(cts.size() * 2 * size_t(N));
    flatten_cts(cts, n, a, b);
    for (size_t i = 0; i < accs.size(); ++i) flatten_rlwe(accs[i], N, &acc[i * 2 * N]);
    check(luxb200_blind_rotate_batch(h_, a.data(), u64(cts[0]->GetModulus()), acc.data(), cts.size()), "BlindRotateBatch");
    for (size_t i = 0; i < accs.size(); ++i) {
        auto& el = accs[i]->GetElements();
        for (int c = 0; c < 2; ++c)
            for (uint32_t k = 0; k < N; ++k) el[c][k] = NativeInteger(acc[(i * 2 + c) * N + k]);
    }
}
void BackendB200::BlindRotate(const std::shared_ptr<RingGSWCryptoParams>& params, const LWECiphertext& ct,
                              const RingGSWACCKey& ek, RLWECiphertext& acc) {
    std::vector<LWECiphertext> cts{ct};
    std::vector<RLWECiphertext> accs{acc};
    BlindRotateBatch(params, cts, ek, accs);
}

// ---- external product: one AddToAccDM step (rgsw-acc-dm.cpp:118-143) -----------------------------------------
void BackendB200::ExternalProductBatch(const std::shared_ptr<RingGSWCryptoParams>&, const std::vector<RingGSWEvalKey>& rgsws,
                                       const std::vector<RLWECiphertext>& rlwes, std::vector<RLWECiphertext>& results) {
    if (rgsws.size() != rlwes.size()) throw std::invalid_argument("BackendB200::ExternalProductBatch: size mismatch");
    if (rgsws.empty()) { results.clear(); return; }
    const uint32_t N = p_.N;
    const size_t per_key = size_t(rows_) * 2 * N;
    std::vector<uint64_t> key(rgsws.size() * per_key), in(rlwes.size() * 2 * size_t(N)), out(in.size());
    for (size_t i = 0; i < rgsws.size(); ++i) {
        const auto& el = rgsws[i]->GetElements();
        if (el.size() != rows_) throw std::invalid_argument("BackendB200::ExternalProductBatch: RGSW ciphertext has the wrong number of rows");
        size_t pos = i * per_key;
        for (const auto& row : el)
            for (const auto& poly : row)
                for (uint32_t t = 0; t < N; ++t) key[pos++] = u64(poly[t]);
        flatten_rlwe(rlwes[i], N, &in[i * 2 * N]);
    }
    check(luxb200_external_product_batch(h_, key.data(), in.data(), out.data(), rgsws.size()), "ExternalProductBatch");
    results.resize(rlwes.size());
    for (size_t i = 0; i < results.size(); ++i) {
        std::vector<NativePoly> res(2);
        for (int c = 0; c < 2; ++c) {
            res[c] = NativePoly(poly_params_, Format::EVALUATION, true);
            for (uint32_t k = 0; k < N; ++k) res[c][k] = NativeInteger(out[(i * 2 + c) * N + k]);
        }
        results[i] = std::make_shared<RLWECiphertextImpl>(std::move(res));
    }
}
void BackendB200::ExternalProduct(const std::shared_ptr<RingGSWCryptoParams>& params, const RingGSWEvalKey& rgsw,
                                  const RLWECiphertext& rlwe, RLWECiphertext& result) {
    std::vector<RingGSWEvalKey> k{rgsw};
    std::vector<RLWECiphertext> in{rlwe}, out;
    ExternalProductBatch(params, k, in, out);
    result = out[0];
}

// ---- LWE key switch (lwe-pke.cpp:348-372): (N, qKS) -> (n, qKS) ----------------------------------------------
void BackendB200::KeySwitchBatch(const std::shared_ptr<LWECryptoParams>&, const std::vector<LWECiphertext>& cts,
                                 const LWESwitchingKey& ks, std::vector<LWECiphertext>& results) {
    if (cts.empty()) { results.clear(); return; }
    ensure_ksk(ks);
    std::vector<uint64_t> a, b, oa(cts.size() * size_t(p_.n)), ob(cts.size());
    flatten_cts(cts, p_.N, a, b);
    check(luxb200_keyswitch_batch(h_, a.data(), b.data(), oa.data(), ob.data(), cts.size()), "KeySwitchBatch");
    build_cts(oa, ob, p_.n, NativeInteger(p_.qKS), cts[0]->GetptModulus(), results);
}
void BackendB200::KeySwitch(const std::shared_ptr<LWECryptoParams>& params, const LWECiphertext& ct, const LWESwitchingKey& ks,
                            LWECiphertext& result) {
    std::vector<LWECiphertext> in{ct}, out;
    KeySwitchBatch(params, in, ks, out);
    result = out[0];
}

// ---- modulus switch to q (lwe-pke.cpp:254-261, RoundqQ :41-46) ------------------------------------------------
void BackendB200::ModSwitchBatch(const std::shared_ptr<LWECryptoParams>& params, const std::vector<LWECiphertext>& cts,
                                 std::vector<LWECiphertext>& results) {
    if (cts.empty()) { results.clear(); return; }
    const uint32_t len = cts[0]->GetLength();
    const uint64_t qfrom = u64(cts[0]->GetModulus()), qto = u64(params->Getq());
    std::vector<uint64_t> a, b;
    flatten_cts(cts, len, a, b);
    std::vector<uint64_t> in(a.size() + b.size()), out(in.size());
    std::memcpy(in.data(), a.data(), a.size() * 8);
    std::memcpy(in.data() + a.size(), b.data(), b.size() * 8);
    check(luxb200_modswitch_batch(h_, qto, qfrom, in.data(), out.data(), in.size()), "ModSwitchBatch");
    std::memcpy(a.data(), out.data(), a.size() * 8);
    std::memcpy(b.data(), out.data() + a.size(), b.size() * 8);
    build_cts(a, b, len, NativeInteger(qto), cts[0]->GetptModulus(), results);
}
void BackendB200::ModSwitch(const std::shared_ptr<LWECryptoParams>& params, const LWECiphertext& ct, LWECiphertext& result) {
    std::vector<LWECiphertext> in{ct}, out;
    ModSwitchBatch(params, in, out);
    result = out[0];
}

// ---- packed formats (backend.h:213-246) -----------------------------------------------------------------------
// Packed wire bytes (backend/packed.h formats, host/packed_b200.cpp) -> canonical flat arrays -> device, without
// rebuilding the RingGSWACCKey / LWESwitchingKey object graphs.
void BackendB200::LoadPacked(const uint8_t* bsk, size_t bsk_size, const uint8_t* ksk, size_t ksk_size) {
    if (!h_) throw std::runtime_error("BackendB200::LoadPacked: Bind() first");
    if (bsk) {
        PackedBSKView v;
        if (!ParsePackedBSK(bsk, bsk_size, v)) throw std::invalid_argument("BackendB200::LoadPacked: invalid packed bootstrapping key");
        if (v.head->rlwe_N != p_.N || v.layout->Q != p_.Q || v.layout->method != p_.method || v.head->lwe_n != p_.n)
            throw std::invalid_argument("BackendB200::LoadPacked: packed bootstrapping key does not match the bound parameter set");
        const uint32_t d1 = v.layout->dims[1], d2 = v.layout->dims[2];
        std::vector<size_t> off(v.slots + 1, 0);   // coefficient offset of every slot
        for (size_t s = 0; s < v.slots; ++s) off[s + 1] = off[s] + size_t(v.rows[s]) * 2 * p_.N;
        std::vector<uint64_t> flat(luxb200_bsk_len(h_), 0);
        size_t pos = 0;
        auto put = [&](uint32_t i, uint32_t j, uint32_t k, uint32_t want_rows) {
            const size_t s = (size_t(i) * d1 + j) * d2 + k;
            if (i >= v.layout->dims[0] || j >= d1 || k >= d2 || v.rows[s] != want_rows || pos + off[s + 1] - off[s] > flat.size())
                throw std::invalid_argument("BackendB200::LoadPacked: packed bootstrapping key has the wrong shape for this method");
            for (size_t t = off[s]; t < off[s + 1]; ++t) flat[pos++] = v.coeff(t);
        };
        switch (p_.method) {
            case LUXB200_GINX:
                for (uint32_t i = 0; i < p_.n; ++i) { put(0, 0, i, rows_); put(0, 1, i, rows_); }
                break;
            case LUXB200_AP:
                for (uint32_t i = 0; i < p_.n; ++i)
                    for (uint32_t j = 0; j < p_.baseR; ++j)
                        for (uint32_t k = 0; k < digitsR_; ++k) {
                            if (j == 0) pos += size_t(rows_) * 2 * p_.N;
                            else put(i, j, k, rows_);
                        }
                break;
            default:
                for (uint32_t i = 0; i < p_.n; ++i) put(0, 0, i, rows_);
                for (uint32_t t = 0; t <= p_.numAutoKeys; ++t) put(0, 1, t, digitsG_ - 1);
        }
        if (pos != flat.size()) throw std::invalid_argument("BackendB200::LoadPacked: packed bootstrapping key is incomplete");
        check(luxb200_load_bsk(h_, flat.data(), 0), "load_bsk");
        bound_bsk_.reset();
        packed_bsk_ = true;
    }
    if (ksk) {
        PackedKSKView v;
        if (!ParsePackedKSK(ksk, ksk_size, v)) throw std::invalid_argument("BackendB200::LoadPacked: invalid packed switching key");
        std::vector<uint64_t> A(luxb200_ksk_len(h_)), B(A.size() / p_.n);
        if (v.head->input_n != p_.N || v.head->output_n != p_.n || v.head->Q != p_.qKS || v.base != p_.baseKS || v.vecs != B.size())
            throw std::invalid_argument("BackendB200::LoadPacked: packed switching key does not match the bound parameter set");
        for (size_t t = 0; t < A.size(); ++t) A[t] = v.coeff(v.A, t);
        for (size_t t = 0; t < B.size(); ++t) B[t] = v.coeff(v.B, t);
        check(luxb200_load_ksk(h_, A.data(), B.data(), 0), "load_ksk");
        bound_ksk_.reset();
        packed_ksk_ = true;
    }
    share_keys();
}

DeviceBuffer BackendB200::PackBootstrappingKey(const RingGSWACCKey& ek) {
    if (!ek) throw std::invalid_argument("BackendB200::PackBootstrappingKey: empty key");
    std::vector<uint64_t> flat;
    flatten_bsk(*ek, flat);
    DeviceBuffer buf = Allocate(flat.size() * 8);
    CopyToDevice(flat.data(), buf, flat.size() * 8);
    return buf;   // accepted as-is by luxb200_load_bsk(ctx, ptr, /*on_device=*/1) and by an NCCL broadcast
}
void BackendB200::UnpackBootstrappingKey(const DeviceBuffer& packed, RingGSWACCKey& ek) {
    std::vector<uint64_t> flat(luxb200_bsk_len(h_));
    if (packed.size != flat.size() * 8) throw std::invalid_argument("BackendB200::UnpackBootstrappingKey: size mismatch");
    CopyToHost(packed, flat.data(), packed.size);
    unflatten_bsk(flat, ek);
}

// canonical flat layout -> the reference's RingGSWACCKey object graph
void BackendB200::unflatten_bsk(const std::vector<uint64_t>& flat, RingGSWACCKey& ek) const {
    const uint32_t N = p_.N, n = p_.n;
    size_t pos = 0;
    auto take = [&]() {
        auto k = std::make_shared<RingGSWEvalKeyImpl>(rows_, 2);
        for (uint32_t r = 0; r < rows_; ++r)
            for (uint32_t c = 0; c < 2; ++c) {
                NativePoly poly(poly_params_, Format::EVALUATION, true);
                for (uint32_t t = 0; t < N; ++t) poly[t] = NativeInteger(flat[pos++]);
                (*k)[r][c] = std::move(poly);
            }
        return k;
    };
    auto take_auto = [&]() {   // automorphism keys have digitsG-1 rows
        auto k = std::make_shared<RingGSWEvalKeyImpl>(digitsG_ - 1, 2);
        for (uint32_t r = 0; r < digitsG_ - 1; ++r)
            for (uint32_t c = 0; c < 2; ++c) {
                NativePoly poly(poly_params_, Format::EVALUATION, true);
                for (uint32_t t = 0; t < N; ++t) poly[t] = NativeInteger(flat[pos++]);
                (*k)[r][c] = std::move(poly);
            }
        return k;
    };
    switch (p_.method) {
        case LUXB200_GINX:
            ek = std::make_shared<RingGSWACCKeyImpl>(1, 2, n);
            for (uint32_t i = 0; i < n; ++i) { (*ek)[0][0][i] = take(); (*ek)[0][1][i] = take(); }
            break;
        case LUXB200_AP:
            ek = std::make_shared<RingGSWACCKeyImpl>(n, p_.baseR, digitsR_);
            for (uint32_t i = 0; i < n; ++i)
                for (uint32_t j = 0; j < p_.baseR; ++j)
                    for (uint32_t k = 0; k < digitsR_; ++k) {
                        if (j == 0) pos += size_t(rows_) * 2 * N;
                        else (*ek)[i][j][k] = take();
                    }
            break;
        default:
            ek = std::make_shared<RingGSWACCKeyImpl>(1, 2, std::max(n, p_.numAutoKeys + 1));
            for (uint32_t i = 0; i < n; ++i) (*ek)[0][0][i] = take();
            for (uint32_t t = 0; t <= p_.numAutoKeys; ++t) (*ek)[0][1][t] = take_auto();
    }
}
DeviceBuffer BackendB200::PackCiphertexts(const std::vector<LWECiphertext>& cts) {
    if (cts.empty()) return DeviceBuffer{};
    const uint32_t len = cts[0]->GetLength();
    std::vector<uint64_t> a, b, flat(cts.size() * size_t(len + 1));
    flatten_cts(cts, len, a, b);
    for (size_t i = 0; i < cts.size(); ++i) {
        std::memcpy(&flat[i * (len + 1)], &a[i * len], len * 8);
        flat[i * (len + 1) + len] = b[i];
    }
    DeviceBuffer buf = Allocate(flat.size() * 8);
    CopyToDevice(flat.data(), buf, flat.size() * 8);
    return buf;
}
void BackendB200::UnpackCiphertexts(const DeviceBuffer& packed, std::vector<LWECiphertext>& cts) {
    // ciphertexts of the bound context: dimension n, modulus q
    const uint32_t len = p_.n;
    if (packed.size % (size_t(len + 1) * 8)) throw std::invalid_argument("BackendB200::UnpackCiphertexts: size mismatch");
    const size_t count = packed.size / (size_t(len + 1) * 8);
    std::vector<uint64_t> flat(count * (len + 1)), a(count * len), b(count);
    CopyToHost(packed, flat.data(), packed.size);
    for (size_t i = 0; i < count; ++i) {
        std::memcpy(&a[i * len], &flat[i * (len + 1)], len * 8);
        b[i] = flat[i * (len + 1) + len];
    }
    build_cts(a, b, len, NativeInteger(p_.q), NativeInteger(4), cts);
}

// ---- gate-level entry points (BinFHEScheme::EvalBinGate / Bootstrap / EvalFunc, binfhe-base-scheme.cpp:76-332) --
// One homogeneous group of 2-input gates: every ct1 one format (SMALL_DIM or extended), every ct2 one format.
// The plaintext modulus handed to the device is always 4: EvalBinGate bootstraps a COPY of ct1 (binfhe-base-scheme.cpp:92)
// and LWECiphertextImpl's copy constructor does not carry m_p (lwe-ciphertext.h:56), so whatever ct1->GetptModulus() says,
// BootstrapGateCore sees 4 (host_parity "EvalBinGateBatch_ptmodulus" checks exactly this against the reference).
void BackendB200::gate2(BINGATE gate, const std::vector<LWECiphertext>& ct1, const std::vector<LWECiphertext>& ct2,
                        std::vector<LWECiphertext>& out) {
    const size_t cnt = ct1.size();
    const uint32_t n = p_.n, N = p_.N;
    const bool e1 = u64(ct1[0]->GetModulus()) == p_.Q, e2 = u64(ct2[0]->GetModulus()) == p_.Q;   // binfhe-base-scheme.cpp:92-93
    const uint32_t l1 = e1 ? N : n, l2 = e2 ? N : n;
    const uint64_t pm = 4;
    const double t0 = now_ms();
    uint64_t *a1 = st_[0].ensure(cnt * size_t(l1)), *b1 = st_[1].ensure(cnt), *a2 = st_[2].ensure(cnt * size_t(l2)),
             *b2 = st_[3].ensure(cnt), *oa = st_[4].ensure(cnt * size_t(n)), *ob = st_[5].ensure(cnt);
    flatten_cts_to(ct1.data(), cnt, l1, e1 ? p_.Q : p_.q, a1, b1);
    flatten_cts_to(ct2.data(), cnt, l2, e2 ? p_.Q : p_.q, a2, b2);
    const double t1 = now_ms();
    const int in_ext = (e1 ? 1 : 0) | (e2 ? 2 : 0);
    run_split(cnt, [&](size_t d, size_t lo, size_t hi) {
        const int rc = luxb200_evalbingate_batch_ex(hs_[d], static_cast<int>(gate), a1 + lo * l1, b1 + lo, a2 + lo * l2, b2 + lo,
                                                    oa + lo * n, ob + lo, hi - lo, 0, pm, in_ext);
        if (rc != LUXB200_OK) throw std::runtime_error(std::string("BackendB200::EvalBinGate: ") + luxb200_last_error());
    });
    const double t2 = now_ms();
    out.resize(cnt);
    build_cts_from(oa, ob, cnt, n, NativeInteger(p_.q), NativeInteger(4), out.data());   // a fresh ciphertext: m_p = 4
    timing_ = Timing{t1 - t0, t2 - t1, now_ms() - t2};
}

void BackendB200::EvalBinGate(BINGATE gate, const std::vector<LWECiphertext>& ct1, const std::vector<LWECiphertext>& ct2,
                              std::vector<LWECiphertext>& out) {
    if (ct1.size() != ct2.size()) throw std::invalid_argument("Input size mismatch");
    if (ct1.empty()) { out.clear(); return; }
    std::lock_guard<std::recursive_mutex> lk(mu_);
    // per-element checks of BinFHEScheme::EvalBinGate (binfhe-base-scheme.cpp:81-86) and the property one device call
    // shares across its batch: the SMALL_DIM / extended format of either operand
    bool uniform = true;
    for (size_t i = 0; i < ct1.size(); ++i) {
        if (!ct1[i]) throw std::invalid_argument("Ciphertext1 is empty");
        if (!ct2[i]) throw std::invalid_argument("Ciphertext2 is empty");
        if (ct1[i] == ct2[i]) throw std::invalid_argument("Input ciphertexts should be independant");
        uniform = uniform && ct1[i]->GetModulus() == ct1[0]->GetModulus() && ct2[i]->GetModulus() == ct2[0]->GetModulus();
    }
    if (uniform) { gate2(gate, ct1, ct2, out); return; }
    // mixed batch: one device call per (format of ct1, format of ct2) group, results scattered back in order
    std::map<std::pair<uint64_t, uint64_t>, std::vector<size_t>> groups;
    for (size_t i = 0; i < ct1.size(); ++i) groups[{u64(ct1[i]->GetModulus()), u64(ct2[i]->GetModulus())}].push_back(i);
    out.assign(ct1.size(), nullptr);
    for (const auto& g : groups) {
        std::vector<LWECiphertext> x, y, r;
        for (size_t i : g.second) { x.push_back(ct1[i]); y.push_back(ct2[i]); }
        gate2(gate, x, y, r);
        for (size_t k = 0; k < g.second.size(); ++k) out[g.second[k]] = r[k];
    }
}

void BackendB200::EvalBinGateMulti(BINGATE gate, const std::vector<std::vector<LWECiphertext>>& cts,
                                   std::vector<LWECiphertext>& out) {
    if (cts.empty()) { out.clear(); return; }
    std::lock_guard<std::recursive_mutex> lk(mu_);
    const uint32_t k = static_cast<uint32_t>(cts[0].size()), n = p_.n;
    if (k == 0) throw std::invalid_argument("Input size mismatch");
    const NativeInteger pm = cts[0][0]->GetptModulus();
    std::vector<uint64_t> a(cts.size() * size_t(k) * n), b(cts.size() * size_t(k)), oa(cts.size() * size_t(n)), ob(cts.size());
    for (size_t i = 0; i < cts.size(); ++i) {
        if (cts[i].size() != k) throw std::invalid_argument("Input size mismatch");
        for (uint32_t x = 0; x < k; ++x)
            for (uint32_t y = x + 1; y < k; ++y)
                if (cts[i][x] == cts[i][y]) throw std::invalid_argument("Input ciphertexts should be independent");
        if (!cts[i][0] || cts[i][0]->GetptModulus() != pm)
            throw std::invalid_argument("BackendB200::EvalBinGateMulti: one batch = one plaintext modulus (split the batch)");
        // inputs mod q; an extended input (N, Q) is not accepted here: switch it down first (KeySwitchBatch + ModSwitchBatch)
        flatten_cts_to(cts[i].data(), k, n, p_.q, &a[i * k * n], &b[i * k]);
    }
    run_split(cts.size(), [&](size_t d, size_t lo, size_t hi) {
        const int rc = luxb200_evalbingate_multi_batch(hs_[d], static_cast<int>(gate), k, &a[lo * k * n], &b[lo * k], &oa[lo * n],
                                                       &ob[lo], hi - lo, u64(pm));
        if (rc != LUXB200_OK) throw std::runtime_error(std::string("BackendB200::EvalBinGateMulti: ") + luxb200_last_error());
    });
    // CMUX ends in a 2-input NAND (a fresh ciphertext, m_p = 4); the sum gates SetptModulus(p) (binfhe-base-scheme.cpp:171)
    build_cts(oa, ob, n, NativeInteger(p_.q), gate == CMUX ? NativeInteger(4) : pm, out);
}

void BackendB200::Bootstrap(const std::vector<LWECiphertext>& in, std::vector<LWECiphertext>& out) {
    if (in.empty()) { out.clear(); return; }
    std::lock_guard<std::recursive_mutex> lk(mu_);
    const uint32_t n = p_.n;
    for (const auto& ct : in) {
        if (!ct) throw std::invalid_argument("Ciphertext is empty");
        if (ct->GetptModulus() != in[0]->GetptModulus())
            throw std::invalid_argument("BackendB200::Bootstrap: one batch = one plaintext modulus (split the batch)");
    }
    uint64_t *a = st_[0].ensure(in.size() * size_t(n)), *b = st_[1].ensure(in.size()), *oa = st_[4].ensure(in.size() * size_t(n)),
             *ob = st_[5].ensure(in.size());
    flatten_cts_to(in.data(), in.size(), n, p_.q, a, b);
    const uint64_t pm = u64(in[0]->GetptModulus());
    run_split(in.size(), [&](size_t d, size_t lo, size_t hi) {
        const int rc = luxb200_bootstrap_batch(hs_[d], a + lo * n, b + lo, oa + lo * n, ob + lo, hi - lo, pm, 0);
        if (rc != LUXB200_OK) throw std::runtime_error(std::string("BackendB200::Bootstrap: ") + luxb200_last_error());
    });
    out.resize(in.size());
    build_cts_from(oa, ob, in.size(), n, NativeInteger(p_.q), in[0]->GetptModulus(), out.data());
}

void BackendB200::EvalCircuit(const std::vector<LWECiphertext>& inputs, const std::vector<luxb200_gate>& gates,
                              const std::vector<uint32_t>& outputs, const std::vector<std::vector<NativeInteger>>& luts,
                              const NativeInteger& beta, std::vector<LWECiphertext>& out) {
    if (outputs.empty()) { out.clear(); return; }
    std::lock_guard<std::recursive_mutex> lk(mu_);
    std::vector<uint64_t> a(inputs.size() * size_t(p_.n)), b(inputs.size()), oa(outputs.size() * size_t(p_.n)), ob(outputs.size());
    flatten_cts_to(inputs.data(), inputs.size(), p_.n, p_.q, a.data(), b.data());   // [n_inputs][count = 1][n]
    std::vector<uint64_t> lt(luts.size() * size_t(p_.q));
    for (size_t i = 0; i < luts.size(); ++i) {
        if (luts[i].size() != p_.q) throw std::invalid_argument("BackendB200::EvalCircuit: a look-up table must have q entries");
        for (size_t k = 0; k < luts[i].size(); ++k) lt[i * p_.q + k] = u64(luts[i][k]);
    }
    check(luxb200_circuit_batch_ex(h_, static_cast<uint32_t>(inputs.size()), gates.data(), static_cast<uint32_t>(gates.size()),
                                   outputs.data(), static_cast<uint32_t>(outputs.size()), lt.empty() ? nullptr : lt.data(),
                                   static_cast<uint32_t>(luts.size()), u64(beta), a.data(), b.data(), oa.data(), ob.data(), 1),
          "EvalCircuit");
    build_cts(oa, ob, p_.n, NativeInteger(p_.q), NativeInteger(4), out);
}

void BackendB200::EvalFunc(const std::vector<LWECiphertext>& in, const std::vector<NativeInteger>& lut, const NativeInteger& beta,
                           std::vector<LWECiphertext>& out) {
    if (in.empty()) { out.clear(); return; }
    std::lock_guard<std::recursive_mutex> lk(mu_);
    const uint32_t n = p_.n;
    if (!in[0]) throw std::invalid_argument("Ciphertext is empty");
    const uint64_t ctmod = u64(in[0]->GetModulus());
    std::vector<uint64_t> l(lut.size());
    for (size_t i = 0; i < lut.size(); ++i) l[i] = u64(lut[i]);
    uint64_t *a = st_[0].ensure(in.size() * size_t(n)), *b = st_[1].ensure(in.size()), *oa = st_[4].ensure(in.size() * size_t(n)),
             *ob = st_[5].ensure(in.size());
    flatten_cts_to(in.data(), in.size(), n, ctmod, a, b);   // every element: dimension n and the modulus of the LUT
    run_split(in.size(), [&](size_t d, size_t lo, size_t hi) {
        const int rc = luxb200_evalfunc_batch(hs_[d], a + lo * n, b + lo, l.data(), l.size(), u64(beta), oa + lo * n, ob + lo,
                                              hi - lo, ctmod);
        if (rc != LUXB200_OK) throw std::runtime_error(std::string("BackendB200::EvalFunc: ") + luxb200_last_error());
    });
    out.resize(in.size());
    build_cts_from(oa, ob, in.size(), n, NativeInteger(ctmod), NativeInteger(4), out.data());   // fresh ciphertexts: m_p = 4
}

// BinFHEContext::BTKeyGen with the keys drawn on the device (luxb200_btkeygen), then handed to the context as the
// reference's own key objects (BTKeyLoad, binfhecontext.h:273), so cc's CPU calls and the device use the same keys.
void BackendB200::BTKeyGen(BinFHEContext& cc, ConstLWEPrivateKey sk, uint64_t seed) {
    if (!sk) throw std::invalid_argument("PrivateKey is empty");
    std::lock_guard<std::recursive_mutex> lk(mu_);
    const luxb200_params p = params_of(cc);
    if (!h_ || std::memcmp(&p, &p_, sizeof(p)) != 0) Bind(cc);   // parameters only: the context holds no keys yet
    const NativeVector& sv = sk->GetElement();
    if (sv.GetLength() != p_.n) throw std::invalid_argument("BackendB200::BTKeyGen: secret key has the wrong dimension");
    const uint64_t mod = u64(sv.GetModulus());
    std::vector<int64_t> s(p_.n);
    for (uint32_t i = 0; i < p_.n; ++i) {   // residues mod q -> signed values (the reference stores -1 as q - 1)
        const uint64_t v = u64(sv[i]);
        s[i] = v > mod / 2 ? int64_t(v) - int64_t(mod) : int64_t(v);
    }
    check(luxb200_btkeygen(h_, s.data(), seed, nullptr), "BTKeyGen");
    share_keys();
    std::vector<uint64_t> flat(luxb200_bsk_len(h_)), A(luxb200_ksk_len(h_)), B(A.size() / p_.n);
    check(luxb200_export_keys(h_, flat.data(), A.data(), B.data()), "export_keys");
    RingGSWBTKey key;
    unflatten_bsk(flat, key.BSkey);
    const uint32_t digitsKS = static_cast<uint32_t>(B.size() / (size_t(p_.N) * p_.baseKS));
    std::vector<std::vector<std::vector<NativeVector>>> kA(p_.N);
    std::vector<std::vector<std::vector<NativeInteger>>> kB(p_.N);
    const NativeInteger qKS(p_.qKS);
#pragma omp parallel for schedule(static)
    for (uint32_t i = 0; i < p_.N; ++i) {
        kA[i].resize(p_.baseKS);
        kB[i].resize(p_.baseKS);
        for (uint32_t j = 0; j < p_.baseKS; ++j)
            for (uint32_t k = 0; k < digitsKS; ++k) {
                const size_t row = (size_t(i) * p_.baseKS + j) * digitsKS + k;
                NativeVector v(p_.n, qKS);
                std::memcpy(&v[0], &A[row * p_.n], size_t(p_.n) * 8);
                kA[i][j].push_back(std::move(v));
                kB[i][j].push_back(NativeInteger(B[row]));
            }
    }
    key.KSkey = std::make_shared<LWESwitchingKeyImpl>(std::move(kA), std::move(kB));
    cc.BTKeyLoad(key);
    bound_bsk_ = key.BSkey;   // these objects ARE what is resident: Attach(cc) will not reload them
    bound_ksk_ = key.KSkey;
    packed_bsk_ = packed_ksk_ = false;
}

BackendB200* RegisterB200(BinFHEContext& cc, const std::vector<int>& devices) {
    auto be = std::make_unique<BackendB200>(devices);
    if (!be->IsAvailable())
        throw std::runtime_error("RegisterB200: CUDA device " + std::to_string(be->device()) +
                                 " (or another requested one) is not present (there is no CPU fallback)");
    be->Bind(cc);
    BackendB200* raw = be.get();
    auto& reg = BackendRegistry::Instance();
    reg.Register(std::move(be));
    reg.SetDefault(BackendType::CUDA);
    return raw;
}
BackendB200* RegisterB200(BinFHEContext& cc, int device) { return RegisterB200(cc, std::vector<int>{device}); }

}  // namespace backend
}  // namespace lux::fhe
