// backend_b200.cpp — see backend_b200.h.  Pack -> one luxb200_* call -> unpack; no arithmetic here.
#include "backend_b200.h"
#include "packed_b200.h"

#include <cstring>
#include <stdexcept>

namespace lux::fhe {
namespace backend {

namespace {

inline uint64_t u64(const NativeInteger& v) { return v.ConvertToInt<uint64_t>(); }

// [count][len] a-parts and [count] b-parts of a batch of LWE ciphertexts (NativeVector storage is uint64 residues)
void flatten_cts(const std::vector<LWECiphertext>& cts, uint32_t len, std::vector<uint64_t>& a, std::vector<uint64_t>& b) {
    a.resize(cts.size() * size_t(len));
    b.resize(cts.size());
    for (size_t i = 0; i < cts.size(); ++i) {
        if (!cts[i] || cts[i]->GetLength() != len)
            throw std::invalid_argument("BackendB200: ciphertext " + std::to_string(i) + " has the wrong dimension");
        const NativeVector& v = cts[i]->GetA();
        for (uint32_t k = 0; k < len; ++k) a[i * len + k] = u64(v[k]);
        b[i] = u64(cts[i]->GetB());
    }
}

void build_cts(const std::vector<uint64_t>& a, const std::vector<uint64_t>& b, uint32_t len, const NativeInteger& mod,
               const NativeInteger& pmod, std::vector<LWECiphertext>& out) {
    out.resize(b.size());
    for (size_t i = 0; i < b.size(); ++i) {
        NativeVector v(len, mod);
        for (uint32_t k = 0; k < len; ++k) v[k] = NativeInteger(a[i * len + k]);
        out[i] = std::make_shared<LWECiphertextImpl>(std::move(v), NativeInteger(b[i]));
        out[i]->SetptModulus(pmod);
    }
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

BackendB200::BackendB200(int device) : device_(device) {}

BackendB200::~BackendB200() {
    if (h_) luxb200_ctx_destroy(h_);
}

void BackendB200::check(int rc, const char* what) const {
    if (rc != LUXB200_OK) throw std::runtime_error(std::string("BackendB200::") + what + ": " + luxb200_last_error());
}

bool BackendB200::IsAvailable() const { return luxb200_device_count() > device_; }

size_t BackendB200::DeviceMemory() const {
    uint64_t f = 0, t = 0;
    return luxb200_dev_mem_info(device_, &f, &t) == LUXB200_OK ? size_t(t) : 0;
}

void BackendB200::Bind(BinFHEContext& cc) {
    auto lwe = cc.GetParams()->GetLWEParams();
    auto rg  = cc.GetParams()->GetRingGSWParams();
    luxb200_params p{};
    p.n = lwe->Getn(); p.N = lwe->GetN();
    p.q = u64(lwe->Getq()); p.Q = u64(lwe->GetQ()); p.qKS = u64(lwe->GetqKS());
    p.baseKS = u64(lwe->GetBaseKS()); p.baseG = rg->GetBaseG(); p.baseR = rg->GetBaseR();
    p.numAutoKeys = rg->GetNumAutoKeys(); p.method = static_cast<uint32_t>(rg->GetMethod());
    p.psi = u64(rg->GetPolyParams()->GetRootOfUnity());
    if (h_) { luxb200_ctx_destroy(h_); h_ = nullptr; }
    check(luxb200_ctx_create(&p, device_, &h_), "Bind");
    p_ = p;
    digitsG_ = rg->GetDigitsG(); digitsR_ = static_cast<uint32_t>(rg->GetDigitsR().size());
    rows_ = 2 * (digitsG_ - 1);
    poly_params_ = rg->GetPolyParams();
    bound_bsk_ = bound_ksk_ = nullptr;
    if (cc.GetRefreshKey() && cc.GetSwitchKey()) load_keys(cc.GetRefreshKey(), cc.GetSwitchKey());
}

// RingGSWACCKeyImpl -> canonical flat layout of include/luxb200.h (the orders of rgsw-acc-cggi.cpp:39-57,
// rgsw-acc-dm.cpp:39-60, rgsw-acc-lmkcdey.cpp:39-68)
void BackendB200::flatten_bsk(const RingGSWACCKeyImpl& ek, std::vector<uint64_t>& out) const {
    out.assign(luxb200_bsk_len(h_), 0);
    const uint32_t N = p_.N, n = p_.n;
    size_t pos = 0;
    auto put = [&](const RingGSWEvalKey& k) {
        if (!k) throw std::invalid_argument("BackendB200: refresh key has an empty entry where one is required");
        for (const auto& row : k->GetElements())
            for (const auto& poly : row)
                for (uint32_t t = 0; t < N; ++t) out[pos++] = u64(poly[t]);
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
}

void BackendB200::ensure_bsk(const RingGSWACCKey& ek) {
    if (!ek) throw std::invalid_argument("Bootstrapping keys have not been generated. Please call BTKeyGen before calling bootstrapping.");
    if (ek.get() == bound_bsk_) return;
    std::vector<uint64_t> flat;
    flatten_bsk(*ek, flat);
    check(luxb200_load_bsk(h_, flat.data(), 0), "load_bsk");
    bound_bsk_ = ek.get();
}

void BackendB200::ensure_ksk(const LWESwitchingKey& ks) {
    if (!ks) throw std::invalid_argument("BackendB200: switching key is empty");
    if (ks.get() == bound_ksk_) return;
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
                for (uint32_t t = 0; t < v.GetLength(); ++t) A[pa++] = u64(v[t]);
                B[pb++] = u64(kB[i][j][k]);
            }
    check(luxb200_load_ksk(h_, A.data(), B.data(), 0), "load_ksk");
    bound_ksk_ = ks.get();
}

void BackendB200::load_keys(const RingGSWACCKey& ek, const LWESwitchingKey& ks) {
    ensure_bsk(ek);
    ensure_ksk(ks);
}

// ---- memory (backend.h:88-113) ----------------------------------------------------------------------------
DeviceBuffer BackendB200::Allocate(size_t bytes) {
    DeviceBuffer b;
    void* p = nullptr;
    check(luxb200_dev_alloc(device_, bytes, &p), "Allocate");
    b.ptr = p; b.size = bytes; b.device = BackendType::CUDA;
    return b;
}
void BackendB200::Free(DeviceBuffer& buffer) {
    if (buffer.ptr) check(luxb200_dev_free(device_, buffer.ptr), "Free");
    buffer.ptr = nullptr; buffer.size = 0;
}
void BackendB200::CopyToDevice(const void* host_ptr, DeviceBuffer& device_buffer, size_t bytes) {
    if (bytes > device_buffer.size) throw std::invalid_argument("BackendB200::CopyToDevice: buffer too small");
    check(luxb200_dev_copy(device_, device_buffer.ptr, host_ptr, bytes, 1), "CopyToDevice");
}
void BackendB200::CopyToHost(const DeviceBuffer& device_buffer, void* host_ptr, size_t bytes) {
    if (bytes > device_buffer.size) throw std::invalid_argument("BackendB200::CopyToHost: buffer too small");
    check(luxb200_dev_copy(device_, host_ptr, device_buffer.ptr, bytes, 2), "CopyToHost");
}
void BackendB200::Synchronize() {
    if (h_) check(luxb200_synchronize(h_), "Synchronize");
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
        bound_bsk_ = bsk;
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
        bound_ksk_ = ksk;
    }
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
void BackendB200::EvalBinGate(BINGATE gate, const std::vector<LWECiphertext>& ct1, const std::vector<LWECiphertext>& ct2,
                              std::vector<LWECiphertext>& out) {
    if (ct1.size() != ct2.size()) throw std::invalid_argument("Input size mismatch");
    if (ct1.empty()) { out.clear(); return; }
    std::vector<uint64_t> a1, b1, a2, b2, oa(ct1.size() * size_t(p_.n)), ob(ct1.size());
    flatten_cts(ct1, p_.n, a1, b1);
    flatten_cts(ct2, p_.n, a2, b2);
    check(luxb200_evalbingate_batch(h_, static_cast<int>(gate), a1.data(), b1.data(), a2.data(), b2.data(), oa.data(),
                                    ob.data(), ct1.size(), 0), "EvalBinGate");
    build_cts(oa, ob, p_.n, NativeInteger(p_.q), NativeInteger(4), out);
}
void BackendB200::EvalBinGateMulti(BINGATE gate, const std::vector<std::vector<LWECiphertext>>& cts,
                                   std::vector<LWECiphertext>& out) {
    if (cts.empty()) { out.clear(); return; }
    const uint32_t k = static_cast<uint32_t>(cts[0].size()), n = p_.n;
    std::vector<uint64_t> a(cts.size() * size_t(k) * n), b(cts.size() * size_t(k)), oa(cts.size() * size_t(n)), ob(cts.size());
    for (size_t i = 0; i < cts.size(); ++i) {
        if (cts[i].size() != k) throw std::invalid_argument("Input size mismatch");
        std::vector<uint64_t> ai, bi;
        flatten_cts(cts[i], n, ai, bi);
        std::memcpy(&a[i * k * n], ai.data(), ai.size() * 8);
        std::memcpy(&b[i * k], bi.data(), bi.size() * 8);
    }
    check(luxb200_evalbingate_multi_batch(h_, static_cast<int>(gate), k, a.data(), b.data(), oa.data(), ob.data(), cts.size(),
                                          u64(cts[0][0]->GetptModulus())), "EvalBinGateMulti");
    build_cts(oa, ob, n, NativeInteger(p_.q), cts[0][0]->GetptModulus(), out);
}
void BackendB200::Bootstrap(const std::vector<LWECiphertext>& in, std::vector<LWECiphertext>& out) {
    if (in.empty()) { out.clear(); return; }
    std::vector<uint64_t> a, b, oa(in.size() * size_t(p_.n)), ob(in.size());
    flatten_cts(in, p_.n, a, b);
    check(luxb200_bootstrap_batch(h_, a.data(), b.data(), oa.data(), ob.data(), in.size(), u64(in[0]->GetptModulus()), 0),
          "Bootstrap");
    build_cts(oa, ob, p_.n, NativeInteger(p_.q), in[0]->GetptModulus(), out);
}
void BackendB200::EvalCircuit(const std::vector<LWECiphertext>& inputs, const std::vector<luxb200_gate>& gates,
                              const std::vector<uint32_t>& outputs, std::vector<LWECiphertext>& out) {
    if (outputs.empty()) { out.clear(); return; }
    std::vector<uint64_t> a, b, oa(outputs.size() * size_t(p_.n)), ob(outputs.size());
    flatten_cts(inputs, p_.n, a, b);   // [n_inputs][count = 1][n]
    check(luxb200_circuit_batch(h_, static_cast<uint32_t>(inputs.size()), gates.data(), static_cast<uint32_t>(gates.size()),
                                outputs.data(), static_cast<uint32_t>(outputs.size()), a.data(), b.data(), oa.data(), ob.data(),
                                1), "EvalCircuit");
    build_cts(oa, ob, p_.n, NativeInteger(p_.q), NativeInteger(4), out);
}
void BackendB200::EvalFunc(const std::vector<LWECiphertext>& in, const std::vector<NativeInteger>& lut, const NativeInteger& beta,
                           std::vector<LWECiphertext>& out) {
    if (in.empty()) { out.clear(); return; }
    std::vector<uint64_t> a, b, l(lut.size()), oa(in.size() * size_t(p_.n)), ob(in.size());
    flatten_cts(in, p_.n, a, b);
    for (size_t i = 0; i < lut.size(); ++i) l[i] = u64(lut[i]);
    const uint64_t ctmod = u64(in[0]->GetModulus());
    check(luxb200_evalfunc_batch(h_, a.data(), b.data(), l.data(), l.size(), u64(beta), oa.data(), ob.data(), in.size(), ctmod),
          "EvalFunc");
    build_cts(oa, ob, p_.n, NativeInteger(ctmod), in[0]->GetptModulus(), out);
}

BackendB200* RegisterB200(BinFHEContext& cc, int device) {
    auto be = std::make_unique<BackendB200>(device);
    if (!be->IsAvailable()) throw std::runtime_error("RegisterB200: no CUDA device " + std::to_string(device) + " (there is no CPU fallback)");
    be->Bind(cc);
    BackendB200* raw = be.get();
    auto& reg = BackendRegistry::Instance();
    reg.Register(std::move(be));
    reg.SetDefault(BackendType::CUDA);
    return raw;
}

}  // namespace backend
}  // namespace lux::fhe
