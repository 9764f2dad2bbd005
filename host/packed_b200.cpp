// packed_b200.cpp — the reference's packed wire formats (src/binfhe/include/backend/packed.h) with the key functions
// filled in.  The reference's packed.cpp implements the LWE formats and the header checks, and leaves
// PackBootstrappingKey / UnpackBootstrappingKey / PackSwitchingKey / UnpackSwitchingKey as empty stubs
// (packed.cpp:284-331).  This translation unit compiles the reference's packed.cpp from where it lies (so PackLWE,
// PackLWEBatch, ValidatePackedHeader, Estimate* are the reference's own code, byte for byte) with its four stubs
// renamed out of the way, and defines the four key functions for real.
//
// Wire format of a packed bootstrapping key (all little-endian, 8-byte aligned):
//   PackedBootstrappingKey                      header.type = BOOTSTRAPPING_KEY, header.flags = key_layout,
//                                               key_layout = flags | KEY_LAYOUT_NTT (keys are EVALUATION-format),
//                                               reserved = bytes per coefficient (4 when Q < 2^32, else 8)
//   uint32 dims[3], method                      the RingGSWACCKeyImpl cube (rgsw-acckey.h:67) and BINFHE_METHOD
//   uint64 Q
//   uint8  rows[dims0*dims1*dims2] (+pad to 8)  rows of entry [i][j][k]; 0 = empty slot (AP digit 0, unused LMKCDEY slots)
//   coefficients                                entries in [i][j][k] order, each rows x 2 polynomials x N coefficients
// and of a packed switching key:
//   PackedSwitchingKey                          input_n = N, output_n = n, decomp_levels = digits, decomp_base_log,
//                                               Q = qKS, reserved[0] = bytes per coefficient (2, 4 or 8), reserved[1] = baseKS
//   A coefficients [N][baseKS][digits][n], then B [N][baseKS][digits]   (order of lwe-pke.cpp:264-344)
// The coefficient areas are exactly the canonical flat orders of include/luxb200.h up to the GINX/LMKCDEY entry
// interleave, so BackendB200::LoadPacked() feeds luxb200_load_bsk / luxb200_load_ksk from the bytes without rebuilding the
// shared_ptr graph.
#define PackBootstrappingKey PackBootstrappingKey_reference_stub
#define UnpackBootstrappingKey UnpackBootstrappingKey_reference_stub
#define PackSwitchingKey PackSwitchingKey_reference_stub
#define UnpackSwitchingKey UnpackSwitchingKey_reference_stub
#include "backend/packed.cpp"   // NOLINT: the reference's own source (src/binfhe/lib/backend/packed.cpp), unmodified
#undef PackBootstrappingKey
#undef UnpackBootstrappingKey
#undef PackSwitchingKey
#undef UnpackSwitchingKey

#include "packed_b200.h"

namespace lux::fhe {
namespace backend {

namespace {

inline uint32_t ilog2(uint64_t v) { uint32_t r = 0; while (v > 1) { v >>= 1; ++r; } return r; }
inline size_t pad8(size_t v) { return (v + 7) & ~size_t(7); }

template <typename T>
inline void put_coeffs(uint8_t*& p, const NativePoly& poly, uint32_t N) {
    T* d = reinterpret_cast<T*>(p);
    for (uint32_t t = 0; t < N; ++t) d[t] = static_cast<T>(poly[t].ConvertToInt<uint64_t>());
    p += size_t(N) * sizeof(T);
}

}  // namespace

std::vector<uint8_t> PackBootstrappingKey(const RingGSWACCKey& ek, const std::shared_ptr<RingGSWCryptoParams>& params,
                                          uint32_t flags) {
    if (!ek || !params) return {};
    const auto& cube = ek->GetElements();
    const uint32_t d0 = cube.size(), d1 = d0 ? cube[0].size() : 0, d2 = d1 ? cube[0][0].size() : 0;
    const uint32_t N = params->GetN();
    const uint64_t Q = params->GetQ().ConvertToInt<uint64_t>();
    const uint32_t cb = Q < (uint64_t(1) << 32) ? 4 : 8;
    const size_t slots = size_t(d0) * d1 * d2;
    size_t coeffs = 0;
    std::vector<uint8_t> rows(slots, 0);
    size_t s = 0;
    for (const auto& a : cube)
        for (const auto& b : a) {
            if (a.size() != d1 || b.size() != d2) throw std::invalid_argument("PackBootstrappingKey: ragged key cube");
            for (const auto& e : b) {
                if (e) {
                    const auto& el = e->GetElements();
                    if (el.empty() || el.size() > 255 || el[0].size() != 2) throw std::invalid_argument("PackBootstrappingKey: unexpected RGSW shape");
                    rows[s] = static_cast<uint8_t>(el.size());
                    coeffs += el.size() * 2 * size_t(N);
                }
                ++s;
            }
        }
    const size_t off_tab = sizeof(PackedBootstrappingKey) + sizeof(PackedBSKLayout);
    const size_t off_dat = off_tab + pad8(slots);
    const size_t total = off_dat + pad8(coeffs * cb);
    std::vector<uint8_t> out(total, 0);
    auto* h = reinterpret_cast<PackedBootstrappingKey*>(out.data());
    h->header = PackedHeader{PACKED_MAGIC, PACKED_VERSION, PackedType::BOOTSTRAPPING_KEY, total, slots, flags | KEY_LAYOUT_NTT, 0};
    uint32_t n = 0;   // LWE dimension = number of secret-key coefficients the key encrypts
    switch (params->GetMethod()) {
        case BINFHE_METHOD::AP: n = d0; break;
        case BINFHE_METHOD::GINX: n = d2; break;
        default: for (uint32_t i = 0; i < d2; ++i) n += cube[0][0][i] != nullptr;
    }
    h->lwe_n = n;
    h->lwe_log_q = ilog2(params->Getq().ConvertToInt<uint64_t>());
    h->rlwe_N = N;
    h->rlwe_num_limbs = 1;
    h->decomp_levels = params->GetDigitsG() - 1;
    h->decomp_base_log = ilog2(params->GetBaseG());
    h->key_size = total - sizeof(PackedBootstrappingKey);
    h->key_layout = flags | KEY_LAYOUT_NTT;
    h->reserved = cb;
    auto* L = reinterpret_cast<PackedBSKLayout*>(out.data() + sizeof(PackedBootstrappingKey));
    *L = PackedBSKLayout{{d0, d1, d2}, static_cast<uint32_t>(params->GetMethod()), Q};
    std::memcpy(out.data() + off_tab, rows.data(), slots);
    uint8_t* p = out.data() + off_dat;
    for (const auto& a : cube)
        for (const auto& b : a)
            for (const auto& e : b) {
                if (!e) continue;
                for (const auto& row : e->GetElements())
                    for (const auto& poly : row) {
                        if (poly.GetLength() != N || poly.GetFormat() != Format::EVALUATION)
                            throw std::invalid_argument("PackBootstrappingKey: key polynomial is not an EVALUATION-format ring element of the parameter set");
                        if (cb == 4) put_coeffs<uint32_t>(p, poly, N); else put_coeffs<uint64_t>(p, poly, N);
                    }
            }
    return out;
}

bool ParsePackedBSK(const uint8_t* data, size_t size, PackedBSKView& v) {
    if (!ValidatePackedHeader(data, size) || size < sizeof(PackedBootstrappingKey) + sizeof(PackedBSKLayout)) return false;
    const auto* h = reinterpret_cast<const PackedBootstrappingKey*>(data);
    if (h->header.type != PackedType::BOOTSTRAPPING_KEY || !(h->key_layout & KEY_LAYOUT_NTT)) return false;
    if (h->reserved != 4 && h->reserved != 8) return false;
    const auto* L = reinterpret_cast<const PackedBSKLayout*>(data + sizeof(PackedBootstrappingKey));
    const size_t slots = size_t(L->dims[0]) * L->dims[1] * L->dims[2];
    if (slots != h->header.element_count) return false;
    const size_t off_tab = sizeof(PackedBootstrappingKey) + sizeof(PackedBSKLayout);
    const size_t off_dat = off_tab + pad8(slots);
    if (off_dat > h->header.total_size) return false;
    size_t coeffs = 0;
    for (size_t s = 0; s < slots; ++s) coeffs += size_t(data[off_tab + s]) * 2 * h->rlwe_N;
    if (off_dat + pad8(coeffs * h->reserved) != h->header.total_size) return false;
    v.head = h; v.layout = L; v.rows = data + off_tab; v.coeffs = data + off_dat; v.coeff_bytes = h->reserved; v.slots = slots;
    return true;
}

RingGSWACCKey UnpackBootstrappingKey(const uint8_t* data, size_t size, const std::shared_ptr<RingGSWCryptoParams>& params) {
    PackedBSKView v;
    if (!params || !ParsePackedBSK(data, size, v)) throw std::invalid_argument("Invalid packed bootstrapping key");
    const uint32_t N = params->GetN();
    if (v.head->rlwe_N != N || v.layout->Q != params->GetQ().ConvertToInt<uint64_t>() ||
        v.layout->method != static_cast<uint32_t>(params->GetMethod()))
        throw std::invalid_argument("Packed bootstrapping key does not match the parameter set");
    auto ek = std::make_shared<RingGSWACCKeyImpl>(v.layout->dims[0], v.layout->dims[1], v.layout->dims[2]);
    const uint8_t* p = v.coeffs;
    size_t s = 0;
    for (uint32_t i = 0; i < v.layout->dims[0]; ++i)
        for (uint32_t j = 0; j < v.layout->dims[1]; ++j)
            for (uint32_t k = 0; k < v.layout->dims[2]; ++k, ++s) {
                const uint32_t rows = v.rows[s];
                if (!rows) continue;
                auto e = std::make_shared<RingGSWEvalKeyImpl>(rows, 2);
                for (uint32_t r = 0; r < rows; ++r)
                    for (uint32_t c = 0; c < 2; ++c) {
                        NativePoly poly(params->GetPolyParams(), Format::EVALUATION, true);
                        if (v.coeff_bytes == 4) {
                            const uint32_t* d = reinterpret_cast<const uint32_t*>(p);
                            for (uint32_t t = 0; t < N; ++t) poly[t] = NativeInteger(uint64_t(d[t]));
                        } else {
                            const uint64_t* d = reinterpret_cast<const uint64_t*>(p);
                            for (uint32_t t = 0; t < N; ++t) poly[t] = NativeInteger(d[t]);
                        }
                        p += size_t(N) * v.coeff_bytes;
                        (*e)[r][c] = std::move(poly);
                    }
                (*ek)[i][j][k] = std::move(e);
            }
    return ek;
}

std::vector<uint8_t> PackSwitchingKey(const LWESwitchingKey& ks, const std::shared_ptr<LWECryptoParams>& params) {
    if (!ks || !params) return {};
    const auto& A = ks->GetElementsA();
    const auto& B = ks->GetElementsB();
    const uint32_t N = A.size(), base = N ? A[0].size() : 0, digits = base ? A[0][0].size() : 0;
    const uint32_t n = digits ? A[0][0][0].GetLength() : 0;
    const uint64_t qKS = params->GetqKS().ConvertToInt<uint64_t>();
    const uint32_t cb = qKS <= (uint64_t(1) << 16) ? 2 : qKS <= (uint64_t(1) << 32) ? 4 : 8;
    const size_t vecs = size_t(N) * base * digits;
    const size_t offA = sizeof(PackedSwitchingKey), offB = offA + pad8(vecs * n * cb), total = offB + pad8(vecs * cb);
    std::vector<uint8_t> out(total, 0);
    auto* h = reinterpret_cast<PackedSwitchingKey*>(out.data());
    h->header = PackedHeader{PACKED_MAGIC, PACKED_VERSION, PackedType::SWITCHING_KEY, total, vecs, 0, 0};
    h->input_n = N; h->output_n = n; h->decomp_levels = digits;
    h->decomp_base_log = ilog2(uint64_t(params->GetBaseKS()));
    h->Q = qKS; h->reserved[0] = cb; h->reserved[1] = base;
    auto put = [cb](uint8_t* p, size_t idx, uint64_t v) {
        if (cb == 2) reinterpret_cast<uint16_t*>(p)[idx] = uint16_t(v);
        else if (cb == 4) reinterpret_cast<uint32_t*>(p)[idx] = uint32_t(v);
        else reinterpret_cast<uint64_t*>(p)[idx] = v;
    };
    size_t s = 0;
    for (uint32_t i = 0; i < N; ++i) {
        if (A[i].size() != base || B[i].size() != base) throw std::invalid_argument("PackSwitchingKey: ragged key");
        for (uint32_t j = 0; j < base; ++j) {
            if (A[i][j].size() != digits || B[i][j].size() != digits) throw std::invalid_argument("PackSwitchingKey: ragged key");
            for (uint32_t k = 0; k < digits; ++k, ++s) {
                const NativeVector& a = A[i][j][k];
                if (a.GetLength() != n) throw std::invalid_argument("PackSwitchingKey: ragged key");
                for (uint32_t t = 0; t < n; ++t) put(out.data() + offA, s * n + t, a[t].ConvertToInt<uint64_t>());
                put(out.data() + offB, s, B[i][j][k].ConvertToInt<uint64_t>());
            }
        }
    }
    return out;
}

bool ParsePackedKSK(const uint8_t* data, size_t size, PackedKSKView& v) {
    if (!ValidatePackedHeader(data, size) || size < sizeof(PackedSwitchingKey)) return false;
    const auto* h = reinterpret_cast<const PackedSwitchingKey*>(data);
    if (h->header.type != PackedType::SWITCHING_KEY) return false;
    const uint32_t cb = h->reserved[0], base = h->reserved[1];
    if ((cb != 2 && cb != 4 && cb != 8) || base == 0) return false;   // baseKS need not be a power of two (TOY: 25); decomp_base_log is floor(log2)
    const size_t vecs = size_t(h->input_n) * base * h->decomp_levels;
    if (vecs != h->header.element_count) return false;
    const size_t offA = sizeof(PackedSwitchingKey), offB = offA + pad8(vecs * h->output_n * cb);
    if (offB + pad8(vecs * cb) != h->header.total_size) return false;
    v.head = h; v.A = data + offA; v.B = data + offB; v.coeff_bytes = cb; v.base = base; v.vecs = vecs;
    return true;
}

LWESwitchingKey UnpackSwitchingKey(const uint8_t* data, size_t size, const std::shared_ptr<LWECryptoParams>& params) {
    PackedKSKView v;
    if (!params || !ParsePackedKSK(data, size, v)) throw std::invalid_argument("Invalid packed switching key");
    const auto* h = v.head;
    if (h->input_n != params->GetN() || h->output_n != params->Getn() || h->Q != params->GetqKS().ConvertToInt<uint64_t>() ||
        v.base != uint64_t(params->GetBaseKS()))
        throw std::invalid_argument("Packed switching key does not match the parameter set");
    const NativeInteger qKS = params->GetqKS();
    std::vector<std::vector<std::vector<NativeVector>>> A(h->input_n);
    std::vector<std::vector<std::vector<NativeInteger>>> B(h->input_n);
    size_t s = 0;
    for (uint32_t i = 0; i < h->input_n; ++i) {
        A[i].resize(v.base); B[i].resize(v.base);
        for (uint32_t j = 0; j < v.base; ++j) {
            A[i][j].reserve(h->decomp_levels); B[i][j].resize(h->decomp_levels);
            for (uint32_t k = 0; k < h->decomp_levels; ++k, ++s) {
                NativeVector a(h->output_n, qKS);
                for (uint32_t t = 0; t < h->output_n; ++t) a[t] = NativeInteger(v.coeff(v.A, s * h->output_n + t));
                A[i][j].push_back(std::move(a));
                B[i][j][k] = NativeInteger(v.coeff(v.B, s));
            }
        }
    }
    return std::make_shared<LWESwitchingKeyImpl>(std::move(A), std::move(B));
}

}  // namespace backend
}  // namespace lux::fhe
