// packed_b200.h — layout records and zero-copy views of the packed key wire formats implemented in packed_b200.cpp
// (the functions themselves are declared by the reference's backend/packed.h:244-274).
#pragma once
#include <cstddef>
#include <cstdint>

#include "backend/packed.h"

namespace lux::fhe {
namespace backend {

// follows PackedBootstrappingKey on the wire
struct PackedBSKLayout {
    uint32_t dims[3];   // RingGSWACCKeyImpl cube (rgsw-acckey.h:67)
    uint32_t method;    // BINFHE_METHOD
    uint64_t Q;
};
static_assert(sizeof(PackedBSKLayout) == 24, "PackedBSKLayout must be 24 bytes");

struct PackedBSKView {
    const PackedBootstrappingKey* head = nullptr;
    const PackedBSKLayout* layout = nullptr;
    const uint8_t* rows = nullptr;     // rows per [i][j][k] slot, 0 = empty
    const uint8_t* coeffs = nullptr;   // present entries in [i][j][k] order
    uint32_t coeff_bytes = 0;
    size_t slots = 0;
    uint64_t coeff(size_t idx) const {
        return coeff_bytes == 4 ? reinterpret_cast<const uint32_t*>(coeffs)[idx] : reinterpret_cast<const uint64_t*>(coeffs)[idx];
    }
};

struct PackedKSKView {
    const PackedSwitchingKey* head = nullptr;
    const uint8_t* A = nullptr;   // [N][baseKS][digits][n]
    const uint8_t* B = nullptr;   // [N][baseKS][digits]
    uint32_t coeff_bytes = 0, base = 0;
    size_t vecs = 0;
    uint64_t coeff(const uint8_t* p, size_t idx) const {
        return coeff_bytes == 2 ? reinterpret_cast<const uint16_t*>(p)[idx]
             : coeff_bytes == 4 ? reinterpret_cast<const uint32_t*>(p)[idx] : reinterpret_cast<const uint64_t*>(p)[idx];
    }
};

// Structural validation (magic, version, type, sizes add up); false = malformed.
bool ParsePackedBSK(const uint8_t* data, size_t size, PackedBSKView& view);
bool ParsePackedKSK(const uint8_t* data, size_t size, PackedKSKView& view);

}  // namespace backend
}  // namespace lux::fhe
