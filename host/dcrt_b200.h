// dcrt_b200.h — the DCRTPoly NTT seam on the B200: DCRTPolyImpl::SwitchFormat
// (src/core/include/lattice/hal/default/dcrtpoly-impl.h:1953-1959 — an OpenMP loop of PolyImpl::SwitchFormat over the RNS
// limbs, poly-impl.h:420-440) as ONE luxb200_ntt_batch call over every limb of every polynomial handed in.
//
// Host-side glue written against the reference's own types (DCRTPoly, ILDCRTParams); no arithmetic here.  The plan (twiddle
// tables of every limb modulus, resident on the device) is cached per parameter object, so repeated calls on polynomials of
// one CryptoContext upload nothing but the coefficients.
#pragma once
#include <luxb200.h>

#include <map>
#include <memory>
#include <mutex>
#include <vector>

#include "lattice/lat-hal.h"

namespace lux::fhe {
namespace backend {

class DCRTNttB200 {
public:
    explicit DCRTNttB200(int device = 0);
    ~DCRTNttB200();
    DCRTNttB200(const DCRTNttB200&) = delete;
    DCRTNttB200& operator=(const DCRTNttB200&) = delete;

    // poly.SwitchFormat(): COEFFICIENT -> EVALUATION (forward NTT of every limb) or back; in place, bit-identical.
    void SwitchFormat(DCRTPoly& poly);
    // The same for a batch of polynomials that share parameters and format (one device call for all limbs of all of them).
    void SwitchFormat(std::vector<DCRTPoly*>& polys);
    // poly.SetFormat(f): no-op when already in format f
    void SetFormat(DCRTPoly& poly, Format f) {
        if (poly.GetFormat() != f) SwitchFormat(poly);
    }

private:
    luxb200_ntt_plan* plan_for(const DCRTPoly& poly);
    int device_;
    std::mutex mu_;
    std::map<std::vector<uint64_t>, luxb200_ntt_plan*> plans_;   // key: N, then the limb moduli
    std::vector<uint64_t> stage_;
};

}  // namespace backend
}  // namespace lux::fhe
