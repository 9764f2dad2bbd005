// dcrt_b200.cpp — see dcrt_b200.h.  Pack limbs -> luxb200_ntt_batch -> unpack; no arithmetic here.
#include "dcrt_b200.h"

#include <cstring>
#include <stdexcept>
#include <string>

namespace lux::fhe {
namespace backend {

DCRTNttB200::DCRTNttB200(int device) : device_(device) {}

DCRTNttB200::~DCRTNttB200() {
    for (auto& kv : plans_) luxb200_ntt_plan_destroy(kv.second);
}

luxb200_ntt_plan* DCRTNttB200::plan_for(const DCRTPoly& poly) {
    const auto& limbs = poly.GetParams()->GetParams();
    const uint32_t N  = poly.GetRingDimension();
    std::vector<uint64_t> key{N}, q, psi;
    for (const auto& lp : limbs) {
        q.push_back(lp->GetModulus().ConvertToInt<uint64_t>());
        psi.push_back(lp->GetRootOfUnity().ConvertToInt<uint64_t>());   // the reference's own root: tables are identical
        key.push_back(q.back());
    }
    auto it = plans_.find(key);
    if (it != plans_.end()) return it->second;
    luxb200_ntt_plan* pl = nullptr;
    if (luxb200_ntt_plan_create(N, static_cast<uint32_t>(q.size()), q.data(), psi.data(), device_, &pl) != LUXB200_OK)
        throw std::runtime_error(std::string("DCRTNttB200: ") + luxb200_last_error());
    plans_[key] = pl;
    return pl;
}

void DCRTNttB200::SwitchFormat(std::vector<DCRTPoly*>& polys) {
    if (polys.empty()) return;
    std::lock_guard<std::mutex> lk(mu_);
    const DCRTPoly& first = *polys[0];
    const Format from     = first.GetFormat();
    const size_t L = first.GetNumOfElements(), N = first.GetRingDimension();
    for (const DCRTPoly* p : polys)
        if (!p || p->GetFormat() != from || p->GetNumOfElements() != L || p->GetRingDimension() != N ||
            p->GetParams() != first.GetParams())
            throw std::invalid_argument("DCRTNttB200::SwitchFormat: polynomials of one batch must share parameters and format");
    luxb200_ntt_plan* pl = plan_for(first);
    stage_.resize(polys.size() * L * N);
#pragma omp parallel for collapse(2)
    for (size_t i = 0; i < polys.size(); ++i)
        for (size_t l = 0; l < L; ++l)
            std::memcpy(&stage_[(i * L + l) * N], &polys[i]->GetAllElements()[l].GetValues()[0], N * 8);
    // inverse = 1 when the data is in EVALUATION format (poly-impl.h:420-440)
    if (luxb200_ntt_batch(pl, stage_.data(), polys.size(), from == Format::EVALUATION ? 1 : 0) != LUXB200_OK)
        throw std::runtime_error(std::string("DCRTNttB200::SwitchFormat: ") + luxb200_last_error());
    const Format to = from == Format::EVALUATION ? Format::COEFFICIENT : Format::EVALUATION;
#pragma omp parallel for collapse(2)
    for (size_t i = 0; i < polys.size(); ++i)
        for (size_t l = 0; l < L; ++l) {
            NativePoly& limb = polys[i]->GetAllElements()[l];
            std::memcpy(&limb[0], &stage_[(i * L + l) * N], N * 8);
            limb.OverrideFormat(to);
        }
    for (DCRTPoly* p : polys) p->OverrideFormat(to);
}

void DCRTNttB200::SwitchFormat(DCRTPoly& poly) {
    std::vector<DCRTPoly*> one{&poly};
    SwitchFormat(one);
}

}  // namespace backend
}  // namespace lux::fhe
