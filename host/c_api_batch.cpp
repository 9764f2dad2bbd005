// c_api_batch.cpp — the reference's C ABI plus the batch entry points of include/lux/fhe/c_api_batch.h.
// LuxFheContext / LuxFheCiphertext are private structs of src/c_api/c_api.cpp, so the additions have to live in
// that translation unit: this file compiles the reference's c_api.cpp from where it lies (host/Makefile passes
// -I$(REF)/src/c_api) and appends the new functions — what a maintainer would paste at the end of c_api.cpp.
#include "c_api.cpp"   // NOLINT: the reference's own source, unmodified

#include "backend_b200.h"
#include "batch/binfhe-batch.h"
#include "lux/fhe/c_api_batch.h"

namespace {

bool gather(const LuxFheCiphertext* const* in, size_t count, std::vector<LWECiphertext>& out) {
    out.resize(count);
    for (size_t i = 0; i < count; ++i) {
        if (!in[i] || !in[i]->ct) return false;
        out[i] = in[i]->ct;
    }
    return true;
}
LuxFheError scatter(const std::vector<LWECiphertext>& v, LuxFheCiphertext** results) {
    for (size_t i = 0; i < v.size(); ++i) {
        auto* r = new LuxFheCiphertext();
        r->ct   = v[i];
        results[i] = r;
    }
    return LUX_FHE_OK;
}
bool gpu_bound() {
    auto* be = dynamic_cast<backend::BackendB200*>(backend::BackendRegistry::Instance().Get(backend::BackendType::CUDA));
    return be && be->IsBound();
}
LuxFheError gate_batch(LuxFheContext* ctx, const LuxFheBootstrapKey* bsk, BINGATE g, const LuxFheCiphertext* const* a,
                       const LuxFheCiphertext* const* b, size_t count, LuxFheCiphertext** results) {
    if (!ctx || !bsk || !a || !b || !results) return LUX_FHE_ERR_NULL_PTR;
    if (!bsk->generated || !gpu_bound()) return LUX_FHE_ERR_NOT_INIT;
    try {
        std::vector<LWECiphertext> va, vb, vo;
        if (!gather(a, count, va) || !gather(b, count, vb)) return LUX_FHE_ERR_NULL_PTR;
        const BatchResult r = EvalBinGateBatch(ctx->cc, g, va, vb, vo);
        if (!r.success) return LUX_FHE_ERR_GATE;
        return scatter(vo, results);
    } catch (...) { return LUX_FHE_ERR_GATE; }
}

}  // namespace

extern "C" {

LuxFheError lux_fhe_gpu_enable(LuxFheContext* ctx, const LuxFheBootstrapKey* bsk, int device) {
    if (!ctx || !bsk) return LUX_FHE_ERR_NULL_PTR;
    if (!bsk->generated) return LUX_FHE_ERR_NOT_INIT;
    try {
        backend::RegisterB200(ctx->cc, device);
        return LUX_FHE_OK;
    } catch (...) { return LUX_FHE_ERR_NOT_INIT; }
}

#define IMPL_GATE_BATCH(name, op)                                                                                 \
    LuxFheError lux_fhe_##name##_batch(LuxFheContext* ctx, const LuxFheBootstrapKey* bsk,                         \
                                       const LuxFheCiphertext* const* a, const LuxFheCiphertext* const* b,        \
                                       size_t count, LuxFheCiphertext** results) {                                \
        return gate_batch(ctx, bsk, op, a, b, count, results);                                                    \
    }
IMPL_GATE_BATCH(and, AND)
IMPL_GATE_BATCH(or, OR)
IMPL_GATE_BATCH(xor, XOR)
IMPL_GATE_BATCH(nand, NAND)
IMPL_GATE_BATCH(nor, NOR)
IMPL_GATE_BATCH(xnor, XNOR)

LuxFheError lux_fhe_mux_batch(LuxFheContext* ctx, const LuxFheBootstrapKey* bsk, const LuxFheCiphertext* const* sel,
                              const LuxFheCiphertext* const* a, const LuxFheCiphertext* const* b, size_t count,
                              LuxFheCiphertext** results) {
    if (!ctx || !bsk || !sel || !a || !b || !results) return LUX_FHE_ERR_NULL_PTR;
    if (!bsk->generated || !gpu_bound()) return LUX_FHE_ERR_NOT_INIT;
    try {
        std::vector<LWECiphertext> vs, va, vb, vo;
        if (!gather(sel, count, vs) || !gather(a, count, va) || !gather(b, count, vb)) return LUX_FHE_ERR_NULL_PTR;
        const BatchResult r = EvalCMUXBatch(ctx->cc, vs, va, vb, vo);   // {sel, a, b} like lux_fhe_mux (c_api.cpp:246-248)
        if (!r.success) return LUX_FHE_ERR_GATE;
        return scatter(vo, results);
    } catch (...) { return LUX_FHE_ERR_GATE; }
}

LuxFheError lux_fhe_bootstrap_batch(LuxFheContext* ctx, const LuxFheBootstrapKey* bsk, const LuxFheCiphertext* const* cts,
                                    size_t count, LuxFheCiphertext** results) {
    if (!ctx || !bsk || !cts || !results) return LUX_FHE_ERR_NULL_PTR;
    if (!bsk->generated || !gpu_bound()) return LUX_FHE_ERR_NOT_INIT;
    try {
        std::vector<LWECiphertext> vi, vo;
        if (!gather(cts, count, vi)) return LUX_FHE_ERR_NULL_PTR;
        const BatchResult r = BootstrapBatch(ctx->cc, vi, vo);
        if (!r.success) return LUX_FHE_ERR_BOOTSTRAP;
        return scatter(vo, results);
    } catch (...) { return LUX_FHE_ERR_BOOTSTRAP; }
}

}  // extern "C"
