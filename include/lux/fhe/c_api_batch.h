/* c_api_batch.h — additive batch entry points for the reference's C ABI (include/lux/fhe/c_api.h:102-321, which has
 * none).  Same conventions as the single-ciphertext calls they extend: LuxFheError return codes, opaque handles,
 * outputs `new`-allocated by the library and released with lux_fhe_ciphertext_free, no exception crosses the boundary
 * (src/c_api/c_api.cpp:82-84,234).  Each call computes, per element, exactly what the single call computes:
 *   lux_fhe_<gate>_batch   = lux_fhe_<gate> (c_api.h:198-250)   -> lux::fhe::EvalBinGateBatch
 *   lux_fhe_mux_batch      = lux_fhe_mux (c_api.h:252-259)      -> lux::fhe::EvalCMUXBatch  ({sel, a, b} vector order)
 *   lux_fhe_bootstrap_batch= lux_fhe_bootstrap (c_api.h:266-271)-> lux::fhe::BootstrapBatch
 * lux_fhe_gpu_enable binds the context's parameters and keys to a B200 (after lux_fhe_keygen_bootstrap); without it, or
 * without a CUDA device, the batch calls fail with LUX_FHE_ERR_NOT_INIT — they never fall back to the CPU loop. */
#ifndef LUX_FHE_C_API_BATCH_H
#define LUX_FHE_C_API_BATCH_H
#include "lux/fhe/c_api.h"

#ifdef __cplusplus
extern "C" {
#endif

LuxFheError lux_fhe_gpu_enable(LuxFheContext* ctx, const LuxFheBootstrapKey* bsk, int device);

#define LUX_FHE_DECL_GATE_BATCH(name)                                                                             \
    LuxFheError lux_fhe_##name##_batch(LuxFheContext* ctx, const LuxFheBootstrapKey* bsk,                         \
                                       const LuxFheCiphertext* const* a, const LuxFheCiphertext* const* b,        \
                                       size_t count, LuxFheCiphertext** results)
LUX_FHE_DECL_GATE_BATCH(and);
LUX_FHE_DECL_GATE_BATCH(or);
LUX_FHE_DECL_GATE_BATCH(xor);
LUX_FHE_DECL_GATE_BATCH(nand);
LUX_FHE_DECL_GATE_BATCH(nor);
LUX_FHE_DECL_GATE_BATCH(xnor);

LuxFheError lux_fhe_mux_batch(LuxFheContext* ctx, const LuxFheBootstrapKey* bsk, const LuxFheCiphertext* const* sel,
                              const LuxFheCiphertext* const* a, const LuxFheCiphertext* const* b, size_t count,
                              LuxFheCiphertext** results);
LuxFheError lux_fhe_bootstrap_batch(LuxFheContext* ctx, const LuxFheBootstrapKey* bsk, const LuxFheCiphertext* const* cts,
                                    size_t count, LuxFheCiphertext** results);

#ifdef __cplusplus
}
#endif
#endif
