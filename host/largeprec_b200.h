// largeprec_b200.h — batched large-precision floor / sign / digit decomposition on the B200 for luxcpp/fhe contexts made by
// BinFHEContext::GenerateBinFHEContext(set, arbFunc, logQ, N, GINX, timeOptimization) (binfhecontext.cpp:53-104).
// Additive to the reference's batch API (binfhe-batch.h has no entry for these; the per-ciphertext calls are
// BinFHEContext::EvalFloor / EvalSign / EvalDecomp, binfhecontext.cpp:346-370): each function computes, per element,
// exactly what the reference call computes (bit-identical ciphertexts), one device call per batch.
//
// One device context per gadget base of the context's key map (cc.GetBTKeyMap(): one entry after a plain BTKeyGen, three —
// baseG = 2^14, 2^18, 2^27 — with timeOptimization, binfhecontext.cpp:292-298); with three the device loop changes the base
// with the shrinking modulus like binfhe-base-scheme.cpp:415-434.  Contexts and keys are cached and follow cc (a new key
// object for a base is reloaded).  No arithmetic here: pack -> luxb200_eval{floor,sign,decomp}_batch -> unpack.
#pragma once
#include <vector>

#include "batch/binfhe-batch.h"
#include "binfhecontext.h"

namespace lux::fhe {

BatchResult EvalFloorBatch(BinFHEContext& cc, const std::vector<LWECiphertext>& ct_in, std::vector<LWECiphertext>& ct_out,
                           uint32_t roundbits = 0, uint32_t flags = BATCH_DEFAULT);
BatchResult EvalSignBatch(BinFHEContext& cc, const std::vector<LWECiphertext>& ct_in, std::vector<LWECiphertext>& ct_out,
                          bool schemeSwitch = false, uint32_t flags = BATCH_DEFAULT);
// ct_out[i] = cc.EvalDecomp(ct_in[i]) (the digits of ciphertext i, most significant last)
BatchResult EvalDecompBatch(BinFHEContext& cc, const std::vector<LWECiphertext>& ct_in,
                            std::vector<std::vector<LWECiphertext>>& ct_out, uint32_t flags = BATCH_DEFAULT);
// release the cached device contexts (keys of every gadget base)
void ReleaseLargePrecisionB200();

}  // namespace lux::fhe
