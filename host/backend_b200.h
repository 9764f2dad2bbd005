// backend_b200.h — lux::fhe::backend::BackendB200: the B200 implementation of the reference's backend seam
// (src/binfhe/include/backend/backend.h:73-247), built on the C ABI of libluxb200.so (include/luxb200.h).
//
// This is host-side glue written against the reference's own headers; it is compiled by host/Makefile with
// -I/root/reference/src/... and linked to the reference library built by oracle/Makefile.  It contains no
// arithmetic: every virtual packs the reference's containers into flat uint64 arrays, calls one luxb200_* entry
// point, and unpacks.  Type() reports BackendType::CUDA, so `BackendRegistry::Instance().Get(BackendType::CUDA)`
// (backend.cpp:266-270) finds it after RegisterB200().
#pragma once
#include <luxb200.h>

#include <functional>
#include <memory>
#include <mutex>
#include <string>
#include <vector>

#include "backend/backend.h"
#include "binfhecontext.h"

namespace lux::fhe {
namespace backend {

class BackendB200 final : public Backend {
public:
    explicit BackendB200(int device = 0);
    // Several GPUs of one box behind one backend: one device context (stream, work buffers) per GPU, the keys uploaded once
    // to devices[0] and copied peer-to-peer over NVLink (luxb200_broadcast_keys), every batch split into contiguous
    // count/G slices that run concurrently, one host thread per GPU.  No collective on the per-gate path.
    explicit BackendB200(std::vector<int> devices);
    ~BackendB200() override;

    // Create the device context(s) for cc's parameters and upload cc's refresh and switching keys (after BTKeyGen).
    // Throws lux::fhe::config_error-style std::runtime_error on failure; there is no CPU fallback.
    void Bind(BinFHEContext& cc);
    // Make the resident parameters and keys those of `cc`: rebinds when cc's parameter set differs from the bound one and
    // reloads whichever of cc's current refresh / switching key objects is not the resident one (a second BTKeyGen, or a
    // second BinFHEContext in the process).  The batch API calls this on every entry, so a batch never runs on stale or
    // foreign keys.  A context without keys keeps keys loaded with LoadPacked.
    void Attach(BinFHEContext& cc);
    bool IsBound() const { return h_ != nullptr; }
    luxb200_ctx* handle() const { return h_; }
    int device() const { return devices_[0]; }
    size_t NumDevices() const { return devices_.size(); }

    BackendType Type() const override { return BackendType::CUDA; }
    std::string Name() const override { return "B200 (libluxb200, sm_100a)"; }
    bool IsAvailable() const override;
    size_t MaxBatchSize() const override { return size_t(1) << 24; }
    size_t DeviceMemory() const override;

    DeviceBuffer Allocate(size_t bytes) override;
    void Free(DeviceBuffer& buffer) override;
    void CopyToDevice(const void* host_ptr, DeviceBuffer& device_buffer, size_t bytes) override;
    void CopyToHost(const DeviceBuffer& device_buffer, void* host_ptr, size_t bytes) override;
    void Synchronize() override;

    void BlindRotate(const std::shared_ptr<RingGSWCryptoParams>& params, const LWECiphertext& ct, const RingGSWACCKey& ek,
                     RLWECiphertext& acc) override;
    void ExternalProduct(const std::shared_ptr<RingGSWCryptoParams>& params, const RingGSWEvalKey& rgsw,
                         const RLWECiphertext& rlwe, RLWECiphertext& result) override;
    void KeySwitch(const std::shared_ptr<LWECryptoParams>& params, const LWECiphertext& ct, const LWESwitchingKey& ks,
                   LWECiphertext& result) override;
    void ModSwitch(const std::shared_ptr<LWECryptoParams>& params, const LWECiphertext& ct, LWECiphertext& result) override;

    void BlindRotateBatch(const std::shared_ptr<RingGSWCryptoParams>& params, const std::vector<LWECiphertext>& cts,
                          const RingGSWACCKey& ek, std::vector<RLWECiphertext>& accs) override;
    void ExternalProductBatch(const std::shared_ptr<RingGSWCryptoParams>& params, const std::vector<RingGSWEvalKey>& rgsws,
                              const std::vector<RLWECiphertext>& rlwes, std::vector<RLWECiphertext>& results) override;
    void KeySwitchBatch(const std::shared_ptr<LWECryptoParams>& params, const std::vector<LWECiphertext>& cts,
                        const LWESwitchingKey& ks, std::vector<LWECiphertext>& results) override;
    void ModSwitchBatch(const std::shared_ptr<LWECryptoParams>& params, const std::vector<LWECiphertext>& cts,
                        std::vector<LWECiphertext>& results) override;

    // Packed formats: the canonical flat layouts of include/luxb200.h, resident in device memory (uint64).
    DeviceBuffer PackBootstrappingKey(const RingGSWACCKey& ek) override;
    void UnpackBootstrappingKey(const DeviceBuffer& packed, RingGSWACCKey& ek) override;
    DeviceBuffer PackCiphertexts(const std::vector<LWECiphertext>& cts) override;     // [count][n+1]: a then b
    void UnpackCiphertexts(const DeviceBuffer& packed, std::vector<LWECiphertext>& cts) override;

    // Keys from the packed wire formats of backend/packed.h (PackBootstrappingKey / PackSwitchingKey bytes, see
    // host/packed_b200.cpp) straight to the device; either pointer may be null.  Requires Bind() (parameters) first.
    void LoadPacked(const uint8_t* bsk, size_t bsk_size, const uint8_t* ksk, size_t ksk_size);

    // Gate-level entry points used by the batch API bodies (host/batch_b200.cpp)
    void EvalBinGate(BINGATE gate, const std::vector<LWECiphertext>& ct1, const std::vector<LWECiphertext>& ct2,
                     std::vector<LWECiphertext>& out);
    void EvalBinGateMulti(BINGATE gate, const std::vector<std::vector<LWECiphertext>>& cts, std::vector<LWECiphertext>& out);
    void Bootstrap(const std::vector<LWECiphertext>& in, std::vector<LWECiphertext>& out);
    void EvalFunc(const std::vector<LWECiphertext>& in, const std::vector<NativeInteger>& lut, const NativeInteger& beta,
                  std::vector<LWECiphertext>& out);
    // A gate DAG in one call (luxb200_circuit_batch, count = 1): wires 0 .. inputs.size()-1 are `inputs`, wire
    // inputs.size() + g the result of gates[g]; out[k] = wire outputs[k].  Intermediates stay on the device.
    // LUXB200_OP_EVALFUNC gates name a table of `luts` (BatchDAG::AddEvalFunc nodes); beta = cc.GetBeta().
    void EvalCircuit(const std::vector<LWECiphertext>& inputs, const std::vector<luxb200_gate>& gates,
                     const std::vector<uint32_t>& outputs, const std::vector<std::vector<NativeInteger>>& luts,
                     const NativeInteger& beta, std::vector<LWECiphertext>& out);

    // BinFHEContext::BTKeyGen (binfhecontext.cpp:285-307) with the refresh and switching keys generated on the device
    // (luxb200_btkeygen; ChaCha20 streams from `seed`) and loaded back into cc as the reference's own key objects
    // (cc.BTKeyLoad), so CPU and device evaluate with the same keys.  Not bit-equal to the reference's Blake2-drawn keys.
    void BTKeyGen(BinFHEContext& cc, ConstLWEPrivateKey sk, uint64_t seed);

    // host time (ms) the last gate-level call spent packing inputs / building outputs, and inside the C ABI call
    struct Timing { double pack_ms = 0, call_ms = 0, unpack_ms = 0; };
    Timing LastTiming() const { return timing_; }

private:
    // grow-only page-locked staging buffer (luxb200_host_alloc)
    struct Pinned {
        uint64_t* p = nullptr;
        size_t cap = 0;
        uint64_t* ensure(size_t elems);
        void release();
    };
    void check(int rc, const char* what) const;
    void flatten_bsk(const RingGSWACCKeyImpl& ek, std::vector<uint64_t>& out) const;
    void unflatten_bsk(const std::vector<uint64_t>& flat, RingGSWACCKey& ek) const;
    void load_keys(const RingGSWACCKey& ek, const LWESwitchingKey& ks);
    void ensure_bsk(const RingGSWACCKey& ek);
    void ensure_ksk(const LWESwitchingKey& ks);
    void share_keys();
    void destroy_contexts();
    // contiguous [lo, hi) slices of a batch over the devices (sizes differ by at most one); body(device index, lo, hi) runs
    // on one host thread per device; the first exception is rethrown
    void run_split(size_t count, const std::function<void(size_t, size_t, size_t)>& body);
    void gate2(BINGATE gate, const std::vector<LWECiphertext>& ct1, const std::vector<LWECiphertext>& ct2,
               std::vector<LWECiphertext>& out);

    std::vector<int> devices_;
    std::vector<luxb200_ctx*> hs_;      // one context per device; h_ == hs_[0]
    luxb200_ctx* h_ = nullptr;
    luxb200_params p_{};
    uint32_t digitsG_ = 0, digitsR_ = 0, rows_ = 0;
    std::shared_ptr<ILNativeParams> poly_params_;
    // the key OBJECTS resident on the device(s), held so that their addresses cannot be recycled by a new key while they
    // serve as the cache identity; packed_* for keys that came from LoadPacked bytes
    RingGSWACCKey bound_bsk_;
    LWESwitchingKey bound_ksk_;
    bool packed_bsk_ = false, packed_ksk_ = false;
    std::recursive_mutex mu_;           // entry points serialise: staging buffers and key binding are shared state
    Pinned st_[6];
    Timing timing_;
};

// Registers a BackendB200 bound to cc with the reference's registry and makes it the default backend.
// Returns the registered instance (owned by the registry).
BackendB200* RegisterB200(BinFHEContext& cc, int device = 0);
// The same over several GPUs of the box (devices = {0, 1, ...}); an empty list means every visible GPU.
BackendB200* RegisterB200(BinFHEContext& cc, const std::vector<int>& devices);

}  // namespace backend
}  // namespace lux::fhe
