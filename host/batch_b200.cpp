// batch_b200.cpp — bodies of the reference's batch API (src/binfhe/include/batch/binfhe-batch.h:48-279) routed to
// the B200 backend.  Linked INSTEAD of src/binfhe/lib/batch/batch.cpp: same names, arguments, BatchResult
// semantics and error strings; each function computes, per element, exactly what the reference body computes
// (a loop over BinFHEContext::EvalBinGate / EvalFunc / Bootstrap, batch.cpp:124-129,197-200,233-239), so the outputs
// are bit-identical.  BATCH_CPU_FORCE keeps the reference's own per-ciphertext loop (an explicit caller request,
// not a fallback): without that flag a missing or unbound B200 backend is an error, never a silent CPU path.
//
// BatchDAG::Execute (sequential in the reference, batch.cpp:429-482) is a level scheduler here.  A DAG made of 2-input
// gates, Bootstrap, CMUX and EvalFunc nodes over ciphertexts mod q goes to the device as ONE luxb200_circuit_batch_ex
// call (levels scheduled there, intermediates resident in HBM); anything else (other ciphertext moduli, a Bootstrap of a
// ciphertext whose plaintext modulus is not 4) is levelled on the host and every level issued as one batched call per
// (operation, gate | LUT).
#include <algorithm>
#include <future>
#include <map>
#include <stdexcept>

#include "backend_b200.h"
#include "batch/binfhe-batch.h"

namespace lux::fhe {

namespace {

// The registered B200 backend, attached to `cc`: its resident parameters and keys are made cc's (BackendB200::Attach), so
// a second BTKeyGen or a second BinFHEContext in the process never runs on stale or foreign keys — the reference always
// uses cc's current keys (binfhecontext.cpp:310-322).
backend::BackendB200* b200(BinFHEContext& cc) {
    auto* be = dynamic_cast<backend::BackendB200*>(backend::BackendRegistry::Instance().Get(backend::BackendType::CUDA));
    if (!be || !be->IsBound())
        throw std::runtime_error("B200 backend is not registered (call backend::RegisterB200 after BTKeyGen); no CPU fallback");
    be->Attach(cc);
    return be;
}

template <typename F>
BatchResult guarded(size_t count, F&& body) {
    BatchResult result{true, 0, 0, ""};
    if (count == 0) return result;
    try {
        body();
        result.processed = count;
    } catch (const std::exception& e) {
        result.success = false;
        result.failed  = count;
        result.error   = e.what();
    }
    return result;
}

}  // namespace

// binfhe-batch.h:48-58.  The reference body is a placeholder that returns its input (batch.cpp:86-94); the
// documented behaviour — "refresh noise for multiple ciphertexts" — is BinFHEContext::Bootstrap per element.
BatchResult BootstrapBatch(BinFHEContext& cc, const std::vector<LWECiphertext>& ct_in, std::vector<LWECiphertext>& ct_out,
                           uint32_t flags) {
    return guarded(ct_in.size(), [&] {
        if (flags & BATCH_CPU_FORCE) {
            ct_out.resize(ct_in.size());
            for (size_t i = 0; i < ct_in.size(); ++i) ct_out[i] = cc.Bootstrap(ct_in[i]);
        } else {
            b200(cc)->Bootstrap(ct_in, ct_out);
        }
    });
}

BatchResult EvalFuncBatch(BinFHEContext& cc, const std::vector<LWECiphertext>& ct_in, const std::vector<NativeInteger>& lut,
                          std::vector<LWECiphertext>& ct_out, uint32_t flags) {
    return guarded(ct_in.size(), [&] {
        if (flags & BATCH_CPU_FORCE) {
            ct_out.resize(ct_in.size());
            for (size_t i = 0; i < ct_in.size(); ++i) ct_out[i] = cc.EvalFunc(ct_in[i], lut);
        } else {
            b200(cc)->EvalFunc(ct_in, lut, cc.GetBeta(), ct_out);
        }
    });
}

// ct_out[i * luts.size() + j] = EvalFunc(ct_in[i], luts[j])   (batch.cpp:151-160)
BatchResult EvalFuncMultiOutputBatch(BinFHEContext& cc, const std::vector<LWECiphertext>& ct_in,
                                     const std::vector<std::vector<NativeInteger>>& luts, std::vector<LWECiphertext>& ct_out,
                                     uint32_t flags) {
    if (luts.empty()) return BatchResult{true, 0, 0, ""};
    return guarded(ct_in.size(), [&] {
        const size_t m = luts.size();
        ct_out.resize(ct_in.size() * m);
        for (size_t j = 0; j < m; ++j) {
            std::vector<LWECiphertext> col;
            if (flags & BATCH_CPU_FORCE) {
                col.resize(ct_in.size());
                for (size_t i = 0; i < ct_in.size(); ++i) col[i] = cc.EvalFunc(ct_in[i], luts[j]);
            } else {
                b200(cc)->EvalFunc(ct_in, luts[j], cc.GetBeta(), col);
            }
            for (size_t i = 0; i < ct_in.size(); ++i) ct_out[i * m + j] = col[i];
        }
    });
}

BatchResult EvalBinGateBatch(BinFHEContext& cc, BINGATE gate, const std::vector<LWECiphertext>& ct1,
                             const std::vector<LWECiphertext>& ct2, std::vector<LWECiphertext>& ct_out, uint32_t flags) {
    if (ct1.size() != ct2.size()) return BatchResult{false, 0, ct1.size(), "Input size mismatch"};
    return guarded(ct1.size(), [&] {
        if (flags & BATCH_CPU_FORCE) {
            ct_out.resize(ct1.size());
            for (size_t i = 0; i < ct1.size(); ++i) ct_out[i] = cc.EvalBinGate(gate, ct1[i], ct2[i]);
        } else {
            b200(cc)->EvalBinGate(gate, ct1, ct2, ct_out);
        }
    });
}

// The reference passes {sel, true, false} to EvalBinGate(CMUX, vector) (batch.cpp:231-239); the same vector order is
// kept here so that results are identical to the reference's EvalCMUXBatch on the same inputs.
BatchResult EvalCMUXBatch(BinFHEContext& cc, const std::vector<LWECiphertext>& ct_sel, const std::vector<LWECiphertext>& ct_true,
                          const std::vector<LWECiphertext>& ct_false, std::vector<LWECiphertext>& ct_out, uint32_t flags) {
    if (ct_sel.size() != ct_true.size() || ct_sel.size() != ct_false.size())
        return BatchResult{false, 0, ct_sel.size(), "Input size mismatch"};
    return guarded(ct_sel.size(), [&] {
        if (flags & BATCH_CPU_FORCE) {
            ct_out.resize(ct_sel.size());
            for (size_t i = 0; i < ct_sel.size(); ++i)
                ct_out[i] = cc.EvalBinGate(CMUX, std::vector<LWECiphertext>{ct_sel[i], ct_true[i], ct_false[i]});
        } else {
            std::vector<std::vector<LWECiphertext>> in(ct_sel.size());
            for (size_t i = 0; i < ct_sel.size(); ++i) in[i] = {ct_sel[i], ct_true[i], ct_false[i]};
            b200(cc)->EvalBinGateMulti(CMUX, in, ct_out);
        }
    });
}

BatchResult KeySwitchBatch(BinFHEContext& cc, const std::vector<LWECiphertext>& ct_in, std::vector<LWECiphertext>& ct_out,
                           uint32_t) {
    return guarded(ct_in.size(), [&] {
        b200(cc)->KeySwitchBatch(cc.GetParams()->GetLWEParams(), ct_in, cc.GetSwitchKey(), ct_out);
    });
}

BatchResult ModSwitchBatch(BinFHEContext& cc, const std::vector<LWECiphertext>& ct_in, std::vector<LWECiphertext>& ct_out,
                           uint32_t) {
    return guarded(ct_in.size(), [&] { b200(cc)->ModSwitchBatch(cc.GetParams()->GetLWEParams(), ct_in, ct_out); });
}

// ---- async variants (binfhe-batch.h:181-223): one std::async thread per call, like batch.cpp:320-349 ------------
struct BatchFuture::Impl {
    std::future<BatchResult> future;
};
BatchFuture::~BatchFuture() = default;
BatchResult BatchFuture::Wait() {
    if (!impl_ || !impl_->future.valid()) return BatchResult{false, 0, 0, "Invalid future"};
    return impl_->future.get();
}
bool BatchFuture::IsReady() const {
    return impl_ && impl_->future.valid() && impl_->future.wait_for(std::chrono::seconds(0)) == std::future_status::ready;
}
bool BatchFuture::Cancel() { return false; }

BatchFuture BootstrapBatchAsync(BinFHEContext& cc, const std::vector<LWECiphertext>& ct_in, std::vector<LWECiphertext>& ct_out) {
    BatchFuture f;
    f.impl_         = std::make_shared<BatchFuture::Impl>();
    f.impl_->future = std::async(std::launch::async, [&cc, &ct_in, &ct_out]() { return BootstrapBatch(cc, ct_in, ct_out, BATCH_ASYNC); });
    return f;
}
BatchFuture EvalFuncBatchAsync(BinFHEContext& cc, const std::vector<LWECiphertext>& ct_in, const std::vector<NativeInteger>& lut,
                               std::vector<LWECiphertext>& ct_out) {
    BatchFuture f;
    f.impl_         = std::make_shared<BatchFuture::Impl>();
    f.impl_->future = std::async(std::launch::async, [&cc, &ct_in, &lut, &ct_out]() { return EvalFuncBatch(cc, ct_in, lut, ct_out, BATCH_ASYNC); });
    return f;
}

// ---- BatchDAG (binfhe-batch.h:255-279) ------------------------------------------------------------------------
struct BatchDAG::Impl {
    BinFHEContext& cc;
    std::vector<LWECiphertext> ciphertexts;
    std::vector<BatchOp> operations;
    explicit Impl(BinFHEContext& context) : cc(context) {}
    size_t add(BatchOp::Type t, std::vector<size_t> in, BINGATE g = AND, const std::vector<NativeInteger>* lut = nullptr) {
        const size_t out = ciphertexts.size();
        ciphertexts.push_back(nullptr);
        BatchOp op;
        op.type = t; op.input_ids = std::move(in); op.output_ids = {out}; op.gate = g;
        if (lut) op.lut = *lut;
        operations.push_back(std::move(op));
        return out;
    }
};
BatchDAG::BatchDAG(BinFHEContext& cc) : impl_(std::make_unique<Impl>(cc)) {}
BatchDAG::~BatchDAG() = default;
size_t BatchDAG::AddCiphertext(const LWECiphertext& ct) {
    impl_->ciphertexts.push_back(ct);
    return impl_->ciphertexts.size() - 1;
}
size_t BatchDAG::AddBootstrap(size_t input_id) { return impl_->add(BatchOp::OP_BOOTSTRAP, {input_id}); }
size_t BatchDAG::AddEvalFunc(size_t input_id, const std::vector<NativeInteger>& lut) {
    return impl_->add(BatchOp::OP_EVAL_FUNC, {input_id}, AND, &lut);
}
size_t BatchDAG::AddBinGate(BINGATE gate, size_t input1_id, size_t input2_id) {
    return impl_->add(BatchOp::OP_BINGATE, {input1_id, input2_id}, gate);
}
size_t BatchDAG::AddCMUX(size_t sel_id, size_t true_id, size_t false_id) {
    return impl_->add(BatchOp::OP_CMUX, {sel_id, true_id, false_id});
}

namespace {
bool gate2(BINGATE g) { return g == OR || g == AND || g == NOR || g == NAND || g == XOR || g == XNOR || g == XOR_FAST || g == XNOR_FAST; }
}  // namespace

// The whole DAG as one device-side circuit.  Returns false (nothing done) when the DAG holds something the circuit entry
// point does not cover; throws what the per-operation path would throw.
static bool execute_on_device(BinFHEContext& cc, std::vector<LWECiphertext>& cts, const std::vector<BatchOp>& ops,
                              BatchResult& result) {
    if (ops.empty()) return false;
    const NativeInteger q = cc.GetParams()->GetLWEParams()->Getq();
    std::vector<char> produced(cts.size(), 0);
    for (const auto& op : ops) {
        if (op.type == BatchOp::OP_BINGATE ? !gate2(op.gate)
                                           : (op.type != BatchOp::OP_BOOTSTRAP && op.type != BatchOp::OP_CMUX && op.type != BatchOp::OP_EVAL_FUNC))
            return false;
        if (op.type == BatchOp::OP_EVAL_FUNC && op.lut.size() != q.ConvertToInt<size_t>()) return false;
        produced[op.output_ids[0]] = 1;
    }
    // wires: the DAG's input ciphertexts first, then one wire per gate (CMUX = NOT + 3 NAND, binfhe-base-scheme.cpp:176-183)
    // pmod: the plaintext modulus a wire's ciphertext carries.  Only Bootstrap reads it (the offset Q/(2p)+1,
    // binfhe-base-scheme.cpp:214); gates and EvalFunc return fresh ciphertexts (m_p = 4), Bootstrap hands its input's on.
    std::vector<uint32_t> wire(cts.size(), 0);
    std::vector<uint64_t> pmod(cts.size(), 4);
    std::vector<LWECiphertext> inputs;
    for (size_t id = 0; id < cts.size(); ++id) {
        if (produced[id]) continue;
        if (!cts[id]) return false;   // never referenced or a caller error: the per-operation path reports it
        if (cts[id]->GetModulus() != q) return false;
        pmod[id] = cts[id]->GetptModulus().ConvertToInt<uint64_t>();
        wire[id] = static_cast<uint32_t>(inputs.size());
        inputs.push_back(cts[id]);
    }
    if (inputs.empty()) return false;
    std::vector<luxb200_gate> gates;
    std::vector<uint32_t> outs;
    std::vector<std::vector<NativeInteger>> luts;   // distinct tables of the EvalFunc nodes
    const uint32_t ni = static_cast<uint32_t>(inputs.size());
    auto emit = [&](uint32_t op, uint32_t a, uint32_t b) {
        gates.push_back(luxb200_gate{op, a, b});
        return ni + static_cast<uint32_t>(gates.size()) - 1;
    };
    for (const auto& op : ops) {
        for (size_t id : op.input_ids)
            if (id >= op.output_ids[0]) throw std::out_of_range("Invalid ciphertext ID");
        uint32_t w;
        if (op.type == BatchOp::OP_BINGATE) {
            w = emit(static_cast<uint32_t>(op.gate), wire[op.input_ids[0]], wire[op.input_ids[1]]);
        } else if (op.type == BatchOp::OP_BOOTSTRAP) {
            if (pmod[op.input_ids[0]] != 4) return false;   // the device circuit bootstraps with p = 4
            w = emit(LUXB200_OP_BOOTSTRAP, wire[op.input_ids[0]], 0);
        } else if (op.type == BatchOp::OP_EVAL_FUNC) {
            size_t li = 0;
            while (li < luts.size() && luts[li] != op.lut) ++li;
            if (li == luts.size()) luts.push_back(op.lut);
            w = emit(LUXB200_OP_EVALFUNC, wire[op.input_ids[0]], static_cast<uint32_t>(li));
        } else {   // EvalBinGate(CMUX, {v0, v1, v2}) = NAND(NAND(v0, NOT v2), NAND(v1, v2))
            const uint32_t v0 = wire[op.input_ids[0]], v1 = wire[op.input_ids[1]], v2 = wire[op.input_ids[2]];
            const uint32_t t  = emit(LUXB200_OP_NOT, v2, 0);
            const uint32_t n1 = emit(NAND, v0, t), n2 = emit(NAND, v1, v2);
            w = emit(NAND, n1, n2);
        }
        wire[op.output_ids[0]] = w;
        pmod[op.output_ids[0]] = 4;
        outs.push_back(w);
    }
    std::vector<LWECiphertext> res;
    b200(cc)->EvalCircuit(inputs, gates, outs, luts, cc.GetBeta(), res);
    for (size_t i = 0; i < ops.size(); ++i) cts[ops[i].output_ids[0]] = res[i];
    result.processed = ops.size();
    return true;
}

BatchResult BatchDAG::Execute(uint32_t flags) {
    BatchResult result{true, 0, 0, ""};
    auto& cts = impl_->ciphertexts;
    auto& ops = impl_->operations;
    try {
        if (!(flags & BATCH_CPU_FORCE) && execute_on_device(impl_->cc, cts, ops, result)) return result;
        // level of a ciphertext = 0 for inputs, 1 + max(level of operands) for an operation's output
        std::vector<size_t> level(cts.size(), 0);
        size_t depth = 0;
        for (const auto& op : ops) {
            size_t l = 0;
            for (size_t id : op.input_ids) {
                if (id >= op.output_ids[0]) throw std::out_of_range("Invalid ciphertext ID");
                l = std::max(l, level[id]);
            }
            level[op.output_ids[0]] = l + 1;
            depth = std::max(depth, l + 1);
        }
        for (size_t lv = 1; lv <= depth && result.success; ++lv) {
            // group this level's operations: (type, gate) and, for EvalFunc, identical LUTs
            std::vector<std::vector<const BatchOp*>> groups;
            for (const auto& op : ops) {
                if (level[op.output_ids[0]] != lv) continue;
                bool placed = false;
                for (auto& g : groups)
                    if (g[0]->type == op.type && (op.type != BatchOp::OP_BINGATE || g[0]->gate == op.gate) &&
                        (op.type != BatchOp::OP_EVAL_FUNC || g[0]->lut == op.lut)) {
                        g.push_back(&op);
                        placed = true;
                        break;
                    }
                if (!placed) groups.push_back({&op});
            }
            for (const auto& g : groups) {
                std::vector<LWECiphertext> in0, in1, in2, out;
                for (const BatchOp* op : g) {
                    in0.push_back(cts[op->input_ids[0]]);
                    if (op->input_ids.size() > 1) in1.push_back(cts[op->input_ids[1]]);
                    if (op->input_ids.size() > 2) in2.push_back(cts[op->input_ids[2]]);
                }
                BatchResult r{true, 0, 0, ""};
                switch (g[0]->type) {
                    case BatchOp::OP_BOOTSTRAP: r = BootstrapBatch(impl_->cc, in0, out, flags); break;
                    case BatchOp::OP_EVAL_FUNC: r = EvalFuncBatch(impl_->cc, in0, g[0]->lut, out, flags); break;
                    case BatchOp::OP_BINGATE: r = EvalBinGateBatch(impl_->cc, g[0]->gate, in0, in1, out, flags); break;
                    case BatchOp::OP_CMUX: r = EvalCMUXBatch(impl_->cc, in0, in1, in2, out, flags); break;
                    default: out = in0; break;
                }
                if (!r.success) {
                    result.success = false;
                    result.error   = r.error;
                    result.failed += g.size();
                    break;
                }
                for (size_t i = 0; i < g.size(); ++i) cts[g[i]->output_ids[0]] = out[i];
                result.processed += g.size();
            }
        }
    } catch (const std::exception& e) {
        result.success = false;
        result.error   = e.what();
    }
    return result;
}

LWECiphertext BatchDAG::GetResult(size_t id) const {
    if (id >= impl_->ciphertexts.size()) throw std::out_of_range("Invalid ciphertext ID");
    return impl_->ciphertexts[id];
}
void BatchDAG::Clear() {
    impl_->ciphertexts.clear();
    impl_->operations.clear();
}

}  // namespace lux::fhe
