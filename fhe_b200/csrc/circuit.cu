// circuit.cu — luxb200_circuit_batch[_ex]: a gate DAG (lux::fhe::BatchDAG, binfhe-batch.h:255-279) evaluated level by level
// on the device.  The reference's BatchDAG::Execute (src/binfhe/lib/batch/batch.cpp:429-482) walks the operations one by one
// on the CPU; here the host only expands the operations into bootstraps and sorts those into dependency levels.  A level is
// three launches over all of its (bootstrap, input set) pairs — mixed gate kinds and look-up tables included, through a
// per-item test-vector table offset (BrParams::lutsel):
//   circuit_gather_kernel  ct = mul * (wire[in0] +- wire[in1]) + const  (mod m)           binfhe-base-scheme.cpp:99-107, :201, :255-332
//   blind rotation         BootstrapGateCore / BootstrapFuncCore + EvalAcc + extraction     :525-626, :111-121
//   key switch             SwitchCTtoqn, written straight into the level's wire slots       lwe-pke.cpp:170-178
// EvalFunc nodes (BatchOp::OP_EVAL_FUNC, batch.cpp:455-460; BinFHEScheme::EvalFunc, binfhe-base-scheme.cpp:241-332) become
// one bootstrap (negacyclic table) or two chained ones (periodic / arbitrary tables: the second runs a level later on
// ct - Boot(ct + beta) + beta - q/4).  Bootstraps of a level that differ in ciphertext modulus or output offset (gates:
// modulus q, b += Q/8+1; function bootstraps: b unchanged; arbitrary functions: modulus 2q) form separate launch groups.
// Intermediate ciphertexts never leave HBM (u32 wire pool [wires][count][n] + [wires][count]).
#include <algorithm>

#include "api_common.h"

typedef uint32_t u32;
typedef uint64_t u64;

namespace {

inline unsigned nblk(u64 total, unsigned per) { return (unsigned)((total + per - 1) / per); }

constexpr u32 NONE = 0xffffffffu;

// one bootstrap of a level: ct = mul * (wire[s0] (+|-) wire[s1]) mod m, b += cst; test-vector table at element offset toff
struct CircItem {
    u32 s0, s1;      // source slots (s1 = NONE: single operand)
    u32 flags;       // bits 0..3 mul (1 or 2), bit 4: subtract wire[s1]
    u32 cst;         // constant added to b (mod m)
    u32 toff;        // element offset of the item's table inside the uploaded tables
    u32 pad[3];
};

__global__ void circuit_gather_kernel(const u32* __restrict__ pool_a, const u32* __restrict__ pool_b,
                                      const CircItem* __restrict__ desc, u32* __restrict__ wa, u32* __restrict__ wb,
                                      u32* __restrict__ lutsel, u32 n, u32 m, u64 count, u64 item0, u64 items) {
    const u64 idx = (u64)blockIdx.x * blockDim.x + threadIdx.x;
    const u64 per = (u64)n + 1;
    if (idx >= items * per) return;
    const u64 it  = idx / per;
    const u32 t   = (u32)(idx % per);
    const u64 git = item0 + it;
    const CircItem d = desc[git / count];
    const u64 i   = git % count;
    const u64 r0  = (u64)d.s0 * count + i;
    u32 x, y = 0;
    if (t < n) {
        x = pool_a[r0 * n + t];
        if (d.s1 != NONE) y = pool_a[((u64)d.s1 * count + i) * n + t];
    } else {
        x = pool_b[r0];
        if (d.s1 != NONE) y = pool_b[(u64)d.s1 * count + i];
    }
    u32 s = x % m;
    if (d.s1 != NONE) s = (d.flags & 16u) ? (s + m - y % m) % m : (s + y % m) % m;
    s = (u32)(((u64)s * (d.flags & 15u)) % m);
    if (t < n) {
        wa[it * n + t] = s;
    } else {
        wb[it]     = (u32)(((u64)s + d.cst) % m);
        lutsel[it] = d.toff;
    }
}

// EvalNOT (binfhe-base-scheme.cpp:223-237) from one wire slot to another: a -> -a, b -> q/4 - b
__global__ void circuit_not_kernel(u32* __restrict__ pool_a, u32* __restrict__ pool_b, u64 src_row, u64 dst_row, u32 n, u32 q,
                                   u64 count) {
    const u64 idx = (u64)blockIdx.x * blockDim.x + threadIdx.x;
    const u64 per = (u64)n + 1;
    if (idx >= count * per) return;
    const u64 ct = idx / per;
    const u32 t  = (u32)(idx % per);
    if (t < n) {
        const u32 v                  = pool_a[(src_row + ct) * n + t];
        pool_a[(dst_row + ct) * n + t] = v == 0 ? 0 : q - v;
    } else {
        const u32 v          = pool_b[src_row + ct];
        pool_b[dst_row + ct] = ((q >> 2) + q - v) % q;
    }
}

// ct->SetModulus(q) (lwe-ciphertext.h:116-120) on a run of wire slots: every entry mod q
__global__ void circuit_mod_kernel(u32* __restrict__ pool_a, u32* __restrict__ pool_b, u64 row0, u64 rows, u32 n, u32 q) {
    const u64 idx = (u64)blockIdx.x * blockDim.x + threadIdx.x;
    const u64 per = (u64)n + 1;
    if (idx >= rows * per) return;
    const u64 r = idx / per;
    const u32 t = (u32)(idx % per);
    if (t < n) pool_a[(row0 + r) * n + t] %= q;
    else pool_b[row0 + r] %= q;
}

__global__ void circuit_narrow_kernel(const u64* __restrict__ in, u32* __restrict__ out, u32 q, u64 len) {
    const u64 idx = (u64)blockIdx.x * blockDim.x + threadIdx.x;
    if (idx < len) out[idx] = (u32)(in[idx] % q);
}

bool is_gate2(u32 op) {
    switch (op) {
        case LUXB200_OR: case LUXB200_AND: case LUXB200_NOR: case LUXB200_NAND: case LUXB200_XOR: case LUXB200_XNOR:
        case LUXB200_XOR_FAST: case LUXB200_XNOR_FAST:
            return true;
        default:
            return false;
    }
}

// table index of an op inside the uploaded [6][q] gate tables (OR, AND, NOR, NAND, XOR, XNOR)
u32 table_of(u32 op) {
    if (op == LUXB200_XOR_FAST) return LUXB200_XOR;
    if (op == LUXB200_XNOR_FAST) return LUXB200_XNOR;
    if (op == LUXB200_OP_BOOTSTRAP) return LUXB200_AND;   // BootstrapGateCore(params, AND, ...), binfhe-base-scheme.cpp:205
    return op;
}

// binfhe-base-scheme.h:245-260
int check_input_function(const uint64_t* lut, uint64_t len, uint64_t mod) {
    const uint64_t mid = len / 2;
    if (lut[0] == mod - lut[mid]) {
        for (uint64_t i = 1; i < mid; ++i)
            if (lut[i] != mod - lut[mid + i]) return 2;
        return 0;
    }
    if (lut[0] == lut[mid]) {
        for (uint64_t i = 1; i < mid; ++i)
            if (lut[i] != lut[mid + i]) return 2;
        return 1;
    }
    return 2;
}

// launch groups of a level: (ciphertext modulus of the bootstrap, modulus of its output, offset added to b)
enum { GRP_GATE = 0, GRP_FUNC = 1, GRP_FUNC2Q = 2, GRP_COUNT = 3 };

// a bootstrap (or an EvalNOT) of the expanded circuit, over internal wire ids
struct Micro {
    u32 kind;        // 0 bootstrap, 1 EvalNOT
    u32 in0, in1;    // internal wires
    u32 flags;       // CircItem::flags
    u32 cst, toff;
    u32 grp;
    u32 mod_q;       // arbitrary functions: reduce the output mod q afterwards (ct4->SetModulus(q))
};

}  // namespace

extern "C" int luxb200_circuit_batch_ex(luxb200_ctx* c, uint32_t n_inputs, const luxb200_gate* gates, uint32_t n_gates,
                                        const uint32_t* outputs, uint32_t n_outputs, const uint64_t* luts, uint32_t n_luts,
                                        uint64_t beta, const uint64_t* in_a, const uint64_t* in_b, uint64_t* out_a,
                                        uint64_t* out_b, uint64_t count) {
    if (!c || !in_a || !in_b || (n_gates && !gates) || (n_outputs && (!outputs || !out_a || !out_b)) || (n_luts && !luts))
        return lb_fail(LUXB200_ERR_INVALID_ARG, "null argument");
    LB_LOCK(c);
    if (n_inputs == 0) return lb_fail(LUXB200_ERR_INVALID_ARG, "a circuit needs at least one input wire");
    if ((u64)n_inputs + 2ull * n_gates >= 0x7fffffffull) return lb_fail(LUXB200_ERR_INVALID_ARG, "too many wires");
    const u32 q = (u32)c->p.q, n = c->p.n;
    const u64 Q = c->p.Q;
    if ((u64)2 * q > (1ull << 31)) return lb_fail(LUXB200_ERR_UNSUPPORTED, "q too large for the u32 wire pool");

    // ---- host: expand the operations into bootstraps over internal wires -------------------------------------------------
    // tables: the six gate tables [6][q] first, then per function table.  f0 (the sign step of periodic / arbitrary functions)
    // is shared by all nodes of a kind.
    std::vector<uint64_t> tables;
    {
        std::vector<uint64_t> one;
        for (int g = LUXB200_OR; g <= LUXB200_XNOR; ++g) {
            lb_gate_lut(c, g, 4, one);
            tables.insert(tables.end(), one.begin(), one.end());
        }
    }
    std::vector<int> lut_kind(n_luts, -1);
    std::vector<u32> lut_toff(n_luts, 0);   // kind 0: T; kind 1 / 2: f2 table (f0 separate)
    u32 f0_toff[3] = {NONE, NONE, NONE};
    auto table_for = [&](u32 li) -> int {   // build on first use; returns the function kind or an error (< 0)
        if (lut_kind[li] >= 0) return lut_kind[li];
        const uint64_t* L = luts + (size_t)li * q;
        const int kind = check_input_function(L, q, q);
        if (kind == 2 && q > c->p.N) return -1;
        const u64 wm = kind == 2 ? 2ull * q : q, scale = Q / wm;
        for (u64 x = 0; x < q; ++x)
            if (L[x] >= wm) return -2;
        if (kind != 0 && f0_toff[kind] == NONE) {
            f0_toff[kind] = (u32)tables.size();
            for (u64 x = 0; x < wm; ++x) tables.push_back(scale * (x < (wm >> 1) ? wm - (wm >> 2) : (wm >> 2)));
        }
        lut_toff[li] = (u32)tables.size();
        for (u64 x = 0; x < wm; ++x) {
            u64 v;
            if (kind == 0) v = L[x];
            else if (kind == 1) v = x < (wm >> 1) ? L[x] : wm - L[x - wm / 2];
            else v = x < (wm >> 1) ? L[x % q] : wm - L[(x - wm / 2) % q];
            tables.push_back(scale * v);
        }
        lut_kind[li] = kind;
        return kind;
    };

    std::vector<Micro> mic;
    std::vector<u32> wire_of(n_gates);   // user gate -> internal wire of its result
    auto user_wire = [&](u32 w) -> u32 { return w < n_inputs ? w : wire_of[w - n_inputs]; };
    auto push = [&](const Micro& m) -> u32 { mic.push_back(m); return n_inputs + (u32)mic.size() - 1; };
    for (u32 g = 0; g < n_gates; ++g) {
        const luxb200_gate& G = gates[g];
        const u64 self = (u64)n_inputs + g;
        Micro m{};
        if (is_gate2(G.op)) {
            if (G.in0 >= self || G.in1 >= self) return lb_fail(LUXB200_ERR_INVALID_ARG, "Invalid ciphertext ID");
            if (G.in0 == G.in1) return lb_fail(LUXB200_ERR_INVALID_ARG, "Input ciphertexts should be independant");
            const bool dbl = G.op == LUXB200_XOR || G.op == LUXB200_XNOR || G.op == LUXB200_XOR_FAST || G.op == LUXB200_XNOR_FAST;
            m = Micro{0, user_wire(G.in0), user_wire(G.in1), dbl ? 2u : 1u, 0, table_of(G.op) * q, GRP_GATE, 0};
            wire_of[g] = push(m);
        } else if (G.op == LUXB200_OP_BOOTSTRAP) {
            if (G.in0 >= self) return lb_fail(LUXB200_ERR_INVALID_ARG, "Invalid ciphertext ID");
            // EvalAddConstEq(cct, q >> 2), binfhe-base-scheme.cpp:201
            m = Micro{0, user_wire(G.in0), NONE, 1u, q >> 2, table_of(G.op) * q, GRP_GATE, 0};
            wire_of[g] = push(m);
        } else if (G.op == LUXB200_OP_NOT) {
            if (G.in0 >= self) return lb_fail(LUXB200_ERR_INVALID_ARG, "Invalid ciphertext ID");
            m = Micro{1, user_wire(G.in0), NONE, 0, 0, 0, 0, 0};
            wire_of[g] = push(m);
        } else if (G.op == LUXB200_OP_EVALFUNC) {
            if (G.in0 >= self) return lb_fail(LUXB200_ERR_INVALID_ARG, "Invalid ciphertext ID");
            if (G.in1 >= n_luts) return lb_fail(LUXB200_ERR_INVALID_ARG, "gate %u names look-up table %u of %u", g, G.in1, n_luts);
            const int kind = table_for(G.in1);
            if (kind == -1)
                return lb_fail(LUXB200_ERR_INVALID_ARG,
                               "ERROR: ciphertext modulus q needs to be <= ring dimension for arbitrary function evaluation");
            if (kind < 0) return lb_fail(LUXB200_ERR_INVALID_ARG, "LUT value out of range for the plaintext space");
            const u32 w = user_wire(G.in0);
            if (kind == 0) {   // one bootstrap of ct + beta (:252-258)
                wire_of[g] = push(Micro{0, w, NONE, 1u, (u32)(beta % q), lut_toff[G.in1], GRP_FUNC, 0});
            } else {
                const u32 wm = kind == 2 ? 2 * q : q, grp = kind == 2 ? GRP_FUNC2Q : GRP_FUNC;
                // ct3 = BootstrapFunc(ct + beta, f0); then BootstrapFunc(ct - ct3 + beta - (q/2 | q/4), LUT') (:283-331)
                const u32 x = push(Micro{0, w, NONE, 1u, (u32)(beta % wm), f0_toff[kind], grp, 0});
                const u32 sub = kind == 2 ? (q >> 1) : (q >> 2);
                wire_of[g] = push(Micro{0, w, x, 1u | 16u, (u32)((beta % wm + wm - sub) % wm), lut_toff[G.in1], grp, kind == 2 ? 1u : 0u});
            }
        } else {
            return lb_fail(LUXB200_ERR_INVALID_ARG, "op %u of gate %u is not a circuit op", G.op, g);
        }
    }
    for (u32 k = 0; k < n_outputs; ++k)
        if (outputs[k] >= (u64)n_inputs + n_gates) return lb_fail(LUXB200_ERR_INVALID_ARG, "Invalid ciphertext ID");
    const u64 wires = (u64)n_inputs + mic.size();
    // level of a wire = 0 for inputs, 1 + max(level of operands) for a bootstrap, the operand's level for NOT
    std::vector<u32> level(wires, 0);
    u32 depth = 0, n_boot = 0;
    for (size_t k = 0; k < mic.size(); ++k) {
        const Micro& m = mic[k];
        u32 lv = level[m.in0];
        if (m.in1 != NONE) lv = std::max(lv, level[m.in1]);
        if (m.kind == 0) { ++lv; ++n_boot; }
        level[n_inputs + k] = lv;
        depth = std::max(depth, lv);
    }
    if (count == 0) return 0;
    if (n_boot && !c->have_bsk)
        return lb_fail(LUXB200_ERR_NO_KEYS, "Bootstrapping keys have not been generated. Please call BTKeyGen before calling bootstrapping.");
    if (n_boot && !c->have_ksk) return lb_fail(LUXB200_ERR_NO_KEYS, "SwitchingKey is empty");
    LB_TRY(lb_set_device(c->device));

    // slots: inputs first, then the bootstraps sorted by (level, launch group) — every (level, group) is one contiguous run
    // that the group's key switch writes in place — and the NOT wires behind them
    std::vector<u32> slot(wires);
    for (u32 i = 0; i < n_inputs; ++i) slot[i] = i;
    const u32 NB = (depth + 1) * GRP_COUNT;
    std::vector<u32> run_first(NB + 1, 0), run_cnt(NB, 0);
    for (size_t k = 0; k < mic.size(); ++k)
        if (mic[k].kind == 0) run_cnt[level[n_inputs + k] * GRP_COUNT + mic[k].grp]++;
    for (u32 r = 0; r < NB; ++r) run_first[r + 1] = run_first[r] + run_cnt[r];
    std::vector<u32> order(n_boot);
    {
        std::vector<u32> fill(run_first.begin(), run_first.end() - 1);
        for (size_t k = 0; k < mic.size(); ++k) {
            if (mic[k].kind != 0) continue;
            const u32 pos = fill[level[n_inputs + k] * GRP_COUNT + mic[k].grp]++;
            order[pos] = (u32)k;
            slot[n_inputs + k] = n_inputs + pos;
        }
    }
    u32 next = n_inputs + n_boot;
    for (size_t k = 0; k < mic.size(); ++k)
        if (mic[k].kind == 1) slot[n_inputs + k] = next++;
    std::vector<CircItem> desc(order.size());
    for (size_t k = 0; k < order.size(); ++k) {
        const Micro& m = mic[order[k]];
        desc[k] = CircItem{slot[m.in0], m.in1 == NONE ? NONE : slot[m.in1], m.flags, m.cst, m.toff, {0, 0, 0}};
    }

    // ---- device ----------------------------------------------------------------------------------------------------
    const u64 launches0 = c->launches;
    const u64 wave = lb_wave(c);
    const u64 CH = std::max<u64>(wave, 65536 / wave * wave);   // chunks of whole blind-rotation waves
    LB_TRY(c->cpa.ensure(wires * count * n * 4));
    LB_TRY(c->cpb.ensure(wires * count * 4));
    u32* pool_a = c->cpa.as<u32>();
    u32* pool_b = c->cpb.as<u32>();
    {   // inputs: u64 host -> u32 wire slots 0 .. n_inputs-1 (landing buffer in chunks of whole wires)
        const u64 per_wire = count * n;
        const u64 wires_per = std::max<u64>(1, (64ull << 20) / (per_wire * 8));
        LB_TRY(c->in0.ensure(std::min<u64>(wires_per, n_inputs) * per_wire * 8));
        LB_TRY(c->in1.ensure((u64)n_inputs * count * 8));
        for (u64 w0 = 0; w0 < n_inputs; w0 += wires_per) {
            const u64 m = std::min<u64>(wires_per, n_inputs - w0) * per_wire;
            LB_CUDA(cudaMemcpyAsync(c->in0.p, in_a + w0 * per_wire, m * 8, cudaMemcpyHostToDevice, c->stream));
            circuit_narrow_kernel<<<nblk(m, 256), 256, 0, c->stream>>>(c->in0.as<u64>(), pool_a + w0 * per_wire, q, m);
            c->launches++;
        }
        LB_CUDA(cudaMemcpyAsync(c->in1.p, in_b, (u64)n_inputs * count * 8, cudaMemcpyHostToDevice, c->stream));
        circuit_narrow_kernel<<<nblk((u64)n_inputs * count, 256), 256, 0, c->stream>>>(c->in1.as<u64>(), pool_b, q,
                                                                                      (u64)n_inputs * count);
        c->launches++;
        LB_CUDA(cudaGetLastError());
    }
    if (n_boot) {
        LB_TRY(lb_upload_lut(c, tables));
        LB_TRY(c->cdesc.ensure(desc.size() * sizeof(CircItem)));
        LB_CUDA(cudaMemcpyAsync(c->cdesc.p, desc.data(), desc.size() * sizeof(CircItem), cudaMemcpyHostToDevice, c->stream));
        LB_CUDA(cudaStreamSynchronize(c->stream));   // desc is a host temporary
        const u64 cap = std::min<u64>(CH, (u64)*std::max_element(run_cnt.begin(), run_cnt.end()) * count);
        LB_TRY(c->wa.ensure(cap * n * 4));
        LB_TRY(c->wb.ensure(cap * 4));
        LB_TRY(c->clsel.ensure(cap * 4));
    }
    std::vector<std::vector<u32>> nots(depth + 1);   // NOT wires by level, in gate order (a NOT may feed a NOT)
    for (size_t k = 0; k < mic.size(); ++k)
        if (mic[k].kind == 1) nots[level[n_inputs + k]].push_back((u32)k);
    auto run_nots = [&](u32 lv) -> int {
        for (u32 k : nots[lv]) {
            circuit_not_kernel<<<nblk(count * (n + 1), 256), 256, 0, c->stream>>>(
                pool_a, pool_b, (u64)slot[mic[k].in0] * count, (u64)slot[n_inputs + k] * count, n, q, count);
            c->launches++;
        }
        LB_CUDA(cudaGetLastError());
        return 0;
    };
    LB_TRY(run_nots(0));
    for (u32 lv = 1; lv <= depth; ++lv) {
        for (u32 grp = 0; grp < GRP_COUNT; ++grp) {
            const u32 r = lv * GRP_COUNT + grp;
            if (run_cnt[r] == 0) continue;
            const u64 m_in  = grp == GRP_FUNC2Q ? 2ull * q : q;             // modulus of the bootstrap's input and output
            const u64 b_add = grp == GRP_GATE ? Q / 8 + 1 : 0;              // (Q >> 3) + 1 = Q / (2 * 4) + 1 (:118, :214)
            const u64 items = (u64)run_cnt[r] * count;
            const u64 row0  = (u64)(n_inputs + run_first[r]) * count;
            for (u64 off = 0; off < items; off += CH) {
                const u64 m = std::min(CH, items - off);
                circuit_gather_kernel<<<nblk(m * (n + 1), 256), 256, 0, c->stream>>>(
                    pool_a, pool_b, c->cdesc.as<CircItem>() + run_first[r], c->wa.as<u32>(), c->wb.as<u32>(), c->clsel.as<u32>(), n,
                    (u32)m_in, count, off, m);
                c->launches++;
                LB_CUDA(cudaGetLastError());
                LB_TRY(lb_boot_dev(c, c->wa.as<u32>(), c->wb.as<u32>(), m_in, b_add, m_in, pool_a + (row0 + off) * n,
                                   pool_b + row0 + off, m, c->clsel.as<u32>()));
            }
            if (grp == GRP_FUNC2Q) {
                // the SECOND bootstrap of an arbitrary function ends with ct4->SetModulus(q); first ones stay mod 2q.  Runs of
                // this group mix both, so reduce item by item
                for (u32 k = 0; k < run_cnt[r]; ++k) {
                    if (!mic[order[run_first[r] + k]].mod_q) continue;
                    circuit_mod_kernel<<<nblk(count * (n + 1), 256), 256, 0, c->stream>>>(pool_a, pool_b, row0 + (u64)k * count, count, n, q);
                    c->launches++;
                }
                LB_CUDA(cudaGetLastError());
            }
        }
        LB_TRY(run_nots(lv));
    }
    for (u32 k = 0; k < n_outputs; ++k) {
        const u64 row = (u64)slot[user_wire(outputs[k])] * count;
        LB_TRY(lb_download_rows(c, pool_a + row * n, n, 0, n, count, c->o64a, out_a + (u64)k * count * n));
        LB_TRY(lb_download_rows(c, pool_b + row, 1, 0, 1, count, c->o64b, out_b + (u64)k * count));
    }
    LB_CUDA(cudaStreamSynchronize(c->stream));
    c->circ_stats[0] = depth;
    c->circ_stats[1] = n_boot;
    c->circ_stats[2] = c->launches - launches0;
    return 0;
}

extern "C" int luxb200_circuit_batch(luxb200_ctx* c, uint32_t n_inputs, const luxb200_gate* gates, uint32_t n_gates,
                                     const uint32_t* outputs, uint32_t n_outputs, const uint64_t* in_a, const uint64_t* in_b,
                                     uint64_t* out_a, uint64_t* out_b, uint64_t count) {
    return luxb200_circuit_batch_ex(c, n_inputs, gates, n_gates, outputs, n_outputs, nullptr, 0, 0, in_a, in_b, out_a, out_b, count);
}

extern "C" int luxb200_circuit_last_stats(luxb200_ctx* c, uint64_t* out) {
    if (!c || !out) return lb_fail(LUXB200_ERR_INVALID_ARG, "null argument");
    LB_LOCK(c);
    out[0] = c->circ_stats[0]; out[1] = c->circ_stats[1]; out[2] = c->circ_stats[2];
    return 0;
}
