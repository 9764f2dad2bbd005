// circuit.cu — luxb200_circuit_batch: a gate DAG (lux::fhe::BatchDAG, binfhe-batch.h:255-279) evaluated level by level on
// the device.  The reference's BatchDAG::Execute (src/binfhe/lib/batch/batch.cpp:429-482) walks the operations one by one
// on the CPU; here the host only sorts the gates into dependency levels, and every level is three launches over all of its
// (gate, input set) pairs — mixed gate kinds included, through a per-item test-vector table offset (BrParams::lutsel):
//   circuit_gather_kernel  ct = mul * (wire[in0] + wire[in1]) (+ q/4 for Bootstrap)     binfhe-base-scheme.cpp:99-107, :201
//   blind rotation         BootstrapGateCore + EvalAcc + extraction                     :525-583, :111-121
//   key switch             SwitchCTtoqn, written straight into the level's wire slots    lwe-pke.cpp:170-178
// Intermediate ciphertexts never leave HBM (u32 wire pool [wires][count][n] + [wires][count]).
#include <algorithm>

#include "api_common.h"

typedef uint32_t u32;
typedef uint64_t u64;

namespace {

inline unsigned nblk(u64 total, unsigned per) { return (unsigned)((total + per - 1) / per); }

// one bootstrapped gate of a level: x = source slot 0, y = source slot 1 (0xffffffff: none), z = mul | table index << 8,
// w = constant added to b
__global__ void circuit_gather_kernel(const u32* __restrict__ pool_a, const u32* __restrict__ pool_b,
                                      const uint4* __restrict__ desc, u32* __restrict__ wa, u32* __restrict__ wb,
                                      u32* __restrict__ lutsel, u32 n, u32 q, u64 count, u64 item0, u64 items) {
    const u64 idx = (u64)blockIdx.x * blockDim.x + threadIdx.x;
    const u64 per = (u64)n + 1;
    if (idx >= items * per) return;
    const u64 it  = idx / per;
    const u32 t   = (u32)(idx % per);
    const u64 git = item0 + it;
    const uint4 d = desc[git / count];
    const u64 i   = git % count;
    const u64 r0  = (u64)d.x * count + i;
    const u32 mul = d.z & 0xffu;
    u32 s;
    if (t < n) {
        s = pool_a[r0 * n + t];
        if (d.y != 0xffffffffu) s += pool_a[((u64)d.y * count + i) * n + t];
    } else {
        s = pool_b[r0];
        if (d.y != 0xffffffffu) s += pool_b[(u64)d.y * count + i];
    }
    s = (u32)(((u64)(s % q) * mul) % q);
    if (t < n) {
        wa[it * n + t] = s;
    } else {
        wb[it]     = (s + d.w) % q;
        lutsel[it] = (d.z >> 8) * q;
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

// table index of an op inside the uploaded [6][q] test-vector tables (OR, AND, NOR, NAND, XOR, XNOR)
u32 table_of(u32 op) {
    if (op == LUXB200_XOR_FAST) return LUXB200_XOR;
    if (op == LUXB200_XNOR_FAST) return LUXB200_XNOR;
    if (op == LUXB200_OP_BOOTSTRAP) return LUXB200_AND;   // BootstrapGateCore(params, AND, ...), binfhe-base-scheme.cpp:205
    return op;
}

}  // namespace

extern "C" int luxb200_circuit_batch(luxb200_ctx* c, uint32_t n_inputs, const luxb200_gate* gates, uint32_t n_gates,
                                     const uint32_t* outputs, uint32_t n_outputs, const uint64_t* in_a, const uint64_t* in_b,
                                     uint64_t* out_a, uint64_t* out_b, uint64_t count) {
    if (!c || !in_a || !in_b || (n_gates && !gates) || (n_outputs && (!outputs || !out_a || !out_b)))
        return lb_fail(LUXB200_ERR_INVALID_ARG, "null argument");
    LB_LOCK(c);
    if (n_inputs == 0) return lb_fail(LUXB200_ERR_INVALID_ARG, "a circuit needs at least one input wire");
    const u64 wires = (u64)n_inputs + n_gates;
    if (wires >= 0xffffffffull) return lb_fail(LUXB200_ERR_INVALID_ARG, "too many wires");
    // ---- host: validate, levels, slots -----------------------------------------------------------------------------
    // level of a wire = 0 for inputs, 1 + max(level of operands) for a bootstrapped gate, the operand's level for NOT
    std::vector<u32> level(wires, 0);
    u32 depth = 0, n_boot = 0;
    for (u32 g = 0; g < n_gates; ++g) {
        const luxb200_gate& G = gates[g];
        const u64 self        = (u64)n_inputs + g;
        if (is_gate2(G.op)) {
            if (G.in0 >= self || G.in1 >= self) return lb_fail(LUXB200_ERR_INVALID_ARG, "Invalid ciphertext ID");
            if (G.in0 == G.in1) return lb_fail(LUXB200_ERR_INVALID_ARG, "Input ciphertexts should be independant");
            level[self] = 1 + std::max(level[G.in0], level[G.in1]);
            ++n_boot;
        } else if (G.op == LUXB200_OP_BOOTSTRAP) {
            if (G.in0 >= self) return lb_fail(LUXB200_ERR_INVALID_ARG, "Invalid ciphertext ID");
            level[self] = 1 + level[G.in0];
            ++n_boot;
        } else if (G.op == LUXB200_OP_NOT) {
            if (G.in0 >= self) return lb_fail(LUXB200_ERR_INVALID_ARG, "Invalid ciphertext ID");
            level[self] = level[G.in0];
        } else {
            return lb_fail(LUXB200_ERR_INVALID_ARG, "op %u of gate %u is not a circuit op", G.op, g);
        }
        depth = std::max(depth, level[self]);
    }
    for (u32 k = 0; k < n_outputs; ++k)
        if (outputs[k] >= wires) return lb_fail(LUXB200_ERR_INVALID_ARG, "Invalid ciphertext ID");
    if (count == 0) return 0;
    if (n_boot && !c->have_bsk)
        return lb_fail(LUXB200_ERR_NO_KEYS, "Bootstrapping keys have not been generated. Please call BTKeyGen before calling bootstrapping.");
    if (n_boot && !c->have_ksk) return lb_fail(LUXB200_ERR_NO_KEYS, "SwitchingKey is empty");
    LB_TRY(lb_set_device(c->device));

    // slots: inputs first, then each level's bootstrapped gates as one contiguous run (the key switch of the level writes
    // the run in place), NOT wires behind them
    std::vector<u32> slot(wires);
    std::vector<u32> lv_first(depth + 3, 0), lv_cnt(depth + 3, 0);
    for (u32 i = 0; i < n_inputs; ++i) slot[i] = i;
    std::vector<u32> order(n_boot);   // bootstrapped gates in level order (counting sort, stable in gate order)
    for (u32 g = 0; g < n_gates; ++g)
        if (gates[g].op != LUXB200_OP_NOT) lv_cnt[level[n_inputs + g]]++;
    for (u32 lv = 1; lv <= depth; ++lv) lv_first[lv + 1] = lv_first[lv] + lv_cnt[lv];
    {
        std::vector<u32> fill(lv_first);
        for (u32 g = 0; g < n_gates; ++g) {
            if (gates[g].op == LUXB200_OP_NOT) continue;
            const u32 k = fill[level[n_inputs + g]]++;
            order[k]           = g;
            slot[n_inputs + g] = n_inputs + k;
        }
    }
    u32 next = n_inputs + n_boot;
    for (u32 g = 0; g < n_gates; ++g)
        if (gates[g].op == LUXB200_OP_NOT) slot[n_inputs + g] = next++;
    const u32 q = (u32)c->p.q, n = c->p.n;
    std::vector<uint4> desc(order.size());
    for (size_t k = 0; k < order.size(); ++k) {
        const luxb200_gate& G = gates[order[k]];
        const bool two        = is_gate2(G.op);
        const bool dbl        = G.op == LUXB200_XOR || G.op == LUXB200_XNOR || G.op == LUXB200_XOR_FAST || G.op == LUXB200_XNOR_FAST;
        desc[k].x = slot[G.in0];
        desc[k].y = two ? slot[G.in1] : 0xffffffffu;
        desc[k].z = (dbl ? 2u : 1u) | (table_of(G.op) << 8);
        desc[k].w = two ? 0u : (q >> 2);   // EvalAddConstEq(cct, q >> 2), binfhe-base-scheme.cpp:201
    }

    // ---- device ----------------------------------------------------------------------------------------------------
    const u64 launches0 = c->launches;
    const u64 wave = lb_wave(c);
    const u64 CH   = std::max<u64>(wave, 65536 / wave * wave);   // chunks of whole blind-rotation waves
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
        std::vector<uint64_t> luts, one;
        for (int g = LUXB200_OR; g <= LUXB200_XNOR; ++g) {
            lb_gate_lut(c, g, 4, one);
            luts.insert(luts.end(), one.begin(), one.end());
        }
        LB_TRY(lb_upload_lut(c, luts));
        LB_TRY(c->cdesc.ensure(desc.size() * sizeof(uint4)));
        LB_CUDA(cudaMemcpyAsync(c->cdesc.p, desc.data(), desc.size() * sizeof(uint4), cudaMemcpyHostToDevice, c->stream));
        const u64 cap = std::min<u64>(CH, (u64)*std::max_element(lv_cnt.begin(), lv_cnt.end()) * count);
        LB_TRY(c->wa.ensure(cap * n * 4));
        LB_TRY(c->wb.ensure(cap * 4));
        LB_TRY(c->clsel.ensure(cap * 4));
    }
    std::vector<std::vector<u32>> nots(depth + 1);   // NOT gates by level, in gate order (a NOT may feed a NOT)
    for (u32 g = 0; g < n_gates; ++g)
        if (gates[g].op == LUXB200_OP_NOT) nots[level[n_inputs + g]].push_back(g);
    auto run_nots = [&](u32 lv) -> int {
        for (u32 g : nots[lv]) {
            circuit_not_kernel<<<nblk(count * (n + 1), 256), 256, 0, c->stream>>>(
                pool_a, pool_b, (u64)slot[gates[g].in0] * count, (u64)slot[n_inputs + g] * count, n, q, count);
            c->launches++;
        }
        LB_CUDA(cudaGetLastError());
        return 0;
    };
    LB_TRY(run_nots(0));
    const u64 b_add = c->p.Q / 8 + 1;   // (Q >> 3) + 1 = Q / (2 * 4) + 1: gates and Bootstrap with p = 4 alike
    for (u32 lv = 1; lv <= depth; ++lv) {
        if (lv_cnt[lv] == 0) continue;
        const u64 items = (u64)lv_cnt[lv] * count;
        const u64 row0  = (u64)slot[n_inputs + order[lv_first[lv]]] * count;
        for (u64 off = 0; off < items; off += CH) {
            const u64 m = std::min(CH, items - off);
            circuit_gather_kernel<<<nblk(m * (n + 1), 256), 256, 0, c->stream>>>(
                pool_a, pool_b, c->cdesc.as<uint4>() + lv_first[lv], c->wa.as<u32>(), c->wb.as<u32>(), c->clsel.as<u32>(), n, q,
                count, off, m);
            c->launches++;
            LB_CUDA(cudaGetLastError());
            LB_TRY(lb_boot_dev(c, c->wa.as<u32>(), c->wb.as<u32>(), q, b_add, q, pool_a + (row0 + off) * n, pool_b + row0 + off,
                               m, c->clsel.as<u32>()));
        }
        LB_TRY(run_nots(lv));
    }
    for (u32 k = 0; k < n_outputs; ++k) {
        const u64 row = (u64)slot[outputs[k]] * count;
        LB_TRY(lb_download_rows(c, pool_a + row * n, n, 0, n, count, c->o64a, out_a + (u64)k * count * n));
        LB_TRY(lb_download_rows(c, pool_b + row, 1, 0, 1, count, c->o64b, out_b + (u64)k * count));
    }
    LB_CUDA(cudaStreamSynchronize(c->stream));
    c->circ_stats[0] = depth;
    c->circ_stats[1] = n_boot;
    c->circ_stats[2] = c->launches - launches0;
    return 0;
}

extern "C" int luxb200_circuit_last_stats(luxb200_ctx* c, uint64_t* out) {
    if (!c || !out) return lb_fail(LUXB200_ERR_INVALID_ARG, "null argument");
    LB_LOCK(c);
    out[0] = c->circ_stats[0]; out[1] = c->circ_stats[1]; out[2] = c->circ_stats[2];
    return 0;
}
