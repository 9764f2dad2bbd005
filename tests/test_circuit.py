"""Gate circuits (luxb200_circuit_batch = lux::fhe::BatchDAG, binfhe-batch.h:255-279, level-scheduled on the device with
device-resident intermediates) against the oracle evaluated gate by gate in BatchDAG's own order
(src/binfhe/lib/batch/batch.cpp:429-482), bit for bit on every wire."""
import numpy as np
import pytest

from fhe_b200.circuits import greater_than8, plain_circuit
from orclib import METHOD, OrcContext, P, hand_params

GATES2 = ["OR", "AND", "NOR", "NAND", "XOR", "XNOR"]


def oracle_twin(ctx, method):
    return OrcContext(hand_params(ctx.n, ctx.N, ctx.q, ctx.Q, ctx.qKS, ctx.baseKS, ctx.baseG, ctx.baseR,
                                  ctx.numAutoKeys, METHOD[method]))


def random_circuit(rng, n_inputs, n_gates):
    """Random DAG in BatchDAG id order: every 2-input gate kind, EvalNOT (also of inputs and of NOTs) and Bootstrap."""
    gates = []
    for g in range(n_gates):
        avail = n_inputs + g
        r = rng.integers(0, 10)
        if r == 0:
            gates.append(("NOT", int(rng.integers(0, avail)), None))
        elif r == 1:
            gates.append(("BOOTSTRAP", int(rng.integers(0, avail)), None))
        else:
            i0, i1 = rng.choice(avail, 2, replace=False)
            gates.append((GATES2[int(rng.integers(0, 6))], int(i0), int(i1)))
    return gates


def oracle_circuit(oc, q, gates, in_a, in_b):
    """Wire by wire, one reference operation per gate (the order BatchDAG::Execute uses on the CPU)."""
    wa = [in_a[i] for i in range(in_a.shape[0])]
    wb = [in_b[i] for i in range(in_b.shape[0])]
    cnt = in_b.shape[1]
    for op, i0, i1 in gates:
        if op == "NOT":     # EvalNOT, binfhe-base-scheme.cpp:223-237
            a = np.where(wa[i0] == 0, 0, q - wa[i0]).astype(np.uint64)
            b = ((q // 4 + q - wb[i0]) % q).astype(np.uint64)
        elif op == "BOOTSTRAP":
            a = np.zeros_like(wa[i0]); b = np.zeros(cnt, np.uint64)
            oc.L.orc_bootstrap(oc.h, P(np.ascontiguousarray(wa[i0])), P(np.ascontiguousarray(wb[i0])), P(a), P(b), cnt, 4, 0, 4)
        else:
            a, b = oc.evalbingate(op, np.ascontiguousarray(wa[i0]), np.ascontiguousarray(wb[i0]),
                                  np.ascontiguousarray(wa[i1]), np.ascontiguousarray(wb[i1]))
        wa.append(a); wb.append(b)
    return wa, wb


@pytest.mark.gpu
@pytest.mark.parametrize("name,method,n_gates,cnt", [("TOY", "GINX", 40, 5), ("TOY", "AP", 16, 3), ("TOY", "LMKCDEY", 16, 3),
                                                     ("STD128", "GINX", 10, 2),
                                                     ("STD256", "GINX", 6, 2)])   # N = 2048: the four-warp kernel under the scheduler
def test_circuit_vs_oracle_every_wire(name, method, n_gates, cnt):
    from fhe_b200.context import BinFHEContext
    ctx = BinFHEContext().GenerateBinFHEContext(name, method)
    oc = oracle_twin(ctx, method)
    sk, bsk, A, B = oc.keygen(seed=7)
    ctx.BTKeyLoad(bsk, A, B)
    rng = np.random.default_rng(11)
    n_inputs = 4
    bits = rng.integers(0, 2, (n_inputs, cnt))
    in_a = np.zeros((n_inputs, cnt, ctx.n), np.uint64); in_b = np.zeros((n_inputs, cnt), np.uint64)
    for i in range(n_inputs):
        for k in range(cnt):
            in_a[i, k], in_b[i, k] = oc.encrypt(sk, int(bits[i, k]), 100 * i + k)
    gates = random_circuit(rng, n_inputs, n_gates - 3)
    # always present: NOT of an input, NOT of a NOT, Bootstrap of a gate result
    gates += [("NOT", 1, None), ("NOT", n_inputs + n_gates - 3, None), ("BOOTSTRAP", n_inputs + n_gates - 4, None)]
    outs = list(range(n_inputs + n_gates))          # every wire, the inputs included
    (oa, ob), res = ctx.EvalCircuitBatch(n_inputs, gates, outs, in_a, in_b)
    assert res.success, res.error
    xa, xb = oracle_circuit(oc, ctx.q, gates, in_a, in_b)
    for w in outs:
        assert (oa[w] == xa[w]).all() and (ob[w] == xb[w]).all(), (name, method, w, gates[w - n_inputs] if w >= n_inputs else "input")
    # and the circuit computes what it says (inputs are real encryptions)
    exp = plain_circuit(gates, bits)
    for w in outs:
        dec = [oc.decrypt(sk, oa[w, k], int(ob[w, k])) for k in range(cnt)]
        assert dec == [int(x) for x in exp[w]], (w, dec)
    st = ctx.CircuitStats()
    assert st["bootstrapped_gates"] == sum(g[0] != "NOT" for g in gates)
    assert st["levels"] <= st["bootstrapped_gates"]
    ctx.close(); oc.close()


@pytest.mark.gpu
def test_greater_than8_circuit_std128():
    """The 8-bit comparison circuit on STD128 GINX: circuit result == the same gates issued level by level through
    luxb200_evalbingate_batch (itself oracle-checked), and decrypts to what the circuit gives on the plaintext bits."""
    from fhe_b200.context import BinFHEContext
    from fhe_b200 import keygen as kg
    ctx = BinFHEContext().GenerateBinFHEContext("STD128", "GINX")
    ks = kg.keygen(ctx, seed=3)
    ctx.BTKeyLoad(ks.bsk, ks.kskA, ks.kskB)
    rng = np.random.default_rng(5)
    pairs = 24
    va = rng.integers(0, 256, pairs); vb = np.where(rng.integers(0, 2, pairs) == 1, rng.integers(0, 256, pairs), va)
    bits = np.concatenate([((va[None, :] >> np.arange(8)[:, None]) & 1), ((vb[None, :] >> np.arange(8)[:, None]) & 1),
                           np.zeros((1, pairs), np.int64)])
    in_a = np.zeros((17, pairs, ctx.n), np.uint64); in_b = np.zeros((17, pairs), np.uint64)
    for i in range(17):
        in_a[i], in_b[i] = kg.encrypt(ctx, ks.sk, bits[i], rng)
    gates, out = greater_than8()
    (oa, ob), res = ctx.EvalCircuitBatch(17, gates, [out], in_a, in_b)
    assert res.success, res.error
    st = ctx.CircuitStats()
    assert st["bootstrapped_gates"] == 48 and st["levels"] == 25
    assert (kg.decrypt(ctx, ks.sk, oa[0], ob[0]) == plain_circuit(gates, bits)[out]).all()
    # the same gates one batched call each
    q = ctx.q
    wa = [in_a[i] for i in range(17)]; wb = [in_b[i] for i in range(17)]
    for op, i0, i1 in gates:
        if op == "NOT":
            wa.append(np.where(wa[i0] == 0, 0, q - wa[i0]).astype(np.uint64)); wb.append(((q // 4 + q - wb[i0]) % q).astype(np.uint64))
        else:
            (a, b), r = ctx.EvalBinGateBatch(op, (wa[i0], wb[i0]), (wa[i1], wb[i1]))
            assert r.success
            wa.append(a); wb.append(b)
    assert (oa[0] == wa[out]).all() and (ob[0] == wb[out]).all()
    ctx.close()


@pytest.mark.gpu
def test_circuit_error_behaviour():
    """BatchDAG::Execute's errors (batch.cpp:429-482: 'Invalid ciphertext ID') and EvalBinGate's
    ('Input ciphertexts should be independant', binfhe-base-scheme.cpp:85-86; no keys, :533)."""
    from fhe_b200.context import BinFHEContext
    ctx = BinFHEContext().GenerateBinFHEContext("TOY", "GINX")
    a = np.zeros((2, 3, ctx.n), np.uint64); b = np.zeros((2, 3), np.uint64)
    (_, _), res = ctx.EvalCircuitBatch(2, [("AND", 0, 1)], [2], a, b)
    assert not res.success and "Bootstrapping keys" in res.error
    (_, _), res = ctx.EvalCircuitBatch(2, [("AND", 0, 0)], [2], a, b)
    assert not res.success and "independant" in res.error
    (_, _), res = ctx.EvalCircuitBatch(2, [("AND", 0, 2)], [2], a, b)
    assert not res.success and "Invalid ciphertext ID" in res.error
    (_, _), res = ctx.EvalCircuitBatch(2, [("AND", 0, 1)], [3], a, b)
    assert not res.success and "Invalid ciphertext ID" in res.error
    (_, _), res = ctx.EvalCircuitBatch(2, [("AND3", 0, 1)], [2], a, b)
    assert not res.success
    # NOT-only circuits need no keys; an empty batch is a no-op success
    (oa, ob), res = ctx.EvalCircuitBatch(2, [("NOT", 1, None)], [2, 0], a + 5, b + 7)
    assert res.success and (oa[0] == ctx.q - 5).all() and (ob[0] == (ctx.q // 4 + ctx.q - 7) % ctx.q).all()
    assert (oa[1] == 5).all() and (ob[1] == 7).all()
    e = np.zeros((2, 0, ctx.n), np.uint64); eb = np.zeros((2, 0), np.uint64)
    (oa, ob), res = ctx.EvalCircuitBatch(2, [("AND", 0, 1)], [2], e, eb)
    assert res.success and oa.shape[1] == 0
    ctx.close()


def test_circuit_helpers_cpu():
    """host logic of this file (no GPU): the comparison circuit has 48 bootstrapped gates and, evaluated on
    plaintext bits, is the ripple the Go code spells out (the last bit visited, bit 0, decides first)."""
    gates, out = greater_than8()
    assert sum(g[0] != "NOT" for g in gates) == 48 and out == 17 + len(gates) - 1
    rng = np.random.default_rng(1)
    va = rng.integers(0, 256, 200); vb = np.where(rng.integers(0, 2, 200) == 1, rng.integers(0, 256, 200), va)
    bits = np.concatenate([((va[None, :] >> np.arange(8)[:, None]) & 1), ((vb[None, :] >> np.arange(8)[:, None]) & 1),
                           np.zeros((1, 200), np.int64)])
    rev = lambda v: sum(((v >> i) & 1) << (7 - i) for i in range(8))
    assert (plain_circuit(gates, bits)[out] == (rev(va) > rev(vb))).all()


@pytest.mark.gpu
@pytest.mark.parametrize("name,method,cnt", [("TOY", "GINX", 4), ("TOY", "LMKCDEY", 3), ("STD128", "GINX", 2), ("STD128_LMKCDEY", "LMKCDEY", 2)])
def test_circuit_with_evalfunc_nodes(name, method, cnt):
    """BatchDAG with OP_EVAL_FUNC nodes (batch.cpp:455-460) mixed with gates, NOT and Bootstrap as ONE device call: a
    negacyclic table (1 bootstrap), a periodic one (2 chained bootstraps) and — where q <= N — an arbitrary one (2 bootstraps
    at modulus 2q), feeding each other and gates; every wire against the oracle applied node by node
    (BinFHEScheme::EvalFunc, binfhe-base-scheme.cpp:241-332).  An arbitrary table with q > N is refused like the reference."""
    from fhe_b200.context import BinFHEContext
    ctx = BinFHEContext().GenerateBinFHEContext(name, method)
    oc = oracle_twin(ctx, method)
    sk, bsk, A, B = oc.keygen(seed=9, key_dist=getattr(ctx, "keyDist", 0))
    ctx.BTKeyLoad(bsk, A, B)
    q, p = ctx.q, 4
    neg = ctx.GenerateLUTviaFunction(lambda x, pp: x + 1 if x < pp // 2 else pp - (x - pp // 2 + 1), p)
    per = ctx.GenerateLUTviaFunction(lambda x, pp: x % (pp // 2), p)
    arb = ctx.GenerateLUTviaFunction(lambda x, pp: (x * x * x + 1) % pp, p)   # 1, 2, 1, 0: neither negacyclic nor periodic
    luts = [neg, per] + ([arb] if q <= ctx.N else [])
    n_inputs = 3
    in_a = np.zeros((n_inputs, cnt, ctx.n), np.uint64); in_b = np.zeros((n_inputs, cnt), np.uint64)
    for i in range(n_inputs):
        for k in range(cnt):
            in_a[i, k], in_b[i, k] = oc.encrypt(sk, (i + k) % p, 100 * i + k, pmod=p)
    # wires: 0 1 2 inputs; 3 = f_neg(0); 4 = f_per(1); 5 = AND(3, 4); 6 = f_per(5); 7 = NOT 6; 8 = XOR(7, 2); 9 = BOOT(4);
    # 10 = f_neg(4) (a function of a function); [11 = f_arb(2); 12 = f_per(11); 13 = NAND(11, 0)]
    gates = [("EVALFUNC", 0, 0), ("EVALFUNC", 1, 1), ("AND", 3, 4), ("EVALFUNC", 5, 1), ("NOT", 6, None), ("XOR", 7, 2),
             ("BOOTSTRAP", 4, None), ("EVALFUNC", 4, 0)]
    if len(luts) == 3:
        gates += [("EVALFUNC", 2, 2), ("EVALFUNC", 11, 1), ("NAND", 11, 0)]
    outs = list(range(n_inputs + len(gates)))
    (oa, ob), res = ctx.EvalCircuitBatch(n_inputs, gates, outs, in_a, in_b, luts)
    assert res.success, res.error
    wa = [in_a[i] for i in range(n_inputs)]; wb = [in_b[i] for i in range(n_inputs)]
    for op, i0, i1 in gates:
        a0, b0 = np.ascontiguousarray(wa[i0]), np.ascontiguousarray(wb[i0])
        if op == "EVALFUNC":
            lut64 = np.ascontiguousarray(luts[i1], np.uint64)
            a = np.zeros_like(a0); b = np.zeros(cnt, np.uint64)
            assert oc.L.orc_evalfunc(oc.h, P(a0), P(b0), P(lut64), len(lut64), ctx.GetBeta(), P(a), P(b), cnt, q, 4) == 0
        elif op == "NOT":
            a = np.where(a0 == 0, 0, q - a0).astype(np.uint64); b = ((q // 4 + q - b0) % q).astype(np.uint64)
        elif op == "BOOTSTRAP":
            a = np.zeros_like(a0); b = np.zeros(cnt, np.uint64)
            oc.L.orc_bootstrap(oc.h, P(a0), P(b0), P(a), P(b), cnt, 4, 0, 4)
        else:
            a, b = oc.evalbingate(op, a0, b0, np.ascontiguousarray(wa[i1]), np.ascontiguousarray(wb[i1]))
        wa.append(a); wb.append(b)
    for w in outs:
        assert (oa[w] == wa[w]).all() and (ob[w] == wb[w]).all(), (name, method, w)
    st = ctx.CircuitStats()
    assert st["bootstrapped_gates"] == sum({"EVALFUNC": None, "NOT": 0}.get(g[0], 1) if g[0] != "EVALFUNC" else (1 if g[2] == 0 else 2) for g in gates)
    if q > ctx.N:   # arbitrary table: refused with the reference's message (binfhe-base-scheme.cpp:264-268)
        (_, _), res = ctx.EvalCircuitBatch(n_inputs, [("EVALFUNC", 0, 0)], [3], in_a, in_b, [arb])
        assert not res.success and "ring dimension" in res.error
    ctx.close(); oc.close()
