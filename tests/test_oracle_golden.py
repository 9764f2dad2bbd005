"""CPU: the plain-C oracle against the golden vectors the REAL reference produced (tests/golden/make_golden.py).
This is what pins the oracle where /root/reference is absent (e.g. on the GPU box)."""
import numpy as np
import pytest

from conftest import GOLD, gold_u64, load_gold
from orclib import GATE, METHOD, OrcContext, P, hand_params, oracle
import os


def ctx_from_gold(d, method):
    c = d["custom"]
    O = oracle()
    Q = O.orc_last_prime(c["bits"], 2 * c["N"])
    oc = OrcContext(hand_params(c["n"], c["N"], c["q"], Q, c["qKS"], c["baseKS"], c["baseG"], c["baseR"],
                                c["numAutoKeys"], METHOD[method]))
    oc.load(gold_u64(d["bsk"]), gold_u64(d["kskA"]), gold_u64(d["kskB"]))
    return oc


def test_primes_and_roots():
    O = oracle()
    z = np.load(os.path.join(GOLD, "chain.npz"))
    for bits, m, Q, psi, prev in z["primes"]:
        assert O.orc_last_prime(int(bits), int(m)) == Q
        assert O.orc_root_of_unity(int(m), int(Q)) == psi
        assert O.orc_previous_prime(int(Q), int(m)) == prev
    q = np.zeros(32, np.uint64); psi = np.zeros(32, np.uint64)
    O.orc_dcrt_chain(65536, 32, 60, P(q), P(psi))
    assert (q == z["dcrt_q"]).all() and (psi == z["dcrt_psi"]).all()


def test_ntt_golden():
    O = oracle()
    z = np.load(os.path.join(GOLD, "ntt.npz"))
    for key in [k for k in z.files if k.startswith("q_")]:
        _, bits, N = key.split("_")
        Q, psi = [int(v) for v in z[key]]
        x, X = z[f"x_{bits}_{N}"], z[f"X_{bits}_{N}"]
        y = x.copy()
        O.orc_ntt(Q, psi, int(N), P(y), len(y), 0, 2)
        assert (y == X).all(), key
        O.orc_ntt(Q, psi, int(N), P(y), len(y), 1, 2)
        assert (y == x).all(), key


@pytest.mark.parametrize("method", ["GINX", "AP", "LMKCDEY"])
def test_binfhe_golden(method):
    d = load_gold(method)
    oc = ctx_from_gold(d, method)
    O, p = oracle(), oc.p
    a1, b1, a2, b2 = gold_u64(d["a1"]), gold_u64(d["b1"]), gold_u64(d["a2"]), gold_u64(d["b2"])
    cnt = len(b1)
    for g in ["AND", "OR", "NAND", "NOR", "XOR", "XNOR"]:
        for ext in (0, 1):
            xa, xb = oc.evalbingate(g, a1, b1, a2, b2, bool(ext))
            assert (xa == d[f"{g}_{ext}_a"]).all() and (xb == d[f"{g}_{ext}_b"]).all(), (g, ext)
    oa = np.zeros((cnt, p.n), np.uint64); ob = np.zeros(cnt, np.uint64)
    O.orc_bootstrap(oc.h, P(a1), P(b1), P(oa), P(ob), cnt, 4, 0, 4)
    assert (oa == d["BOOT_a"]).all() and (ob == d["BOOT_b"]).all()
    for g, k, pm in [("AND3", 3, 6), ("OR3", 3, 6), ("MAJORITY", 3, 4), ("AND4", 4, 8), ("OR4", 4, 8), ("CMUX", 3, 4)]:
        am, bm = gold_u64(d[f"{g}_in_a"]), gold_u64(d[f"{g}_in_b"])
        O.orc_evalbingate_multi(oc.h, GATE[g], k, P(am), P(bm), P(oa), P(ob), cnt, pm, 4)
        assert (oa == d[f"{g}_a"]).all() and (ob == d[f"{g}_b"]).all(), g
    fa, fb = gold_u64(d["F_in_a"]), gold_u64(d["F_in_b"])
    for name in ["neg", "per", "arb"]:
        lut = gold_u64(d[f"F_{name}_lut"])
        rc = O.orc_evalfunc(oc.h, P(fa), P(fb), P(lut), len(lut), 128, P(oa), P(ob), cnt, p.q, 4)
        assert rc == int(d[f"F_{name}_rc"][0])
        if rc == 0:
            assert (oa == d[f"F_{name}_a"]).all() and (ob == d[f"F_{name}_b"]).all(), name
    acc = gold_u64(d["ACC_in"]).copy()
    O.orc_evalacc(oc.h, P(a1[0].copy()), p.q, P(acc))
    assert (acc == d["ACC_out"]).all()
    dec = np.zeros(p.rows * p.N, np.uint64)
    O.orc_decompose(oc.h, P(gold_u64(d["ACC_in"])), P(dec))
    assert (dec == d["DEC_out"]).all()
    ka, kb = gold_u64(d["KS_in_a"]), gold_u64(d["KS_in_b"])
    for i in range(2):
        ra = np.zeros(p.n, np.uint64); rb = np.zeros(1, np.uint64)
        O.orc_keyswitch(oc.h, P(ka[i].copy()), int(kb[i]), P(ra), P(rb))
        assert (ra == d["KS_a"][i]).all() and rb[0] == d["KS_b"][i]
    ms = np.zeros(p.N, np.uint64)
    O.orc_modswitch(p.qKS, p.Q, P(gold_u64(d["MS_in"][0])), P(ms), p.N)
    assert (ms == d["MS_out"][0]).all()
    oc.close()
