"""Small batches for compute-sanitizer (one tool per run): TOY GINX / LMKCDEY / AP gates + a STD128 GINX wave (the staged-key
kernel), checked against the oracle so the run also proves the sanitized execution is the correct one.
  compute-sanitizer --tool racecheck python tools/sanitize_toy.py [GINX|LMKCDEY|AP|STD128]"""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
from orclib import METHOD, OrcContext, hand_params  # noqa: E402
from fhe_b200.context import BinFHEContext  # noqa: E402

which = sys.argv[1] if len(sys.argv) > 1 else "GINX"
name, method, cnt = ("STD128", "GINX", 7) if which == "STD128" else ("TOY", which, 13)
ctx = BinFHEContext().GenerateBinFHEContext(name, method)
oc = OrcContext(hand_params(ctx.n, ctx.N, ctx.q, ctx.Q, ctx.qKS, ctx.baseKS, ctx.baseG, ctx.baseR, ctx.numAutoKeys, METHOD[method]))
sk, bsk, A, B = oc.keygen(seed=7)
ctx.BTKeyLoad(bsk, A, B)
a1 = np.zeros((cnt, ctx.n), np.uint64); b1 = np.zeros(cnt, np.uint64); a2 = a1.copy(); b2 = b1.copy()
for i in range(cnt):
    a1[i], b1[i] = oc.encrypt(sk, i & 1, 10 + i); a2[i], b2[i] = oc.encrypt(sk, (i >> 1) & 1, 50 + i)
(oa, ob), res = ctx.EvalBinGateBatch("NAND", (a1, b1), (a2, b2))
assert res.success, res.error
xa, xb = oc.evalbingate("NAND", a1, b1, a2, b2)
assert (oa == xa).all() and (ob == xb).all()
print(f"sanitize_toy {name}/{method}: {cnt} gates bit-exact vs oracle")
