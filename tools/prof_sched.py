"""Small driver for ncu: one batch of LMKCDEY or AP gates (python tools/prof_sched.py LMKCDEY|AP [gates])."""
import os
import sys

import numpy as np

sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), ".."))
from fhe_b200 import keygen as kg  # noqa: E402
from fhe_b200.context import BinFHEContext  # noqa: E402

method = sys.argv[1] if len(sys.argv) > 1 else "LMKCDEY"
G = int(sys.argv[2]) if len(sys.argv) > 2 else 1776
name = {"LMKCDEY": "STD128_LMKCDEY", "AP": "STD128_AP", "GINX": "STD128"}[method]
ctx = BinFHEContext().GenerateBinFHEContext(name, method)
ks = kg.keygen(ctx, seed=31)
ctx.BTKeyLoad(ks.bsk, ks.kskA, ks.kskB)
rng = np.random.default_rng(8)
m1, m2 = rng.integers(0, 2, G), rng.integers(0, 2, G)
a1, b1 = kg.encrypt(ctx, ks.sk, m1, rng)
a2, b2 = kg.encrypt(ctx, ks.sk, m2, rng)
for it in range(3):
    (oa, ob), res = ctx.EvalBinGateBatch("AND", (a1, b1), (a2, b2))
    assert res.success, res.error
    print(it, ctx.LastTiming())
print("decrypt errors", int((kg.decrypt(ctx, ks.sk, oa, ob) != (m1 & m2)).sum()))
ctx.close()
