"""One batch of STD256 (N = 2048, 29-bit Q) or STD192 (37-bit Q) GINX gates through brw_kernel, for ncu:
  ncu --set full --clock-control none --import-source on -k regex:brw_kernel -c 1 -o gpurun_out/wide python tools/prof_wide.py STD256 444"""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from fhe_b200.context import BinFHEContext  # noqa: E402

name = sys.argv[1] if len(sys.argv) > 1 else "STD256"
G = int(sys.argv[2]) if len(sys.argv) > 2 else 444
method = "LMKCDEY" if name.endswith("LMKCDEY") else "GINX"
rng = np.random.default_rng(1)
ctx = BinFHEContext().GenerateBinFHEContext(name, method)
bsk = rng.integers(0, ctx.Q, size=ctx.bsk_len(), dtype=np.uint64)
A = rng.integers(0, ctx.qKS, size=ctx.ksk_rows() * ctx.n, dtype=np.uint64)
B = rng.integers(0, ctx.qKS, size=ctx.ksk_rows(), dtype=np.uint64)
ctx.BTKeyLoad(bsk, A, B)
a1 = rng.integers(0, ctx.q, size=(G, ctx.n), dtype=np.uint64); b1 = rng.integers(0, ctx.q, size=G, dtype=np.uint64)
a2 = rng.integers(0, ctx.q, size=(G, ctx.n), dtype=np.uint64); b2 = rng.integers(0, ctx.q, size=G, dtype=np.uint64)
(oa, ob), res = ctx.EvalBinGateBatch("AND", (a1, b1), (a2, b2))
assert res.success, res.error
t = ctx.LastTiming()
print({"set": name, "gates": G, "br_ms": t["br_ms"], "bootstraps_per_s": G / (t["total_ms"] * 1e-3)})
ctx.close()
