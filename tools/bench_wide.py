"""Throughput of the shared-memory blind rotation (br_wide.cuh) on the N = 2048 / 64-bit parameter sets: random keys and
inputs (throughput does not depend on their values), device-side kernel times from luxb200_last_timing.
  python tools/bench_wide.py [gates]"""
import json
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from fhe_b200.context import BinFHEContext  # noqa: E402

G = int(sys.argv[1]) if len(sys.argv) > 1 and sys.argv[1].isdigit() else 1184
rng = np.random.default_rng(1)
SETS = [("STD128_3", "GINX"), ("STD128Q", "GINX"), ("STD128_4_LMKCDEY", "LMKCDEY")] if "--n1024" in sys.argv else None
for name, method in SETS or [("STD256", "GINX"), ("STD256_3", "GINX"), ("STD256Q_3", "GINX"), ("STD192", "GINX"), ("STD192Q_3", "GINX"), ("STD128Q_3", "GINX"),
                     ("STD192_LMKCDEY", "LMKCDEY"), ("STD256_LMKCDEY", "LMKCDEY")]:
    ctx = BinFHEContext().GenerateBinFHEContext(name, method)
    bsk = rng.integers(0, ctx.Q, size=ctx.bsk_len(), dtype=np.uint64)
    A = rng.integers(0, ctx.qKS, size=ctx.ksk_rows() * ctx.n, dtype=np.uint64)
    B = rng.integers(0, ctx.qKS, size=ctx.ksk_rows(), dtype=np.uint64)
    ctx.BTKeyLoad(bsk, A, B)
    a1 = rng.integers(0, ctx.q, size=(G, ctx.n), dtype=np.uint64); b1 = rng.integers(0, ctx.q, size=G, dtype=np.uint64)
    a2 = rng.integers(0, ctx.q, size=(G, ctx.n), dtype=np.uint64); b2 = rng.integers(0, ctx.q, size=G, dtype=np.uint64)
    ctx.EvalBinGateBatch("AND", (a1[:148], b1[:148]), (a2[:148], b2[:148]))
    (oa, ob), res = ctx.EvalBinGateBatch("AND", (a1, b1), (a2, b2))
    assert res.success, res.error
    t = ctx.LastTiming()
    print(json.dumps({"set": name, "method": method, "n": ctx.n, "N": ctx.N, "Q_bits": int(ctx.Q).bit_length(), "digits": ctx.digitsG - 1,
                      "gates": G, "br_ms": t["br_ms"], "ks_ms": t["ks_ms"], "bootstraps_per_s": G / (t["total_ms"] * 1e-3)}))
    ctx.close()
