"""A/B of the scheduled blind-rotation kernel (AP / LMKCDEY) between library builds: LUXB200_SO=<path> python tools/ab_sched.py"""
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench  # noqa: E402
from fhe_b200 import lib as _l  # noqa: E402

L = _l.load()
for name, method, G in (("STD128_AP", "AP", 3552), ("STD128_LMKCDEY", "LMKCDEY", 14208)):
    r = bench.method_metric(L, 0, None, name, method, G, 18.56e12)
    print(json.dumps({"so": os.environ.get("LUXB200_SO", "main"), "set": name, "gates": G, "value": r["value"],
                      "kernel_ms": r["roofline"]["kernel_ms"], "frac": r["roofline"]["frac"], "errors": r["decrypt_errors"]}))
