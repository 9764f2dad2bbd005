"""One device-resident hybrid key switch batch (N = 2^16, 24 + 8 limbs, 3 digits, 32 polynomials) for ncu:
  ncu --metrics gpu__time_duration.sum --clock-control none --csv --log-file gpurun_out/hks_launches.csv python tools/prof_hks.py"""
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
from fhe_b200.context import HybridKeySwitchPlan  # noqa: E402
from orclib import P, oracle  # noqa: E402

O = oracle()
N, sizeQ, sizeP, alpha, count = 65536, 24, 8, 8, 32
chain = np.zeros(sizeQ + sizeP, np.uint64); psi = chain.copy()
O.orc_dcrt_chain(N, sizeQ + sizeP, 60, P(chain), P(psi))
q, p = np.ascontiguousarray(chain[:sizeQ]), np.ascontiguousarray(chain[sizeQ:])
pl = HybridKeySwitchPlan(N, q, p, alpha, sizeQ)
nd = pl.num_digits
dev = torch.device("cuda:0")
g = torch.Generator(device="cpu").manual_seed(1)
d_c = torch.randint(0, 1 << 59, (count, sizeQ, N), dtype=torch.int64, generator=g).to(dev)
d_kb = torch.randint(0, 1 << 59, (nd, sizeQ + sizeP, N), dtype=torch.int64, generator=g).to(dev)
d_ka = torch.randint(0, 1 << 59, (nd, sizeQ + sizeP, N), dtype=torch.int64, generator=g).to(dev)
d_out = torch.empty((count, 2, sizeQ, N), dtype=torch.int64, device=dev)
torch.cuda.synchronize()
for _ in range(int(sys.argv[1]) if len(sys.argv) > 1 else 2):
    pl.KeySwitchDev(d_c.data_ptr(), d_kb.data_ptr(), d_ka.data_ptr(), d_out.data_ptr(), count, 0)
print("done")
