"""Small driver for ncu: a few forward+inverse DCRT NTTs at N=2^16, L=32 (python tools/prof_ntt64.py [N] [L] [cnt])."""
import os
import sys

import torch

sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), ".."))
from fhe_b200 import lib as _l  # noqa: E402
from fhe_b200.context import NTTPlan  # noqa: E402

N = int(sys.argv[1]) if len(sys.argv) > 1 else 65536
Lm = int(sys.argv[2]) if len(sys.argv) > 2 else 32
cnt = int(sys.argv[3]) if len(sys.argv) > 3 else 64
L = _l.load()
q = [L.luxb200_last_prime(60, 2 * N)]
for _ in range(Lm - 1):
    q.append(L.luxb200_previous_prime(q[-1], 2 * N))
pl = NTTPlan(N, q)
x = torch.randint(0, 2**59, (cnt, Lm, N), dtype=torch.int64, device="cuda")
for it in range(3):
    for inv in (0, 1):
        pl.SwitchFormatDev(x.data_ptr(), cnt, inv, 8)
        print(it, inv, pl.LastMs())
pl.close()
