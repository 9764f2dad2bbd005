"""Small driver for ncu: batched u32 NTT (python tools/prof_ntt32.py [N] [bits] [count])."""
import os
import sys

import torch

sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), ".."))
from fhe_b200 import lib as _l  # noqa: E402
from fhe_b200.context import NTTPlan  # noqa: E402

N = int(sys.argv[1]) if len(sys.argv) > 1 else 1024
bits = int(sys.argv[2]) if len(sys.argv) > 2 else 27
cnt = int(sys.argv[3]) if len(sys.argv) > 3 else 65536
L = _l.load()
Q = L.luxb200_last_prime(bits, 2 * N)
pl = NTTPlan(N, [Q])
x = torch.randint(0, Q, (cnt, N), dtype=torch.int32, device="cuda")
for it in range(4):
    for inv in (0, 1):
        pl.SwitchFormatDev(x.data_ptr(), cnt, inv, 4)
        print(it, inv, pl.LastMs())
pl.close()
