import os
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
GOLD = os.path.join(ROOT, "tests", "golden")


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box with -m gpu)")


def load_gold(method):
    """Golden vectors written by tests/golden/make_golden.py from the real reference."""
    d = dict(np.load(os.path.join(GOLD, f"binfhe_{method.lower()}.npz")))
    bits, cyc, n, q, modKS, baseKS, baseG, baseR, nauto, keyDist = [int(x) for x in d["params"]]
    d["custom"] = dict(bits=bits, N=cyc // 2, n=n, q=q, qKS=modKS, baseKS=baseKS, baseG=baseG, baseR=baseR,
                       numAutoKeys=nauto)
    return d


def gold_u64(x):
    return np.ascontiguousarray(x, dtype=np.uint64)


@pytest.fixture(scope="session")
def orc():
    import orclib
    return orclib


def has_gpu():
    try:
        from fhe_b200 import lib
        return lib.load().luxb200_device_count() > 0
    except Exception:
        return False
