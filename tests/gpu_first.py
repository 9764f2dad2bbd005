"""First GPU parity run (scratch driver; the real tests are tests/test_*.py)."""
import sys
import time

import numpy as np

sys.path.insert(0, "tests")
sys.path.insert(0, ".")
from orclib import OrcContext, hand_params, oracle, P  # noqa: E402

from fhe_b200.context import BinFHEContext, NTTPlan, PARAM_SETS  # noqa: E402
from fhe_b200 import lib as L  # noqa: E402


def orc_for(ctx, method):
    p = hand_params(ctx.n, ctx.N, ctx.q, ctx.Q, ctx.qKS, ctx.baseKS, ctx.baseG, ctx.baseR, ctx.numAutoKeys,
                    L.METHOD[method])
    return OrcContext(p)


def main():
    lib = L.load()
    print("devices", lib.luxb200_device_count())
    O = oracle()
    rng = np.random.default_rng(5)
    # ---- NTT parity
    for bits, N in [(27, 1024), (27, 512), (28, 1024), (29, 2048), (27, 256), (25, 1024)]:
        Q = lib.luxb200_last_prime(bits, 2 * N)
        psi = lib.luxb200_root_of_unity(2 * N, Q)
        pl = NTTPlan(N, [Q])
        x = rng.integers(0, Q, size=(64, N), dtype=np.uint64)
        ref = x.copy()
        O.orc_ntt(Q, psi, N, P(ref), 64, 0, 4)
        got = pl.SwitchFormat(x, 0)
        okf = (got == ref).all()
        back = pl.SwitchFormat(got, 1)
        print("NTT", bits, N, "fwd", okf, "inv", (back == x).all())
        pl.close()
    # ---- gates
    for name, method, cnt in [("TOY", "GINX", 16), ("STD128", "GINX", 8)]:
        ctx = BinFHEContext().GenerateBinFHEContext(name, method)
        oc = orc_for(ctx, method)
        t = time.time()
        sk, bsk, A, B = oc.keygen(seed=7, key_dist=0)
        print(name, "oracle keygen", time.time() - t)
        ctx.BTKeyLoad(bsk, A, B)
        a1 = np.zeros((cnt, ctx.n), np.uint64); b1 = np.zeros(cnt, np.uint64)
        a2 = a1.copy(); b2 = b1.copy()
        for i in range(cnt):
            a1[i], b1[i] = oc.encrypt(sk, i & 1, 100 + i)
            a2[i], b2[i] = oc.encrypt(sk, (i >> 1) & 1, 200 + i)
        # raw blind rotation
        acc = rng.integers(0, ctx.Q, size=(2, 2, ctx.N), dtype=np.uint64)
        exp = acc.copy()
        for i in range(2):
            O.orc_evalacc(oc.h, P(a1[i]), ctx.q, P(exp[i]))
        got = ctx.BlindRotateBatch(a1[:2], acc)
        print(name, "BlindRotateBatch exact:", (got == exp).all())
        for g in ["AND", "XOR"]:
            for ext in (True, False):
                t = time.time()
                xa, xb = oc.evalbingate(g, a1, b1, a2, b2, ext)
                t1 = time.time() - t
                (oa, ob), res = ctx.EvalBinGateBatch(g, (a1, b1), (a2, b2), ext)
                ok = res.success and (oa == xa).all() and (ob == xb).all()
                print(name, g, "ext" if ext else "std", "exact:", ok, res.error, "oracle s", round(t1, 2),
                      ctx.LastTiming())
                if not ok and res.success:
                    bad = np.argwhere(oa != xa)
                    print("  mismatches", len(bad), bad[:5], "b ok", (ob == xb).all())
        dec = [oc.decrypt(sk, oa[i], ob[i]) for i in range(cnt)]
        print(name, "decrypt XOR:", dec)
        ctx.close(); oc.close()


if __name__ == "__main__":
    main()
