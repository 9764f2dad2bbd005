#!/usr/bin/env python
"""Secondary workloads of BASELINE.json (configs[1], [3], [4]); one JSON line per workload.
  python bench_extra.py [ntt32] [ntt64] [methods]
Device-resident data, CUDA events on the library stream, >= 3 warm-ups; data sets exceed the 126 MB L2."""
import json
import os
import sys
import time

import numpy as np
import torch

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)
from fhe_b200 import keygen as kg, lib as _l  # noqa: E402
from fhe_b200.context import BinFHEContext, NTTPlan  # noqa: E402

PEAKS = {}
try:
    PEAKS = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
except Exception:
    pass
HBM = PEAKS.get("hbm_gbs", 6650.0)


def ntt32(N, bits, count, iters=10):
    L = _l.load()
    Q = L.luxb200_last_prime(bits, 2 * N)
    pl = NTTPlan(N, [Q])
    x = torch.randint(0, Q, (count, N), dtype=torch.int32, device="cuda")
    torch.cuda.synchronize()
    res = {}
    for inv in (0, 1):
        for _ in range(3):
            pl.SwitchFormatDev(x.data_ptr(), count, inv, 4)
        pl.Sync()
        ms = []
        for _ in range(iters):
            pl.SwitchFormatDev(x.data_ptr(), count, inv, 4)
            ms.append(pl.LastMs())
        t = float(np.median(ms)) * 1e-3
        bf = count * (N // 2) * (N.bit_length() - 1)
        gbs = count * N * 4 * 2 / t / 1e9
        res["inverse" if inv else "forward"] = {"ms": t * 1e3, "Gbutterfly_s": bf / t / 1e9, "GBs": gbs, "frac_hbm": gbs / HBM}
    pl.close()
    return {"workload": f"batched NTT u32 N={N} Q={bits}-bit x{count}", "bound": "hbm", "peak_gbs": HBM, **res}


def ntt64(N, Lm, cnt, bits=60, iters=5):
    L = _l.load()
    q = [L.luxb200_last_prime(bits, 2 * N)]
    for _ in range(Lm - 1):
        q.append(L.luxb200_previous_prime(q[-1], 2 * N))
    pl = NTTPlan(N, q)
    x = torch.randint(0, 2**59, (cnt, Lm, N), dtype=torch.int64, device="cuda")
    torch.cuda.synchronize()
    res = {}
    for inv in (0, 1):
        for _ in range(3):
            pl.SwitchFormatDev(x.data_ptr(), cnt, inv, 8)
        pl.Sync()
        ms = []
        for _ in range(iters):
            pl.SwitchFormatDev(x.data_ptr(), cnt, inv, 8)
            ms.append(pl.LastMs())
        t = float(np.median(ms)) * 1e-3
        bf = cnt * Lm * (N // 2) * (N.bit_length() - 1)
        gbs = cnt * Lm * N * 8 * 2 / t / 1e9
        res["inverse" if inv else "forward"] = {"ms": t * 1e3, "Gbutterfly_s": bf / t / 1e9, "algorithmic_GBs": gbs,
                                                 "frac_hbm_algorithmic": gbs / HBM}
    pl.close()
    import ctypes
    pk = ctypes.c_double(0)
    L.luxb200_intpipe_probe64(0, ctypes.byref(pk))
    for k in res:
        res[k]["frac_intpipe"] = res[k]["Gbutterfly_s"] * 1e9 / pk.value if pk.value else None
    return {"workload": f"DCRT NTT u64 N={N} L={Lm} x{cnt} ({bits}-bit primes)",
            "intpipe_peak_Gbutterfly_s": pk.value / 1e9, **res}


def method_gates(name, method, G, fn="gate"):
    ctx = BinFHEContext().GenerateBinFHEContext(name, method)
    ks = kg.keygen(ctx, seed=77)
    ctx.BTKeyLoad(ks.bsk, ks.kskA, ks.kskB)
    rng = np.random.default_rng(5)
    out = {"workload": f"{name}/{method} {fn} x{G}"}
    if fn == "gate":
        m1 = rng.integers(0, 2, G); m2 = rng.integers(0, 2, G)
        a1, b1 = kg.encrypt(ctx, ks.sk, m1, rng); a2, b2 = kg.encrypt(ctx, ks.sk, m2, rng)
        ctx.EvalBinGateBatch("AND", (a1[:64], b1[:64]), (a2[:64], b2[:64]))
        t0 = time.perf_counter()
        (oa, ob), res = ctx.EvalBinGateBatch("AND", (a1, b1), (a2, b2))
        dt = time.perf_counter() - t0
        assert res.success, res.error
        out["decrypt_errors"] = int((kg.decrypt(ctx, ks.sk, oa, ob) != (m1 & m2)).sum())
        out["bootstraps_per_s_e2e"] = G / dt
        out.update(ctx.LastTiming())
    else:  # EvalFunc, periodic LUT (2 bootstraps per call), p = q / 256
        p = max(2, ctx.q // 256)
        lut = ctx.GenerateLUTviaFunction(lambda m, pp: m % (pp // 2) if pp > 2 else 0, p)
        m = rng.integers(0, p, G)
        a, b = kg.encrypt(ctx, ks.sk, m, rng, p=p)
        ctx.EvalFuncBatch((a[:64], b[:64]), lut)
        t0 = time.perf_counter()
        (oa, ob), res = ctx.EvalFuncBatch((a, b), lut)
        dt = time.perf_counter() - t0
        assert res.success, res.error
        exp = np.array([(int(v) % (p // 2)) if p > 2 else 0 for v in m])
        out["decrypt_errors"] = int((kg.decrypt(ctx, ks.sk, oa, ob, p=p) != exp).sum())
        out["evalfunc_per_s_e2e"] = G / dt
        out["bootstraps_per_s_e2e"] = 2 * G / dt
    ctx.close()
    return out


def compare8(name, method, pairs):
    """BASELINE configs[3]: the 8-bit GreaterThan8 circuit of go/tfhe/compare.go:106-139 over `pairs` independent
    comparisons, level-batched on the device: one batch of 24*pairs carry-independent gates (a AND NOT b, NOT a AND b,
    a XNOR b per bit), then 3 dependent levels per bit (24 batches of `pairs` gates); 48 bootstrapped gates per
    comparison.  Ciphertexts stay in HBM as uint32; EvalNOT (binfhe-base-scheme.cpp:222-237) is the linear map
    a -> -a, b -> q/4 - b.  Bit-exactness of this circuit against the reference is checked by host/host_parity.cpp;
    here the result is decrypted and compared with the same circuit on the plaintext bits."""
    ctx = BinFHEContext().GenerateBinFHEContext(name, method)
    ks = kg.keygen(ctx, seed=31)
    ctx.BTKeyLoad(ks.bsk, ks.kskA, ks.kskB)
    rng = np.random.default_rng(8)
    q, n = ctx.q, ctx.n
    va = rng.integers(0, 256, pairs); vb = np.where(rng.integers(0, 2, pairs) == 1, rng.integers(0, 256, pairs), va ^ (1 << rng.integers(0, 8, pairs)))
    bits_a = ((va[:, None] >> np.arange(8)) & 1).reshape(-1)          # [pairs*8], index p*8 + i
    bits_b = ((vb[:, None] >> np.arange(8)) & 1).reshape(-1)

    def dev(m):
        a, b = kg.encrypt(ctx, ks.sk, m, rng)
        return (torch.from_numpy(a.astype(np.uint32).view(np.int32)).cuda(), torch.from_numpy(b.astype(np.uint32).view(np.int32)).cuda())

    def NOT(ct):
        a, b = ct
        return (torch.where(a == 0, a, q - a), torch.remainder(q // 4 - b, q))

    def gate(g, x, y):
        cnt = x[1].numel()
        oa = torch.empty((cnt, n), dtype=torch.int32, device="cuda"); ob = torch.empty(cnt, dtype=torch.int32, device="cuda")
        xa, xb, ya, yb = x[0].contiguous(), x[1].contiguous(), y[0].contiguous(), y[1].contiguous()
        ctx.EvalBinGateBatchDev(g, xa.data_ptr(), xb.data_ptr(), ya.data_ptr(), yb.data_ptr(), oa.data_ptr(), ob.data_ptr(), cnt)
        ctx.Synchronize()   # inputs of this level are temporaries
        return (oa, ob)

    A, B, init = dev(bits_a), dev(bits_b), dev(np.zeros(pairs, np.int64))

    def circuit():
        aG, bG, eq = gate("AND", A, NOT(B)), gate("AND", NOT(A), B), gate("XNOR", A, B)
        res = init
        for i in range(7, -1, -1):
            sel = torch.arange(pairs, device="cuda") * 8 + i
            e, g, nb = (eq[0][sel], eq[1][sel]), (aG[0][sel], aG[1][sel]), NOT((bG[0][sel], bG[1][sel]))
            res = gate("AND", e, res)
            res = gate("OR", g, res)
            res = gate("AND", res, nb)
        return res

    circuit()   # warm-up
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    res = circuit()
    torch.cuda.synchronize()
    dt = time.perf_counter() - t0
    oa = res[0].cpu().numpy().view(np.uint32).astype(np.uint64); ob = res[1].cpu().numpy().view(np.uint32).astype(np.uint64)
    dec = kg.decrypt(ctx, ks.sk, oa, ob)
    exp = np.zeros(pairs, np.int64)
    for p in range(pairs):
        r = 0
        for i in range(7, -1, -1):
            ai, bi = (int(va[p]) >> i) & 1, (int(vb[p]) >> i) & 1
            r = (ai & (1 - bi)) | (int(ai == bi) & r)
            r = r & (1 - ((1 - ai) & bi))
        exp[p] = r
    ctx.close()
    return {"workload": f"{name}/{method} GreaterThan8 circuit x{pairs} (48 gates each, 25 batched levels)",
            "comparisons_per_s": pairs / dt, "bootstraps_per_s": 48 * pairs / dt, "seconds": dt,
            "decrypt_errors": int((dec != exp).sum())}


def compare8_circuit(name, method, pairs):
    """The same GreaterThan8 circuit through luxb200_circuit_batch (BatchDAG on the device): one call with HOST buffers;
    the 17 input wires go up, 25 levels run back to back on the context stream with every intermediate in HBM, one wire
    comes back.  Timed end to end (H2D + D2H inside)."""
    from fhe_b200.circuits import greater_than8, plain_circuit
    ctx = BinFHEContext().GenerateBinFHEContext(name, method)
    ks = kg.keygen(ctx, seed=31)
    ctx.BTKeyLoad(ks.bsk, ks.kskA, ks.kskB)
    rng = np.random.default_rng(8)
    va = rng.integers(0, 256, pairs); vb = np.where(rng.integers(0, 2, pairs) == 1, rng.integers(0, 256, pairs), va ^ (1 << rng.integers(0, 8, pairs)))
    bits = np.concatenate([((va[None, :] >> np.arange(8)[:, None]) & 1), ((vb[None, :] >> np.arange(8)[:, None]) & 1),
                           np.zeros((1, pairs), np.int64)])
    in_a = torch.empty((17, pairs, ctx.n), dtype=torch.int64).pin_memory().numpy().view(np.uint64)
    in_b = torch.empty((17, pairs), dtype=torch.int64).pin_memory().numpy().view(np.uint64)
    for i in range(17):
        in_a[i], in_b[i] = kg.encrypt(ctx, ks.sk, bits[i], rng)
    gates, out = greater_than8()
    ctx.EvalCircuitBatch(17, gates, [out], in_a[:, :64], in_b[:, :64])   # warm-up
    t0 = time.perf_counter()
    (oa, ob), res = ctx.EvalCircuitBatch(17, gates, [out], in_a, in_b)
    dt = time.perf_counter() - t0
    assert res.success, res.error
    st = ctx.CircuitStats()
    dec = kg.decrypt(ctx, ks.sk, oa[0], ob[0])
    exp = plain_circuit(gates, bits)[out]
    ctx.close()
    return {"workload": f"{name}/{method} GreaterThan8 x{pairs} as ONE luxb200_circuit_batch call (host buffers)",
            "comparisons_per_s": pairs / dt, "bootstraps_per_s": 48 * pairs / dt, "seconds": dt, "levels": st["levels"],
            "launches": st["launches"], "decrypt_errors": int((dec != exp).sum())}


if __name__ == "__main__":
    which = sys.argv[1:] or ["ntt32", "ntt64", "methods", "compare"]
    if "ntt32" in which:
        print(json.dumps(ntt32(1024, 27, 65536)))
        print(json.dumps(ntt32(2048, 29, 65536)))
    if "ntt64" in which:
        print(json.dumps(ntt64(65536, 32, 64)))
        print(json.dumps(ntt64(16384, 16, 64)))
        print(json.dumps(ntt64(4096, 8, 512)))
    if "compare" in which:
        print(json.dumps(compare8("STD128", "GINX", 4096)))
        print(json.dumps(compare8("STD128_LMKCDEY", "LMKCDEY", 4096)))
    if "circuit" in which:
        print(json.dumps(compare8_circuit("STD128", "GINX", 4096)))
        print(json.dumps(compare8_circuit("STD128_LMKCDEY", "LMKCDEY", 4096)))
    if "methods" in which:
        print(json.dumps(method_gates("STD128_LMKCDEY", "LMKCDEY", 16384)))
        print(json.dumps(method_gates("STD128_AP", "AP", 4096)))
        print(json.dumps(method_gates("STD128_LMKCDEY", "LMKCDEY", 16384, "func")))
        print(json.dumps(method_gates("STD128_AP", "AP", 4096, "func")))
