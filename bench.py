#!/usr/bin/env python
"""bench.py — headline benchmark: batched STD128 GINX EvalBinGate (1 bootstrap per gate) on N B200s.

  python bench.py [--gpus N] [--steps K] [--warmup W] [--gates G] [--impl ours|reference]
  torchrun ... bench.py --gpus N ...      (one rank per GPU; keys broadcast once with NCCL, no per-gate collective)

A "step" = one pass of the hot path over one batch of G synthetic gates per GPU (default 2^16, BASELINE.json
configs[2]).  Prints ONE JSON line (see DESIGN.md "Measurement").
"""
import argparse
import ctypes
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

# algorithmic integer work per STD128 GINX gate: 3021 transforms x 5120 butterflies x 3 IMAD
# + 503 x 20 x 1024 pointwise modmuls x 4 IMAD
IMAD_PER_GATE = 3021 * 5120 * 3 + 503 * 20 * 1024 * 4


def env_int(k, d):
    return int(os.environ.get(k, d))


class ClockSampler:
    """nvidia-smi clocks / throttle reasons during the timed region (B200_PROFILING.md recipe)."""
    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")

    def __init__(self, gpu):
        self.gpu, self.proc, self.lines = gpu, None, []

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", f"--id={self.gpu}", f"--query-gpu={self.Q}",
                                          "--format=csv,noheader,nounits", "-lms", "200"], stdout=subprocess.PIPE,
                                         stderr=subprocess.DEVNULL, text=True)
            self.th = threading.Thread(target=self._read, daemon=True)
            self.th.start()
        except Exception:
            self.proc = None

    def _read(self):
        for ln in self.proc.stdout:
            self.lines.append(ln.strip())

    def stop(self):
        if not self.proc:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        self.proc.terminate()
        try:
            self.proc.wait(timeout=2)
        except Exception:
            pass
        sm, mx, reasons = [], None, set()
        for ln in self.lines:
            f = [x.strip() for x in ln.split(",")]
            if len(f) < 9:
                continue
            try:
                sm.append(float(f[1])); mx = float(f[2])
            except ValueError:
                continue
            for name, v in zip(["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"], f[5:9]):
                if v.lower().startswith("active"):
                    reasons.add(name)
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": mx, "reasons": sorted(reasons),
                "samples": len(sm)}


def load_checker():
    """CPU baseline library: the real reference (oracle/_ref/libref_shim.so) if present, else the C port."""
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    import orclib
    if orclib.ref_available():
        try:
            return "reference", orclib
        except OSError:
            pass
    return "port", orclib


def cpu_time_gates(kind, orclib, gates, threads):
    """Time `gates` STD128 GINX AND gates on the host cores. Returns seconds."""
    P = orclib.P
    rng = np.random.default_rng(11)
    if kind == "reference":
        rc = orclib.RefContext("STD128", "GINX")
        rc.keygen()
        p = rc.p
        a1 = np.zeros((gates, p.n), np.uint64); b1 = np.zeros(gates, np.uint64); a2 = a1.copy(); b2 = b1.copy()
        for i in range(gates):
            a1[i], b1[i] = rc.encrypt(1); a2[i], b2[i] = rc.encrypt(i & 1)
        oa = np.zeros((gates, p.n), np.uint64); ob = np.zeros(gates, np.uint64)
        fn = lambda: orclib.ref().ref_time_bingate(rc.h, 1, P(a1), P(b1), P(a2), P(b2), P(oa), P(ob), gates, threads)
    else:
        hp = orclib.hand_params(503, 1024, 1024, 134215681, 16384, 32, 512, 32, 10, 2)
        oc = orclib.OrcContext(hp)
        sk, _, _, _ = oc.keygen(seed=3)
        a1 = np.zeros((gates, 503), np.uint64); b1 = np.zeros(gates, np.uint64); a2 = a1.copy(); b2 = b1.copy()
        for i in range(gates):
            a1[i], b1[i] = oc.encrypt(sk, 1, 10 + i); a2[i], b2[i] = oc.encrypt(sk, i & 1, 1000 + i)
        oa = np.zeros((gates, 503), np.uint64); ob = np.zeros(gates, np.uint64)
        fn = lambda: orclib.oracle().orc_time_bingate(oc.h, 1, P(a1), P(b1), P(a2), P(b2), P(oa), P(ob), gates, threads)
    return fn


def run_reference(args, rank, world):
    if rank != 0:
        return
    kind, orclib = load_checker()
    cores = os.cpu_count() or 1
    gates = args.ref_gates or cores * 16
    fn = cpu_time_gates(kind, orclib, gates, cores)
    for _ in range(args.warmup):
        fn()
    t = [fn() for _ in range(args.steps)]
    tot = sum(t)
    val = gates * args.steps / tot
    sample = f"{gates} STD128 GINX AND gates per step, OpenMP over gates, {cores} threads"
    print(json.dumps({
        "impl": "reference", "metric": "STD128 GINX bootstraps/sec", "value": val, "unit": "bootstraps/s",
        "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup, "ms_per_step": 1e3 * tot / args.steps,
        "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "u64", "data": "synthetic",
        "config": {"workload": "binfhe STD128 GINX EvalBinGate(AND) batch", "gates_per_step": gates, "device": "cpu"},
        "cpu_baseline": {"value": val, "unit": "bootstraps/s", "cores": cores, "kind": kind, "sample": sample},
        "e2e": {"value": val, "unit": "bootstraps/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0}))


def ntt_metric(lib, local, dev, world, barrier, N=65536, Lm=32, cnt=32, iters=5, rank0_cpu=False):
    """Batched forward / inverse DCRTPoly NTT (N=2^16, 32 limbs of 60 bits, `cnt` ciphertexts per GPU = 512 MB,
    larger than L2), device-resident, CUDA events on the library stream, max over ranks.  Roofline: the 64-bit
    butterfly rate of luxb200_intpipe_probe64 (integer pipe) and the algorithmic 2*N*8 bytes per limb-NTT (HBM)."""
    import torch
    import torch.distributed as dist
    from fhe_b200.context import NTTPlan
    q = [lib.luxb200_last_prime(60, 2 * N)]
    for _ in range(Lm - 1):
        q.append(lib.luxb200_previous_prime(q[-1], 2 * N))
    pl = NTTPlan(N, q, device=local)
    x = torch.randint(0, 2**59, (cnt, Lm, N), dtype=torch.int64, device=dev)
    x0 = x.clone()
    torch.cuda.synchronize(dev)
    out = {"workload": f"DCRT NTT u64 N={N} L={Lm} x{cnt} per GPU (60-bit primes, {cnt * Lm * N * 8 >> 20} MB > L2)"}
    bf = cnt * Lm * (N // 2) * (N.bit_length() - 1)
    for inv in (0, 1):
        for _ in range(3):
            pl.SwitchFormatDev(x.data_ptr(), cnt, inv, 8)
        pl.Sync()
        barrier()
        ms = []
        for _ in range(iters):
            pl.SwitchFormatDev(x.data_ptr(), cnt, inv, 8)
            ms.append(pl.LastMs())
        t = float(np.median(ms)) * 1e-3
        if world > 1:
            tt = torch.tensor([t], dtype=torch.float64, device=dev)
            dist.all_reduce(tt, op=dist.ReduceOp.MAX)
            t = float(tt.item())
        out["inverse" if inv else "forward"] = {"ms": t * 1e3, "Gbutterfly_s": world * bf / t / 1e9,
                                                 "algorithmic_GBs_per_gpu": cnt * Lm * N * 16 / t / 1e9}
    # parity of the timed buffer: 8 forward then 8 inverse passes must give the input back, and one more forward pass of
    # ciphertext 0 is compared limb by limb with the oracle (NumberTheoreticTransformNat restatement)
    ok = bool((x == x0).all().item())
    kind, orclib = load_checker()
    O = orclib.oracle()
    qa = np.array(q, np.uint64); psi = np.zeros(Lm, np.uint64); q2 = np.zeros(Lm, np.uint64)
    O.orc_dcrt_chain(N, Lm, 60, orclib.P(q2), orclib.P(psi))
    ok = ok and bool((qa == q2).all())
    y = x0[:1].clone()
    ref = y.cpu().numpy().view(np.uint64).copy()
    pl.SwitchFormatDev(y.data_ptr(), 1, 0, 8); pl.Sync()
    for l in range(Lm):
        O.orc_ntt(int(q2[l]), int(psi[l]), N, orclib.P(ref[0, l]), 1, 0, 1)
    ok = ok and bool((y.cpu().numpy().view(np.uint64) == ref).all())
    out["parity"] = ok
    pl.close()
    pk = ctypes.c_double(0)
    lib.luxb200_intpipe_probe64(local, ctypes.byref(pk))
    imad, _ = _probe_imad(lib, local)
    alg = imad / 14.0   # 14 IMAD-pipe issue slots per 64-bit Shoup butterfly: the algorithmic denominator
    f, i = out["forward"]["Gbutterfly_s"] / world * 1e9, out["inverse"]["Gbutterfly_s"] / world * 1e9
    out["roofline"] = {"bound": "int-pipe (IMAD)", "peak_Gbutterfly_s_per_gpu": alg / 1e9, "frac_forward": f / alg if alg else None,
                       "frac_inverse": i / alg if alg else None,
                       "peak_source": "luxb200_intpipe_probe IMAD rate / 14 issue slots per 64-bit butterfly (algorithmic)",
                       "own_butterfly_probe": {"peak_Gbutterfly_s_per_gpu": pk.value / 1e9,
                                               "frac_forward": f / pk.value if pk.value else None,
                                               "frac_inverse": i / pk.value if pk.value else None,
                                               "peak_source": "luxb200_intpipe_probe64: the kernel's own compiled butterfly on registers"}}
    if rank0_cpu:
        out["cpu_baseline"] = cpu_ntt_baseline(N, int(q2[0]), int(psi[0]), 64, os.cpu_count() or 1)
    return out


def _probe_imad(lib, local):
    imad = ctypes.c_double(0); mixed = ctypes.c_double(0)
    lib.luxb200_intpipe_probe(local, ctypes.byref(imad), ctypes.byref(mixed))
    return imad.value, mixed.value


def _peaks():
    try:
        return json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
    except Exception:
        return {}


def _traffic(kernel):
    """DRAM bytes per launch of `kernel` from the tracked ncu capture of this bench command (profiles/r2_traffic.json,
    written by tools/traffic_from_ncu.py from an `ncu --metrics dram__bytes_read.sum,dram__bytes_write.sum` pass)."""
    try:
        t = json.load(open(os.path.join(ROOT, "profiles", "r2_traffic.json")))
        k = t["kernels"][kernel]
        return k["bytes_per_launch"], f"profiles/r2_traffic.json ({t.get('source', '?')}; {k.get('note', '')})"
    except Exception:
        return None, "no tracked capture"


def cpu_ntt_baseline(N, Q, psi, polys, cores):
    """the reference's NumberTheoreticTransformNat (lib-benchmark.cpp:151-168 shape) on the host cores,
    forward, one thread and all threads; bounded sample.  None when the reference library did not travel."""
    kind, orclib = load_checker()
    if kind != "reference":
        return None
    R = orclib.ref()
    rng = np.random.default_rng(3)
    x = rng.integers(0, Q, size=(polys, N), dtype=np.uint64)
    one = min(polys, max(4, polys // cores))
    t1 = R.ref_time_ntt(Q, psi, N, orclib.P(x[:one].copy()), one, 0, 1)
    tn = R.ref_time_ntt(Q, psi, N, orclib.P(x.copy()), polys, 0, cores)
    bf = (N // 2) * (N.bit_length() - 1)
    return {"kind": "reference", "unit": "Gbutterfly/s", "one_thread": one * bf / t1 / 1e9, "value": polys * bf / tn / 1e9,
            "cores": cores, "sample": f"{polys} forward NTTs of N={N} (NativeVector, ChineseRemainderTransformFTT), {tn * 1e3:.1f} ms on {cores} threads, {one} in {t1 * 1e3:.1f} ms on one"}


def ntt32_metric(lib, local, dev, N, bits, count=65536, iters=10):
    """BASELINE configs[1]: `count` polynomials of the binfhe modulus, u32 warp-per-polynomial kernel, device-resident,
    CUDA events on the plan stream.  HBM roofline: algorithmic 2*N*4 bytes per transform against MEASURED_PEAKS hbm_gbs.
    Parity of the timed buffer: round trip back to the input after equally many forward and inverse passes, and one more
    forward pass of 64 polynomials against the oracle."""
    import torch
    from fhe_b200.context import NTTPlan
    Q = lib.luxb200_last_prime(bits, 2 * N)
    pl = NTTPlan(N, [Q], device=local)
    x = torch.randint(0, Q, (count, N), dtype=torch.int32, device=dev)
    x0 = x.clone()
    torch.cuda.synchronize(dev)
    peaks = _peaks()
    hbm = peaks.get("hbm_gbs", 6550.0)
    out = {"workload": f"batched NTT u32 N={N} Q={bits}-bit x{count} ({count * N * 4 >> 20} MB > L2)", "bound": "hbm",
           "peak_gbs": hbm, "peak_source": "MEASURED_PEAKS.json hbm_gbs" if "hbm_gbs" in peaks else "fallback"}
    for inv in (0, 1):
        for _ in range(3):
            pl.SwitchFormatDev(x.data_ptr(), count, inv, 4)
        pl.Sync()
        ms = []
        for _ in range(iters):
            pl.SwitchFormatDev(x.data_ptr(), count, inv, 4)
            ms.append(pl.LastMs())
        t = float(np.median(ms)) * 1e-3
        gbs = count * N * 4 * 2 / t / 1e9
        out["inverse" if inv else "forward"] = {"ms": t * 1e3, "Gbutterfly_s": count * (N // 2) * (N.bit_length() - 1) / t / 1e9,
                                                 "algorithmic_GBs": gbs, "frac_hbm": gbs / hbm}
    ok = bool((x == x0).all().item())
    kind, orclib = load_checker()
    O = orclib.oracle()
    psi = O.orc_root_of_unity(2 * N, Q)
    sample = x0[:64].cpu().numpy().view(np.uint32).astype(np.uint64)
    ref = sample.copy(); O.orc_ntt(Q, psi, N, orclib.P(ref), 64, 0, 8)
    y = x0[:64].clone()
    pl.SwitchFormatDev(y.data_ptr(), 64, 0, 4); pl.Sync()
    ok = ok and bool((y.cpu().numpy().view(np.uint32).astype(np.uint64) == ref).all())
    out["parity"] = ok
    out["cpu_baseline"] = cpu_ntt_baseline(N, Q, psi, 4096, os.cpu_count() or 1)
    pl.close()
    return out


def _lmkcdey_ops(a, q, N, w):
    """(external products, automorphisms) of one LMKCDEY blind rotation, counted as rgsw-acc-lmkcdey.cpp:83-157 walks it:
    masks are bucketed by log_5 of (-a_i | 1) mod 2N; the two sweeps i = N/2-1 .. 1 emit an automorphism before every
    occupied bucket that follows skipped ones and after every `w` skips."""
    M, Nh = 2 * N, N // 2
    assert q == M, "op count written for q = 2N (STD128_LMKCDEY)"
    table, gp = {1: (0, 1), M - 1: (0, -1)}, 1
    for i in range(1, Nh):
        gp = (gp * 5) % M
        table[gp] = (i, 1); table[M - gp] = (i, -1)
    occ = {1: set(), -1: set()}
    for ai in a:
        ai = int(ai) % q
        i, sgn = table[((0 if ai == 0 else M - ai) | 1) % M]
        occ[sgn].add(i)
    autos = 0
    for sgn in (-1, 1):
        skips = 0
        for i in range(Nh - 1, 0, -1):
            if i in occ[sgn] and skips != 0:
                autos += 1; skips = 0
            skips += 1
            if skips == w or i == 1:
                autos += 1; skips = 0
    return len(a), autos + 1   # + the automorphism by 2N - 5 between the sweeps


def method_metric(lib, local, dev, name, method, G, imad_peak):
    """BASELINE configs[3] shapes: G gates of STD128_LMKCDEY / STD128_AP through the scheduled blind-rotation kernel.
    value = device-timed gates/s of the whole call's kernels, e2e = host buffers through the C ABI.  Roofline of
    br_sched_kernel against the live IMAD probe with the algorithmic count written out: per external product 6 transforms
    x 5120 butterflies x 3 IMAD + 8 x 1024 pointwise x 4 IMAD = 124 928; per LMKCDEY automorphism 3 x 5120 x 3 + 4 x 1024 x 4 =
    62 464; op counts are those of the timed inputs (AP: non-zero base-baseR digits of -a_i; LMKCDEY: a sample of 256 op
    lists re-derived on the host)."""
    from fhe_b200 import keygen as kg
    from fhe_b200.context import BinFHEContext
    ctx = BinFHEContext(device=local).GenerateBinFHEContext(name, method)
    ks = kg.keygen(ctx, seed=77)
    ctx.BTKeyLoad(ks.bsk, ks.kskA, ks.kskB)
    rng = np.random.default_rng(5)
    m1 = rng.integers(0, 2, G); m2 = rng.integers(0, 2, G)
    a1, b1 = kg.encrypt(ctx, ks.sk, m1, rng); a2, b2 = kg.encrypt(ctx, ks.sk, m2, rng)
    for _ in range(2):
        ctx.EvalBinGateBatch("AND", (a1[:1024], b1[:1024]), (a2[:1024], b2[:1024]))
    br = tot = wall = 0.0
    reps = 3
    # pinned host buffers through the C ABI call, like the headline e2e (H2D + D2H inside the timed call)
    import torch
    from fhe_b200 import lib as _l
    VP = _l.VP
    h = [torch.from_numpy(np.ascontiguousarray(x).view(np.int64)).pin_memory() for x in (a1, b1, a2, b2)]
    h_oa = torch.empty((G, ctx.n), dtype=torch.int64).pin_memory()
    h_ob = torch.empty(G, dtype=torch.int64).pin_memory()
    for _ in range(reps):
        t0 = time.perf_counter()
        rc = lib.luxb200_evalbingate_batch(ctx.h, _l.GATE["AND"], VP(h[0].data_ptr()), VP(h[1].data_ptr()), VP(h[2].data_ptr()),
                                           VP(h[3].data_ptr()), VP(h_oa.data_ptr()), VP(h_ob.data_ptr()), G, 0)
        wall += time.perf_counter() - t0
        assert rc == 0, lib.luxb200_last_error().decode()
        tm = ctx.LastTiming()
        br += tm["br_ms"]; tot += tm["total_ms"]
    oa = h_oa.numpy().view(np.uint64); ob = h_ob.numpy().view(np.uint64)
    errors = int((kg.decrypt(ctx, ks.sk, oa, ob) != (m1 & m2)).sum())
    s = (a1 + a2) % ctx.q          # the blind rotation runs on ct1 + ct2 (AND)
    EXT, AUTO = 6 * 5120 * 3 + 8 * 1024 * 4, 3 * 5120 * 3 + 4 * 1024 * 4
    if method == "AP":
        neg = (ctx.q - s) % ctx.q
        ops = sum(int(((neg // (ctx.baseR ** k)) % ctx.baseR != 0).sum()) for k in range(ctx.digitsR)) / G
        imad = ops * EXT
        counts = {"external_products_per_gate": ops, "automorphisms_per_gate": 0}
    else:
        sample = [_lmkcdey_ops(s[i], ctx.q, ctx.N, ctx.numAutoKeys) for i in range(0, G, max(1, G // 256))]
        ne = float(np.mean([x[0] for x in sample])); na = float(np.mean([x[1] for x in sample]))
        imad = ne * EXT + na * AUTO
        counts = {"external_products_per_gate": ne, "automorphisms_per_gate": na}
    br_s = br * 1e-3 / reps
    achieved = G * imad / br_s
    out = {"workload": f"{name} EvalBinGate(AND) x{G}", "value": G / (tot * 1e-3 / reps), "unit": "bootstraps/s",
           "e2e": {"value": G * reps / wall, "unit": "bootstraps/s"}, "decrypt_errors": errors,
           "roofline": {"bound": "int32-pipe", "kernel": "br_sched_kernel", "achieved": achieved / 1e12, "peak": imad_peak / 1e12,
                        "unit": "TIMAD/s", "frac": achieved / imad_peak if imad_peak else None,
                        "algorithmic_imad_per_gate": imad, "imad_per_external_product": EXT, "imad_per_automorphism": AUTO,
                        "kernel_ms": br_s * 1e3, **counts}}
    if method == "AP":
        # SURVEY 8(d) expected the key gather to bind AP: bytes of RGSW keys the op lists touch / kernel time vs HBM
        kb = counts["external_products_per_gate"] * ctx.rows * 2 * ctx.N * 4 * G
        peaks = _peaks()
        out["roofline"]["key_stream_GBs_if_from_hbm"] = kb / br_s / 1e9
        out["roofline"]["hbm_peak_gbs"] = peaks.get("hbm_gbs")
        out["roofline"]["note"] = ("even if every key byte the op lists touch came from DRAM the stream would need about a third of "
                                   "the HBM peak; the CTAs walk i together, so most of the 1 GB key is served from L2 (ncu "
                                   "r1_sched_ap_v0: L2 hit 93 %): the integer pipe binds, not HBM (SURVEY 8(d) expected HBM)")
    ctx.close()
    return out


def n2048_metric(lib, local, name, G, imad_peak):
    """The N = 2048 GINX parameter sets (judge-added row of SURVEY 8): G gates on random keys and inputs (throughput does not
    depend on their values; parity of these sets is tests/test_gpu_parity.py::test_wide_sets_vs_oracle), device-timed.
    Roofline for the 32-bit sets against the live IMAD probe: per step (2 INTT + rows NTT) x N/2 log2 N butterflies x 3 IMAD +
    2 keys x rows x 2 x N pointwise x 4 IMAD; the 64-bit sets count 14 slots per butterfly and 8 per multiply-accumulate."""
    from fhe_b200.context import BinFHEContext
    rng = np.random.default_rng(1)
    ctx = BinFHEContext(device=local).GenerateBinFHEContext(name, "GINX")
    bsk = rng.integers(0, ctx.Q, size=ctx.bsk_len(), dtype=np.uint64)
    A = rng.integers(0, ctx.qKS, size=ctx.ksk_rows() * ctx.n, dtype=np.uint64)
    B = rng.integers(0, ctx.qKS, size=ctx.ksk_rows(), dtype=np.uint64)
    ctx.BTKeyLoad(bsk, A, B)
    del bsk, A
    a1 = rng.integers(0, ctx.q, size=(G, ctx.n), dtype=np.uint64); b1 = rng.integers(0, ctx.q, size=G, dtype=np.uint64)
    a2 = rng.integers(0, ctx.q, size=(G, ctx.n), dtype=np.uint64); b2 = rng.integers(0, ctx.q, size=G, dtype=np.uint64)
    ctx.EvalBinGateBatch("AND", (a1[:148], b1[:148]), (a2[:148], b2[:148]))
    (oa, ob), res = ctx.EvalBinGateBatch("AND", (a1, b1), (a2, b2))
    assert res.success, res.error
    t = ctx.LastTiming()
    N, n, rows = ctx.N, ctx.n, ctx.rows
    out = {"workload": f"{name} GINX EvalBinGate(AND) x{G} (n={n}, N={N}, {int(ctx.Q).bit_length()}-bit Q, {rows // 2} digits, random keys)",
           "value": G / (t["total_ms"] * 1e-3), "unit": "bootstraps/s", "br_ms": t["br_ms"]}
    if ctx.Q < (1 << 30):
        steps = float(((a1 + a2) % ctx.q != 0).sum()) / G     # X^0 - 1 = 0 steps are skipped by the shared-memory kernel
        per_step = (2 + rows) * (N // 2) * (N.bit_length() - 1) * 3 + 2 * rows * 2 * N * 4
        achieved = G * n * per_step / (t["br_ms"] * 1e-3)
        out["roofline"] = {"bound": "int32-pipe", "kernel": "br_quad_kernel (br11.cu)" if rows in (4, 6) else "brw_kernel<u32>",
                           "achieved": achieved / 1e12, "peak": imad_peak / 1e12, "unit": "TIMAD/s",
                           "frac": achieved / imad_peak if imad_peak else None, "algorithmic_imad_per_gate": n * per_step,
                           "nonzero_steps_per_gate": steps}
    else:
        # 64-bit residues: 14 IMAD issue slots per butterfly (the algorithmic count of the N = 2^16 NTT metric) and 8 per
        # 64 x 64 -> 128-bit multiply-accumulate (four 32 x 32 -> 64 products of two slots each)
        per_step = (2 + rows) * (N // 2) * (N.bit_length() - 1) * 14 + 2 * rows * 2 * N * 8
        achieved = G * n * per_step / (t["br_ms"] * 1e-3)
        out["roofline"] = {"bound": "int32-pipe", "kernel": "brw_kernel<u64>", "achieved": achieved / 1e12, "peak": imad_peak / 1e12,
                           "unit": "T issue slots/s", "frac": achieved / imad_peak if imad_peak else None,
                           "algorithmic_slots_per_gate": n * per_step}
    ctx.close()
    return out


def hks_metric(lib, local, dev, peaks, imad_peak, count=32, iters=5):
    """SURVEY 8(f) rank 4 at the BASELINE configs[4] ring: hybrid key switching (KeySwitchHYBRID::KeySwitchCore,
    keyswitch-hybrid.cpp:328-436) of `count` polynomials at N = 2^16 over 24 + 8 sixty-bit limbs, 3 digits of 8 limbs, key and data
    resident in HBM (luxb200_hks_keyswitch_dev).  Parity of the timed call: polynomial 0 against the oracle.  Work per key
    switch: 3 x (8 INTT + 24 NTT) + 2 x (8 INTT + 24 NTT) = 160 limb transforms; algorithmic traffic = c in, digits out and in,
    key in, both accumulators out and in, result out."""
    import torch
    from fhe_b200.context import HybridKeySwitchPlan
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    from orclib import P, oracle
    O = oracle()
    N, sizeQ, sizeP, alpha = 65536, 24, 8, 8
    chain = np.zeros(sizeQ + sizeP, np.uint64); psi = chain.copy()
    O.orc_dcrt_chain(N, sizeQ + sizeP, 60, P(chain), P(psi))
    q, p = np.ascontiguousarray(chain[:sizeQ]), np.ascontiguousarray(chain[sizeQ:])
    pl = HybridKeySwitchPlan(N, q, p, alpha, sizeQ, device=local)
    nd, L = pl.num_digits, sizeQ + sizeP
    rng = np.random.default_rng(16)
    c = np.stack([np.stack([rng.integers(0, int(m), size=N, dtype=np.uint64) for m in q]) for _ in range(count)])
    kb = np.stack([np.stack([rng.integers(0, int(m), size=N, dtype=np.uint64) for m in chain]) for _ in range(nd)])
    ka = np.stack([np.stack([rng.integers(0, int(m), size=N, dtype=np.uint64) for m in chain]) for _ in range(nd)])
    d_c = torch.from_numpy(c.view(np.int64)).to(dev); d_kb = torch.from_numpy(kb.view(np.int64)).to(dev)
    d_ka = torch.from_numpy(ka.view(np.int64)).to(dev)
    d_out = torch.empty((count, 2, sizeQ, N), dtype=torch.int64, device=dev)
    torch.cuda.synchronize()
    run = lambda: pl.KeySwitchDev(d_c.data_ptr(), d_kb.data_ptr(), d_ka.data_ptr(), d_out.data_ptr(), count, 0)
    run(); run()
    times = []
    for _ in range(iters):
        t0 = time.perf_counter(); run(); times.append(time.perf_counter() - t0)
    ms = 1e3 * min(times) / count
    ref = np.zeros((1, 2, sizeQ, N), np.uint64)
    O.orc_hybrid_keyswitch(N, sizeQ, P(q), sizeP, P(p), alpha, sizeQ, P(np.ascontiguousarray(c[:1])), P(kb), P(ka), P(ref), 1, 0)
    parity = bool((d_out[:1].cpu().numpy().view(np.uint64) == ref).all())
    transforms = nd * L + 2 * L
    bfly = transforms * (N // 2) * 16
    limbs_moved = sizeQ + 2 * nd * L + 2 * nd * L + 2 * 2 * L + 2 * sizeQ   # c, digits w + r, key b + a, accumulators w + r, result
    gbs = limbs_moved * N * 8 / (ms * 1e-3) / 1e9
    pl.close()
    return {"workload": f"hybrid key switch, N={N}, {sizeQ}+{sizeP} limbs of 60 bits, {nd} digits x {alpha} limbs, {count} polynomials resident in HBM",
            "ms_per_keyswitch": ms, "keyswitches_per_s": 1e3 / ms, "parity": parity,
            "Gbutterfly_s": bfly / (ms * 1e-3) / 1e9, "limb_transforms": transforms,
            "roofline": {"bound": "int-pipe (the NTTs)", "peak_Gbutterfly_s": imad_peak / 14 / 1e9 if imad_peak else None,
                         "frac": (bfly / (ms * 1e-3)) / (imad_peak / 14) if imad_peak else None,
                         "algorithmic_GBs": gbs, "hbm_peak_gbs": peaks.get("hbm_gbs"),
                         "frac_hbm": gbs / peaks["hbm_gbs"] if peaks.get("hbm_gbs") else None,
                         "note": "functional composite: existing kernels with a stream synchronisation between steps (DESIGN 4.8, 8)"}}


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=3)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--gates", type=int, default=65536, help="gates per GPU per step")
    ap.add_argument("--impl", default="ours")
    ap.add_argument("--ref-gates", type=int, default=0)
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--no-ntt", action="store_true")
    ap.add_argument("--no-strong", action="store_true")
    ap.add_argument("--no-extras", action="store_true")
    args = ap.parse_args()
    rank, local, world = env_int("RANK", 0), env_int("LOCAL_RANK", 0), env_int("WORLD_SIZE", 1)

    if args.impl == "reference":
        run_reference(args, rank, world)
        return

    import torch
    import torch.distributed as dist
    from fhe_b200 import keygen as kg
    from fhe_b200.context import BinFHEContext

    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)

    ctx = BinFHEContext(device=local).GenerateBinFHEContext("STD128", "GINX")
    G, n = args.gates, ctx.n

    # ---- keys: generated once on rank 0, broadcast with NCCL over NVLink, loaded from device memory
    bsk_len, ksk_rows = ctx.bsk_len(), ctx.ksk_rows()
    if rank == 0:
        ks = kg.keygen(ctx, seed=2026)
        sk = ks.sk
    if world > 1:
        from fhe_b200 import parallel as par
        lengths = {"bsk": bsk_len, "kskA": ksk_rows * n, "kskB": ksk_rows, "sk": n}
        arrays = {"bsk": ks.bsk, "kskA": ks.kskA, "kskB": ks.kskB, "sk": ks.sk} if rank == 0 else None
        keys = par.broadcast_keys(arrays, lengths, src=0, device=dev)      # NCCL over NVLink, once
        torch.cuda.synchronize()
        ctx.BTKeyLoad(keys["bsk"].data_ptr(), keys["kskA"].data_ptr(), keys["kskB"].data_ptr(), on_device=True)
        sk = keys["sk"].cpu().numpy()
        del keys
        torch.cuda.empty_cache()
    else:
        ctx.BTKeyLoad(ks.bsk, ks.kskA, ks.kskB)

    # ---- synthetic inputs: half real encryptions (decrypt-checkable), half uniform random (SURVEY §8d cfg 3)
    rng = np.random.default_rng(1000 + rank)
    half = G // 2
    m1 = rng.integers(0, 2, size=half); m2 = rng.integers(0, 2, size=half)
    ea1, eb1 = kg.encrypt(ctx, sk, m1, rng); ea2, eb2 = kg.encrypt(ctx, sk, m2, rng)
    a1 = np.concatenate([ea1, rng.integers(0, ctx.q, size=(G - half, n), dtype=np.uint64)])
    a2 = np.concatenate([ea2, rng.integers(0, ctx.q, size=(G - half, n), dtype=np.uint64)])
    b1 = np.concatenate([eb1, rng.integers(0, ctx.q, size=G - half, dtype=np.uint64)])
    b2 = np.concatenate([eb2, rng.integers(0, ctx.q, size=G - half, dtype=np.uint64)])
    # pinned host copies (e2e) and device-resident u32 copies (value)
    h = [torch.from_numpy(x.view(np.int64)).pin_memory() for x in (a1, b1, a2, b2)]
    d = [torch.from_numpy(x.astype(np.uint32).view(np.int32)).to(dev) for x in (a1, b1, a2, b2)]
    d_oa = torch.empty((G, n), dtype=torch.int32, device=dev)
    d_ob = torch.empty(G, dtype=torch.int32, device=dev)
    h_oa = torch.empty((G, n), dtype=torch.int64).pin_memory()
    h_ob = torch.empty(G, dtype=torch.int64).pin_memory()
    torch.cuda.synchronize()

    def barrier():
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def step_dev():
        ctx.EvalBinGateBatchDev("AND", d[0].data_ptr(), d[1].data_ptr(), d[2].data_ptr(), d[3].data_ptr(),
                                d_oa.data_ptr(), d_ob.data_ptr(), G)

    lib = ctx.L

    def step_e2e():
        from fhe_b200 import lib as _l
        _l.check(lib.luxb200_evalbingate_batch(ctx.h, _l.GATE["AND"], _l.VP(h[0].data_ptr()), _l.VP(h[1].data_ptr()),
                                               _l.VP(h[2].data_ptr()), _l.VP(h[3].data_ptr()), _l.VP(h_oa.data_ptr()),
                                               _l.VP(h_ob.data_ptr()), G, 0))

    # ---- warm-up
    for _ in range(args.warmup):
        step_dev()
    ctx.Synchronize()

    # ---- timed region: device-resident inputs (inputs 264 MB + ext 269 MB + KSK gather >> 126 MB L2: no flush needed)
    sampler = ClockSampler(local)
    sampler.start()
    launches0 = ctx.LaunchCount()
    barrier()
    br_ms = ks_ms = ev_ms = 0.0
    t0 = time.perf_counter()
    for _ in range(args.steps):
        step_dev()                      # synchronises per 65536-gate chunk to read its CUDA events
        tm = ctx.LastTiming()
        br_ms += tm["br_ms"]; ks_ms += tm["ks_ms"]; ev_ms += tm["total_ms"]
    ctx.Synchronize()
    barrier()
    t_dev = time.perf_counter() - t0
    launches = ctx.LaunchCount() - launches0
    clocks = sampler.stop()
    # device-side duration of the steps themselves (CUDA events on the library stream)
    if world > 1:
        tt = torch.tensor([t_dev], dtype=torch.float64, device=dev)
        dist.all_reduce(tt, op=dist.ReduceOp.MAX)
        t_dev = float(tt.item())

    # ---- correctness of what was timed: decrypt the encrypted half
    ctx.Synchronize()
    oa = d_oa[:half].cpu().numpy().view(np.uint32).astype(np.uint64)
    ob = d_ob[:half].cpu().numpy().view(np.uint32).astype(np.uint64)
    dec = kg.decrypt(ctx, sk, oa, ob)
    errors = int((dec != (m1 & m2)).sum())       # every rank decrypts the encrypted half of ITS shard
    if world > 1:
        te = torch.tensor([errors], dtype=torch.int64, device=dev)
        dist.all_reduce(te)                         # summed over ranks before rank 0 prints
        errors = int(te.item())

    # ---- e2e through the host-buffer C ABI call (pinned host memory, H2D + D2H inside)
    e2e = None
    if not args.no_e2e:
        step_e2e()
        barrier()
        t0 = time.perf_counter()
        for _ in range(args.steps):
            step_e2e()
        barrier()
        t_e2e = time.perf_counter() - t0
        if world > 1:
            tt = torch.tensor([t_e2e], dtype=torch.float64, device=dev)
            dist.all_reduce(tt, op=dist.ReduceOp.MAX)
            t_e2e = float(tt.item())
        same = bool((h_oa[:half].numpy().view(np.uint64) == oa).all())
        e2e = {"value": G * world * args.steps / t_e2e, "unit": "bootstraps/s",
               "h2d_bytes_per_step": int(2 * G * (n + 1) * 8), "d2h_bytes_per_step": int(G * (n + 1) * 8),
               "matches_device_path": same}

    # ---- second headline metric: DCRT RNS NTT at N=2^16, 32 x 60-bit limbs (BASELINE configs[4]), weak-scaled
    ntt = None
    if not args.no_ntt:
        ntt = ntt_metric(lib, local, dev, world, barrier, rank0_cpu=(rank == 0 and world == 1 and not args.no_cpu_baseline))

    # ---- BASELINE configs[2] as written: 2^16 gates TOTAL, sharded over the N GPUs (strong scaling; the headline above is
    # the weak curve, 2^16 per GPU).  parallel.shard_range gives each rank its contiguous slice of the same inputs.
    strong = None
    if not args.no_strong:
        from fhe_b200 import parallel as par
        TOT = 65536
        lo, hi = par.shard_range(TOT, rank, world)
        m = hi - lo
        sl = [x[:m].contiguous() for x in d]
        hs = [x[:m] for x in h]
        so_a = d_oa[:m]; so_b = d_ob[:m]

        def s_dev():
            ctx.EvalBinGateBatchDev("AND", sl[0].data_ptr(), sl[1].data_ptr(), sl[2].data_ptr(), sl[3].data_ptr(),
                                    so_a.data_ptr(), so_b.data_ptr(), m)

        def s_e2e():
            from fhe_b200 import lib as _l
            _l.check(lib.luxb200_evalbingate_batch(ctx.h, _l.GATE["AND"], _l.VP(hs[0].data_ptr()), _l.VP(hs[1].data_ptr()),
                                                   _l.VP(hs[2].data_ptr()), _l.VP(hs[3].data_ptr()), _l.VP(h_oa.data_ptr()),
                                                   _l.VP(h_ob.data_ptr()), m, 0))
        res = {}
        for nm, fn in (("value", s_dev), ("e2e", s_e2e)):
            fn(); ctx.Synchronize()
            barrier()
            t0 = time.perf_counter()
            for _ in range(args.steps):
                fn()
            ctx.Synchronize()
            barrier()
            t = time.perf_counter() - t0
            if world > 1:
                tt = torch.tensor([t], dtype=torch.float64, device=dev)
                dist.all_reduce(tt, op=dist.ReduceOp.MAX)
                t = float(tt.item())
            res[nm] = TOT * args.steps / t
        strong = {"workload": f"{TOT} STD128 GINX AND gates in total, {m} per GPU (parallel.shard_range)", "scaling": "strong",
                  "value": res["value"], "e2e": res["e2e"], "unit": "bootstraps/s", "gates_per_gpu": m,
                  "waves_per_gpu": m / 888.0}

    if rank != 0:
        if world > 1:
            dist.destroy_process_group()
        return

    # ---- roofline of the dominant kernel (fused blind rotation): integer-pipe bound
    imad_v, mixed_v = _probe_imad(lib, local)
    br_s = br_ms * 1e-3 / args.steps
    achieved = G * IMAD_PER_GATE / br_s
    peaks = _peaks()
    traffic, traffic_src = _traffic("br_ginx_kernel") if G == 65536 else (None, "only captured for 65536 gates")
    roofline = {"bound": "int32-pipe", "kernel": "br_ginx_kernel", "achieved": achieved / 1e12, "peak": imad_v / 1e12,
                "unit": "TIMAD/s", "frac": achieved / imad_v if imad_v else None,
                # DRAM bytes of one br_ginx launch at this batch size (algorithmic: 132 MB masks + 269 MB extracted
                # ciphertexts + 41 MB keys/monomials = 442 MB)
                "traffic": traffic, "traffic_source": traffic_src,
                "peak_source": "luxb200_intpipe_probe (dependent IMAD chains, measured live on this GPU)",
                "algorithmic_imad_per_gate": IMAD_PER_GATE, "kernel_ms_per_step": br_ms / args.steps,
                "kernel_share_of_step": br_ms * 1e-3 / t_dev,
                "butterfly_peak_per_s": mixed_v, "hbm_peak_gbs": peaks.get("hbm_gbs")}

    # ---- the other BASELINE configs, N = 1 only (a few steps each, after the headline): configs[1] u32 NTT batches,
    # configs[3] LMKCDEY / AP gates.  Each carries its own roofline fraction.
    extras = {}
    if world == 1 and not args.no_extras:
        ctx.close()
        del d, d_oa, d_ob
        torch.cuda.empty_cache()
        extras["ntt32_n1024"] = ntt32_metric(lib, local, dev, 1024, 27)
        extras["ntt32_n2048"] = ntt32_metric(lib, local, dev, 2048, 29)
        extras["lmkcdey"] = method_metric(lib, local, dev, "STD128_LMKCDEY", "LMKCDEY", 16384, imad_v)
        extras["ap"] = method_metric(lib, local, dev, "STD128_AP", "AP", 4096, imad_v)
        extras["hks_n65536"] = hks_metric(lib, local, dev, _peaks(), imad_v)
        # 1776 gates = whole waves for 3 (two digits), 2 (three digits) and 3 (shared-memory kernel) ciphertexts per SM
        extras["n2048"] = {k: n2048_metric(lib, local, k, 1776, imad_v) for k in ("STD256", "STD256_3", "STD192")}

    cpu_baseline = None
    if not args.no_cpu_baseline:
        kind, orclib = load_checker()
        cores = os.cpu_count() or 1
        gates = cores * 32
        fn = cpu_time_gates(kind, orclib, gates, cores)
        sec = fn()
        cpu_baseline = {"value": gates / sec, "unit": "bootstraps/s", "cores": cores, "kind": kind,
                        "sample": f"{gates} STD128 GINX AND gates, OpenMP over gates, {cores} threads, {sec:.1f}s"}

    value = G * world * args.steps / t_dev
    print(json.dumps({
        "metric": "STD128 GINX bootstraps/sec", "value": value, "unit": "bootstraps/s", "n_gpus": world,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": 1e3 * t_dev / args.steps, "higher_is_better": True,
        "scaling": "weak", "vs_baseline": None, "dtype": "u32", "data": "synthetic",
        "config": {"workload": "binfhe STD128 GINX EvalBinGate(AND) batch (BASELINE configs[2])", "gates_per_gpu": G,
                   "global_gates": G * world, "n": n, "N": ctx.N, "Q": ctx.Q, "parallelism": f"batch-shard x{world}",
                   "l2": "working set per step (inputs 264 MB, extracted cts 269 MB, KSK 100 MB) exceeds the 126 MB L2"},
        "clocks": clocks, "e2e": e2e, "gpu_launches": int(launches), "roofline": roofline,
        "cpu_baseline": cpu_baseline, "ntt_n65536": ntt, "strong_2p16": strong, **extras,
        "decrypt_errors": errors, "decrypt_errors_scope": f"encrypted half of every rank's shard, summed over {world} rank(s)",
        "br_ms_per_step": br_ms / args.steps,
        "ks_ms_per_step": ks_ms / args.steps,
        # CUDA events on the library stream around every kernel of the step (prep + blind rotation + key switch), rank 0;
        # `value` uses the wall clock between the two barrier + synchronize brackets (max over ranks), which contains it
        "device_ms_per_step": ev_ms / args.steps}))
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
