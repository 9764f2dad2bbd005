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


def ntt_metric(lib, local, dev, world, barrier, N=65536, Lm=32, cnt=32, iters=5):
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
    pl.close()
    pk = ctypes.c_double(0)
    lib.luxb200_intpipe_probe64(local, ctypes.byref(pk))
    out["roofline"] = {"bound": "int-pipe (IMAD)", "peak_Gbutterfly_s_per_gpu": pk.value / 1e9,
                       "frac_forward": out["forward"]["Gbutterfly_s"] / world / (pk.value / 1e9) if pk.value else None,
                       "frac_inverse": out["inverse"]["Gbutterfly_s"] / world / (pk.value / 1e9) if pk.value else None,
                       "peak_source": "luxb200_intpipe_probe64: the kernel's own 64-bit Shoup butterfly on registers"}
    return out


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
    errors = int((dec != (m1 & m2)).sum())

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
        ntt = ntt_metric(lib, local, dev, world, barrier)

    if rank != 0:
        if world > 1:
            dist.destroy_process_group()
        return

    # ---- roofline of the dominant kernel (fused blind rotation): integer-pipe bound
    imad = ctypes.c_double(0); mixed = ctypes.c_double(0)
    lib.luxb200_intpipe_probe(local, ctypes.byref(imad), ctypes.byref(mixed))
    br_s = br_ms * 1e-3 / args.steps
    achieved = G * IMAD_PER_GATE / br_s
    peaks = {}
    try:
        peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
    except Exception:
        pass
    roofline = {"bound": "int32-pipe", "kernel": "br_ginx_kernel", "achieved": achieved / 1e12, "peak": imad.value / 1e12,
                "unit": "TIMAD/s", "frac": achieved / imad.value if imad.value else None,
                # DRAM bytes of one br_ginx launch at this batch size, from the ncu pass in profiles/r1_launches_v3.csv
                # (196 MB read + 252 MB written; algorithmic: 132 MB masks + 269 MB extracted cts + 41 MB keys/monomials)
                "traffic": 438.7e6 if G == 65536 else None, "traffic_unit": "bytes/launch (profiles/r1_launches_v3.csv: 189.3 MB read + 249.4 MB written)",
                "peak_source": "luxb200_intpipe_probe (dependent IMAD chains, measured live on this GPU)",
                "algorithmic_imad_per_gate": IMAD_PER_GATE, "kernel_ms_per_step": br_ms / args.steps,
                "kernel_share_of_step": br_ms * 1e-3 / t_dev,
                "butterfly_peak_per_s": mixed.value, "hbm_peak_gbs": peaks.get("hbm_gbs")}

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
        "cpu_baseline": cpu_baseline, "ntt_n65536": ntt, "decrypt_errors": errors, "br_ms_per_step": br_ms / args.steps,
        "ks_ms_per_step": ks_ms / args.steps,
        # CUDA events on the library stream around every kernel of the step (prep + blind rotation + key switch), rank 0;
        # `value` uses the wall clock between the two barrier + synchronize brackets (max over ranks), which contains it
        "device_ms_per_step": ev_ms / args.steps}))
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
