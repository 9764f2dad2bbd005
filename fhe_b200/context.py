"""Host-side mirror of the reference's batch interface over libluxb200.so.

Names and semantics follow the reference so that parity tests read like its own:
  BinFHEContext.GenerateBinFHEContext(set, method)      src/binfhe/lib/binfhecontext.cpp:109-187 (parameter table :113-159)
  BinFHEContext.BTKeyLoad(keys)                          src/binfhe/include/binfhecontext.h:273
  EvalBinGateBatch / BootstrapBatch / EvalFuncBatch / EvalCMUXBatch / KeySwitchBatch / ModSwitchBatch
                                                         src/binfhe/include/batch/binfhe-batch.h:48-174
  BatchResult{success, processed, failed, error}         binfhe-batch.h:29-40
  BlindRotateBatch                                       src/binfhe/include/backend/backend.h:177-183
Ciphertext batches are flat arrays: a [count][n] uint64, b [count] uint64 (NativeVector element type).
Key generation stays with the caller (the reference's BTKeyGen, or fhe_b200.keygen for synthetic keys).
"""
import ctypes as C
from dataclasses import dataclass

import numpy as np

from . import lib as _l

PRIME = 0
# { name: (numberBits, cyclOrder, latticeParam n, mod q, modKS, baseKS, gadgetBase, baseRK, numAutoKeys, keyDist) }
# binfhecontext.cpp:117-158 (keyDist: 0 = UNIFORM_TERNARY, 1 = GAUSSIAN)
PARAM_SETS = {
    "TOY": (27, 1024, 64, 512, PRIME, 25, 512, 23, 9, 0),
    "MEDIUM": (28, 2048, 422, 1024, 16384, 128, 1024, 32, 10, 0),
    "STD128_AP": (27, 2048, 503, 1024, 16384, 32, 512, 32, 10, 0),
    "STD128": (27, 2048, 503, 1024, 16384, 32, 512, 32, 10, 0),
    "STD128_3": (27, 2048, 595, 1024, 65536, 64, 128, 32, 10, 0),
    "STD128_4": (27, 2048, 595, 2048, 65536, 64, 128, 64, 10, 0),
    "STD128Q": (25, 2048, 534, 1024, 16384, 32, 128, 32, 10, 0),
    "STD128_LMKCDEY": (28, 2048, 447, 2048, 16384, 32, 1024, 64, 10, 1),
    "STD128_3_LMKCDEY": (27, 2048, 556, 2048, 32768, 32, 512, 64, 10, 0),
    "STD128_4_LMKCDEY": (27, 2048, 595, 2048, 65536, 64, 128, 64, 10, 0),
    "STD128Q_LMKCDEY": (27, 2048, 483, 2048, 16384, 32, 512, 64, 10, 1),
    "STD128Q_3_LMKCDEY": (25, 2048, 643, 2048, 65536, 64, 128, 64, 10, 0),
    # N = 2048 and / or Q above 2^30: the shared-memory blind rotation (csrc/br_wide.cuh)
    "STD128Q_3": (50, 4096, 600, 2048, 32768, 32, 33554432, 64, 10, 0),
    "STD128Q_4": (50, 4096, 641, 2048, 65536, 64, 33554432, 64, 10, 0),
    "STD192": (37, 4096, 790, 2048, 16384, 32, 524288, 64, 10, 0),
    "STD192_3": (37, 4096, 875, 4096, 65536, 64, 524288, 64, 10, 0),
    "STD192_4": (37, 4096, 875, 4096, 65536, 64, 8192, 64, 10, 0),
    "STD192Q": (35, 4096, 875, 1024, 32768, 32, 4096, 32, 10, 0),
    "STD192Q_3": (34, 4096, 922, 2048, 65536, 16, 4096, 64, 10, 0),
    "STD192Q_4": (34, 4096, 980, 2048, 131072, 16, 4096, 64, 10, 0),
    "STD256": (29, 4096, 1076, 2048, 32768, 32, 1024, 64, 10, 0),
    "STD256_3": (29, 4096, 1145, 2048, 65536, 64, 256, 64, 10, 0),
    "STD256_4": (29, 4096, 1145, 4096, 65536, 64, 256, 64, 10, 0),
    "STD256Q": (27, 4096, 1225, 1024, 65536, 16, 128, 32, 10, 0),
    "STD256Q_3": (27, 4096, 1400, 4096, 65536, 21, 64, 64, 10, 0),
    "STD256Q_4": (27, 4096, 1625, 4096, 2097152, 16, 64, 64, 10, 0),
    "STD128Q_4_LMKCDEY": (50, 4096, 641, 4096, 65536, 64, 33554432, 64, 10, 0),
    "STD192_LMKCDEY": (39, 4096, 716, 2048, 32768, 32, 1048576, 64, 10, 1),
    "STD192_3_LMKCDEY": (39, 4096, 771, 4096, 65536, 64, 1048576, 64, 10, 1),
    "STD192_4_LMKCDEY": (37, 4096, 875, 4096, 65536, 64, 8192, 64, 10, 0),
    "STD192Q_LMKCDEY": (36, 4096, 776, 4096, 32768, 32, 262144, 64, 10, 1),
    "STD192Q_3_LMKCDEY": (36, 4096, 834, 4096, 65536, 64, 4096, 64, 10, 1),
    "STD192Q_4_LMKCDEY": (34, 4096, 949, 4096, 65536, 64, 4096, 64, 10, 0),
    "STD256_LMKCDEY": (30, 4096, 939, 2048, 32768, 32, 1024, 64, 10, 1),
    "STD256_3_LMKCDEY": (29, 4096, 1076, 4096, 32768, 32, 256, 64, 10, 0),
    "STD256_4_LMKCDEY": (29, 4096, 1145, 4096, 65536, 64, 256, 64, 10, 0),
    "STD256Q_LMKCDEY": (28, 4096, 1019, 4096, 32768, 32, 1024, 64, 10, 1),
    "STD256Q_3_LMKCDEY": (26, 4096, 1242, 4096, 65536, 64, 128, 64, 10, 0),
    "STD256Q_4_LMKCDEY": (26, 4096, 1320, 4096, 131072, 64, 64, 64, 10, 0),
    "LPF_STD128": (27, 2048, 556, 2048, 32768, 32, 128, 64, 10, 0),
    "LPF_STD128Q": (25, 2048, 645, 2048, 65536, 64, 128, 64, 10, 0),
    "LPF_STD128Q_LMKCDEY": (25, 2048, 600, 2048, 32768, 32, 128, 64, 10, 0),
    "LPF_STD128_LMKCDEY": (27, 2048, 556, 2048, 32768, 32, 512, 64, 10, 0),
    "SIGNED_MOD_TEST": (28, 2048, 512, 1024, PRIME, 25, 128, 23, 10, 0),
}


@dataclass
class BatchResult:
    success: bool = True
    processed: int = 0
    failed: int = 0
    error: str = ""


class BinFHEContext:
    """Device-side counterpart of lux::fhe::BinFHEContext for the batched hot path (one context = one GPU)."""

    def __init__(self, device=0):
        self.L = _l.load()
        self.device = device
        self.h = None
        self.info = None

    # -- parameters ---------------------------------------------------------------------------------------
    def GenerateBinFHEContext(self, paramset, method="GINX"):
        if paramset not in PARAM_SETS:
            raise _l.LuxB200Error(-4, f"unknown parameter set {paramset}")
        bits, cyc, n, q, modKS, baseKS, baseG, baseR, nauto, keyDist = PARAM_SETS[paramset]
        Q = self.L.luxb200_last_prime(bits, cyc)
        self.keyDist = keyDist
        return self.GenerateFromParams(n=n, N=cyc // 2, q=q, Q=Q, qKS=(Q if modKS == PRIME else modKS), baseKS=baseKS,
                                       baseG=baseG, baseR=baseR, numAutoKeys=nauto, method=method)

    def GenerateFromParams(self, n, N, q, Q, qKS, baseKS, baseG, baseR, numAutoKeys, method, psi=0):
        self.close()
        p = _l.Params(n=n, N=N, q=q, Q=Q, qKS=qKS, baseKS=baseKS, baseG=baseG, baseR=baseR, numAutoKeys=numAutoKeys,
                      method=_l.METHOD[method] if isinstance(method, str) else int(method), reserved=0, psi=psi)
        h = _l.VP()
        _l.check(self.L.luxb200_ctx_create(C.byref(p), self.device, C.byref(h)))
        self.h = h
        arr = np.zeros(16, np.uint64)
        _l.check(self.L.luxb200_ctx_info(self.h, _l.ptr(arr)))
        names = ["n", "N", "q", "Q", "qKS", "baseKS", "baseG", "baseR", "digitsG", "numAutoKeys", "_", "psi", "digitsKS",
                 "method", "digitsR", "device"]
        for k, v in zip(names, arr):
            setattr(self, k, int(v))
        self.rows = 2 * (self.digitsG - 1)
        return self

    def GetBeta(self):
        return 128  # binfhecontext.h:445-447

    def bsk_len(self):
        return int(self.L.luxb200_bsk_len(self.h))

    def ksk_rows(self):
        return self.N * self.baseKS * self.digitsKS

    # -- keys -----------------------------------------------------------------------------------------------
    def BTKeyLoad(self, bsk, kskA, kskB, on_device=False):
        """bsk / kskA / kskB: canonical flat uint64 arrays (numpy) or device pointers (ints) when on_device."""
        _l.check(self.L.luxb200_load_bsk(self.h, _l.ptr(bsk), int(on_device)))
        _l.check(self.L.luxb200_load_ksk(self.h, _l.ptr(kskA), _l.ptr(kskB), int(on_device)))

    def BTKeyGen(self, sk, seed, want_skN=False):
        """BinFHEContext::BTKeyGen (binfhecontext.cpp:285-307) on the device: sk = LWE secret (n signed values); the ring
        secret, the switching key and the refresh key are drawn on the GPU (ChaCha20 streams from `seed`) straight into
        the packed layouts.  Not bit-equal to the reference's Blake2-seeded keys; same structure and noise."""
        sk = np.ascontiguousarray(sk, np.int64)
        assert sk.shape == (self.n,)
        skN = np.zeros(self.N, np.int64) if want_skN else None
        _l.check(self.L.luxb200_btkeygen(self.h, _l.ptr(sk), int(seed) & (2**64 - 1), _l.ptr(skN) if want_skN else None))
        return skN

    def ExportKeys(self):
        """(bsk, kskA, kskB) of the resident keys in the canonical flat uint64 layouts."""
        bsk = np.empty(self.bsk_len(), np.uint64)
        A = np.empty(self.ksk_rows() * self.n, np.uint64)
        B = np.empty(self.ksk_rows(), np.uint64)
        _l.check(self.L.luxb200_export_keys(self.h, _l.ptr(bsk), _l.ptr(A), _l.ptr(B)))
        return bsk, A, B

    # -- batched operators (host buffers) -----------------------------------------------------------------
    def _u64(self, x):
        return np.ascontiguousarray(x, dtype=np.uint64)

    def EvalBinGateBatch(self, gate, ct1, ct2, extended=False):
        (a1, b1), (a2, b2) = ct1, ct2
        a1, b1, a2, b2 = map(self._u64, (a1, b1, a2, b2))
        cnt = len(b1)
        oa = np.empty((cnt, self.N if extended else self.n), np.uint64)
        ob = np.empty(cnt, np.uint64)
        rc = self.L.luxb200_evalbingate_batch(self.h, _l.GATE[gate], _l.ptr(a1), _l.ptr(b1), _l.ptr(a2), _l.ptr(b2),
                                              _l.ptr(oa), _l.ptr(ob), cnt, int(extended))
        return (oa, ob), self._result(rc, cnt)

    def EvalBinGateMultiBatch(self, gate, a, b, p=4):
        a, b = self._u64(a), self._u64(b)
        cnt, k = b.shape
        oa = np.empty((cnt, self.n), np.uint64)
        ob = np.empty(cnt, np.uint64)
        rc = self.L.luxb200_evalbingate_multi_batch(self.h, _l.GATE[gate], k, _l.ptr(a), _l.ptr(b), _l.ptr(oa),
                                                    _l.ptr(ob), cnt, p)
        return (oa, ob), self._result(rc, cnt)

    def EvalCMUXBatch(self, ct_sel, ct_true, ct_false):
        """lux::fhe::EvalCMUXBatch(cc, ct_sel, ct_true, ct_false, ...) (binfhe-batch.h:113-125): per element
        EvalBinGate(CMUX, {sel, true, false}) — the vector order of batch.cpp:231-239, which host/batch_b200.cpp keeps too."""
        a = np.stack([self._u64(ct_sel[0]), self._u64(ct_true[0]), self._u64(ct_false[0])], axis=1)
        b = np.stack([self._u64(ct_sel[1]), self._u64(ct_true[1]), self._u64(ct_false[1])], axis=1)
        return self.EvalBinGateMultiBatch("CMUX", a, b)

    def BootstrapBatch(self, ct, p=4, extended=False):
        a, b = map(self._u64, ct)
        cnt = len(b)
        oa = np.empty((cnt, self.N if extended else self.n), np.uint64)
        ob = np.empty(cnt, np.uint64)
        rc = self.L.luxb200_bootstrap_batch(self.h, _l.ptr(a), _l.ptr(b), _l.ptr(oa), _l.ptr(ob), cnt, p, int(extended))
        return (oa, ob), self._result(rc, cnt)

    def EvalFuncBatch(self, ct, lut, ctmod=None):
        a, b = map(self._u64, ct)
        lut = self._u64(lut)
        cnt = len(b)
        oa = np.empty((cnt, self.n), np.uint64)
        ob = np.empty(cnt, np.uint64)
        rc = self.L.luxb200_evalfunc_batch(self.h, _l.ptr(a), _l.ptr(b), _l.ptr(lut), len(lut), self.GetBeta(),
                                           _l.ptr(oa), _l.ptr(ob), cnt, ctmod or self.q)
        return (oa, ob), self._result(rc, cnt)

    def BlindRotateBatch(self, a, acc, amod=None):
        a = self._u64(a)
        acc = self._u64(acc).copy()
        _l.check(self.L.luxb200_blind_rotate_batch(self.h, _l.ptr(a), amod or self.q, _l.ptr(acc), a.shape[0]))
        return acc

    def ExternalProductBatch(self, rgsw, rlwe):
        """Backend::ExternalProductBatch (backend.h:185-194): rgsw [count][rows][2][N], rlwe [count][2][N], EVALUATION."""
        rgsw, rlwe = self._u64(rgsw), self._u64(rlwe)
        out = np.empty_like(rlwe)
        _l.check(self.L.luxb200_external_product_batch(self.h, _l.ptr(rgsw), _l.ptr(rlwe), _l.ptr(out), rlwe.shape[0]))
        return out

    def KeySwitchBatch(self, ct):
        a, b = map(self._u64, ct)
        cnt = len(b)
        oa = np.empty((cnt, self.n), np.uint64)
        ob = np.empty(cnt, np.uint64)
        _l.check(self.L.luxb200_keyswitch_batch(self.h, _l.ptr(a), _l.ptr(b), _l.ptr(oa), _l.ptr(ob), cnt))
        return oa, ob

    def ModSwitchBatch(self, x, qto, qfrom):
        x = self._u64(x)
        out = np.empty_like(x)
        _l.check(self.L.luxb200_modswitch_batch(self.h, qto, qfrom, _l.ptr(x), _l.ptr(out), x.size))
        return out

    def GenerateLUTviaFunction(self, f, p):
        """binfhecontext.cpp:372-387"""
        q = self.q
        out = np.zeros(q, np.uint64)
        for i in range(q):
            v = (q // p) * f((i * p) // q, p)
            if v >= q:
                raise ValueError("input function should output in Z_{p_output}")
            out[i] = v
        return out

    # -- large precision (binfhe-base-scheme.cpp:335-521) -------------------------------------------------
    def EvalFloorBatch(self, ct, mod, roundbits=0):
        """BinFHEContext::EvalFloor (binfhecontext.cpp:346-357) per element; ct = (a [count][n], b [count]) mod `mod`."""
        a, b = map(self._u64, ct)
        cnt = len(b)
        oa = np.empty((cnt, self.n), np.uint64)
        ob = np.empty(cnt, np.uint64)
        rc = self.L.luxb200_evalfloor_batch(self.h, _l.ptr(a), _l.ptr(b), _l.ptr(oa), _l.ptr(ob), cnt, mod, self.GetBeta(),
                                            roundbits)
        return (oa, ob), self._result(rc, cnt)

    # -- device-resident operators (uint32 tensors already in HBM; pointers as ints) ------------------------
    # -- gate circuits (BatchDAG, binfhe-batch.h:255-279) ------------------------------------------------
    def EvalCircuitBatch(self, n_inputs, gates, outputs, in_a, in_b, luts=None):
        """One circuit over `count` independent input sets, level-scheduled on the device with device-resident
        intermediates (luxb200_circuit_batch[_ex]).  gates: [(op, in0, in1)], op a 2-input gate name, "NOT", "BOOTSTRAP" or
        "EVALFUNC" (in1 = index into `luts`, a list of length-q tables as GenerateLUTviaFunction returns them);
        wire ids as BatchDAG hands them out (inputs 0..n_inputs-1, then one per gate).  in_a [n_inputs][count][n],
        in_b [n_inputs][count]; returns ((out_a [n_outputs][count][n], out_b [n_outputs][count]), BatchResult)."""
        ops = {"NOT": 100, "BOOTSTRAP": 101, "EVALFUNC": 102}
        g = np.zeros((max(len(gates), 1), 3), np.uint32)
        for k, (op, i0, i1) in enumerate(gates):
            g[k] = (ops[op] if op in ops else _l.GATE[op], i0, i1 if i1 is not None else 0)
        outs = np.ascontiguousarray(outputs, dtype=np.uint32)
        in_a, in_b = self._u64(in_a), self._u64(in_b)
        assert in_b.shape[0] == n_inputs and in_a.shape[:2] == in_b.shape
        cnt = in_b.shape[1]
        oa = np.empty((len(outs), cnt, self.n), np.uint64)
        ob = np.empty((len(outs), cnt), np.uint64)
        lt = self._u64(np.stack(luts)) if luts else None
        rc = self.L.luxb200_circuit_batch_ex(self.h, n_inputs, _l.ptr(g), len(gates), _l.ptr(outs), len(outs),
                                             _l.ptr(lt) if lt is not None else None, len(luts) if luts else 0, self.GetBeta(),
                                             _l.ptr(in_a), _l.ptr(in_b), _l.ptr(oa), _l.ptr(ob), cnt)
        return (oa, ob), self._result(rc, len(gates) * cnt)

    def CircuitStats(self):
        t = np.zeros(3, np.uint64)
        _l.check(self.L.luxb200_circuit_last_stats(self.h, _l.ptr(t)))
        return {"levels": int(t[0]), "bootstrapped_gates": int(t[1]), "launches": int(t[2])}

    def EvalBinGateBatchDev(self, gate, a1, b1, a2, b2, oa, ob, count):
        _l.check(self.L.luxb200_evalbingate_batch_dev(self.h, _l.GATE[gate], _l.VP(a1), _l.VP(b1), _l.VP(a2), _l.VP(b2),
                                                      _l.VP(oa), _l.VP(ob), count))

    def Synchronize(self):
        _l.check(self.L.luxb200_synchronize(self.h))

    def LastTiming(self):
        t = np.zeros(8, np.float32)
        _l.check(self.L.luxb200_last_timing(self.h, _l.ptr(t)))
        return {"prep_ms": float(t[0]), "br_ms": float(t[1]), "ks_ms": float(t[2]), "total_ms": float(t[3]),
                "launches": int(t[4])}

    def LaunchCount(self):
        return int(self.L.luxb200_launch_count(self.h))

    def _result(self, rc, cnt):
        if rc == 0:
            return BatchResult(True, cnt, 0, "")
        return BatchResult(False, 0, cnt, self.L.luxb200_last_error().decode())

    def close(self):
        if self.h:
            self.L.luxb200_ctx_destroy(self.h)
            self.h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


class NTTPlan:
    """PolyImpl::SwitchFormat / DCRTPolyImpl::SwitchFormat on flat [count][L][N] data."""

    def __init__(self, N, moduli, roots=None, device=0):
        self.L = _l.load()
        q = np.ascontiguousarray(moduli, np.uint64)
        psi = np.ascontiguousarray(roots, np.uint64) if roots is not None else None
        h = _l.VP()
        _l.check(self.L.luxb200_ntt_plan_create(N, len(q), _l.ptr(q), _l.ptr(psi) if psi is not None else None, device,
                                                C.byref(h)))
        self.h, self.N, self.nl = h, N, len(q)

    def SwitchFormat(self, data, inverse):
        data = np.ascontiguousarray(data, np.uint64).copy()
        count = data.size // (self.N * self.nl)
        _l.check(self.L.luxb200_ntt_batch(self.h, _l.ptr(data), count, int(inverse)))
        return data

    def SwitchFormatDev(self, dptr, count, inverse, elem_bytes):
        _l.check(self.L.luxb200_ntt_batch_dev(self.h, _l.VP(dptr), count, int(inverse), elem_bytes))

    def Sync(self):
        _l.check(self.L.luxb200_ntt_plan_sync(self.h))

    def LastMs(self):
        return float(self.L.luxb200_ntt_last_ms(self.h))

    def close(self):
        if self.h:
            self.L.luxb200_ntt_plan_destroy(self.h)
            self.h = None


# ---- large-precision contexts: GenerateBinFHEContext(set, arbFunc, logQ, N, GINX, timeOptimization) ------------------
def GenerateLargePrecisionContexts(paramset, arbFunc=False, logQ=11, N=0, timeOptimization=False, device=0):
    """binfhecontext.cpp:53-104: Q = LastPrime(54, 2N) (27 bits for logQ = 11), N = max(N, FindRingDim(ternary, 128 classic,
    logQ')) — 2048 for 54 bits, 1024 for 27 (stdlatticeparms.cpp:169-172) — q = N (arbFunc) or 2N, qKS = 2^35, baseKS = 32,
    n = 32 (TOY) or 1305 (STD128), baseR = 23, GINX.  Returns the list of device contexts, the current gadget base first;
    with timeOptimization (and logQ != 11) one per base of the reference's key map (2^14, 2^18, 2^27,
    rgsw-cryptoparameters.cpp:49-66), as BTKeyGen fills it (binfhecontext.cpp:292-298)."""
    if paramset not in ("TOY", "STD128"):
        raise _l.LuxB200Error(-1, "STD128 and TOY are the only supported sets")
    if logQ > 29 or logQ < 11:
        raise _l.LuxB200Error(-1, "logQ must be in [11, 29]")
    logQp = 54
    if logQ > 25:
        baseG = 1 << 14
    elif logQ > 16:
        baseG = 1 << 18
    elif logQ > 11:
        baseG = 1 << 27
    else:
        baseG, logQp = 1 << 5, 27
    ring = max(N, 2048 if logQp == 54 else 1024)
    L = _l.load()
    Q = L.luxb200_last_prime(logQp, 2 * ring)
    q = ring if arbFunc else 2 * ring
    n = 32 if paramset == "TOY" else 1305
    bases = [baseG]
    if timeOptimization and logQ != 11:
        bases += [b for b in (1 << 14, 1 << 18, 1 << 27) if b != baseG]
    out = []
    for bg in bases:
        c = BinFHEContext(device=device)
        c.GenerateFromParams(n=n, N=ring, q=q, Q=Q, qKS=1 << 35, baseKS=32, baseG=bg, baseR=23, numAutoKeys=10, method="GINX")
        c.keyDist = 0
        out.append(c)
    return out


def _ctx_array(ctxs):
    return (C.c_void_p * len(ctxs))(*[c.h for c in ctxs])


def EvalSignBatch(ctxs, ct, mod, schemeSwitch=False):
    """BinFHEContext::EvalSign (binfhecontext.cpp:359-364) per element over the key map `ctxs` (current base first; three
    contexts = the reference's dynamic mode).  Returns ((a, b) mod q, BatchResult)."""
    c0 = ctxs[0]
    a, b = map(c0._u64, ct)
    cnt = len(b)
    oa = np.empty((cnt, c0.n), np.uint64)
    ob = np.empty(cnt, np.uint64)
    rc = c0.L.luxb200_evalsign_batch(_ctx_array(ctxs), len(ctxs), _l.ptr(a), _l.ptr(b), _l.ptr(oa), _l.ptr(ob), cnt, mod,
                                     c0.GetBeta(), int(schemeSwitch))
    return (oa, ob), c0._result(rc, cnt)


def EvalDecompBatch(ctxs, ct, mod):
    """BinFHEContext::EvalDecomp (binfhecontext.cpp:366-370): ((a [k][count][n], b [k][count], moduli [k]), BatchResult)."""
    c0 = ctxs[0]
    a, b = map(c0._u64, ct)
    cnt = len(b)
    k = int(c0.L.luxb200_evaldecomp_count(c0.q, mod, c0.GetBeta()))
    oa = np.empty((k, cnt, c0.n), np.uint64)
    ob = np.empty((k, cnt), np.uint64)
    om = np.zeros(k, np.uint64)
    nout = C.c_uint32(0)
    rc = c0.L.luxb200_evaldecomp_batch(_ctx_array(ctxs), len(ctxs), _l.ptr(a), _l.ptr(b), _l.ptr(oa), _l.ptr(ob), _l.ptr(om), k,
                                       C.byref(nout), cnt, mod, c0.GetBeta())
    return (oa, ob, om), c0._result(rc, cnt)


class RNSPlan:
    """DCRTPolyImpl::ApproxSwitchCRTBasis / ApproxModUp / ApproxModDown (dcrtpoly-impl.h:910-1030) on flat [count][limbs][N]
    uint64 data: the basis-conversion steps of CKKS/BGV hybrid key switching either side of the DCRT NTT."""

    def __init__(self, N, q, p, device=0):
        self.L = _l.load()
        self.q = np.ascontiguousarray(q, np.uint64)
        self.p = np.ascontiguousarray(p, np.uint64)
        h = _l.VP()
        _l.check(self.L.luxb200_rns_plan_create(N, len(self.q), _l.ptr(self.q), len(self.p), _l.ptr(self.p), device, C.byref(h)))
        self.h, self.N = h, N

    def ApproxSwitchCRTBasis(self, x, direction=0):
        x = np.ascontiguousarray(x, np.uint64)
        cnt = x.shape[0]
        out = np.empty((cnt, len(self.q) if direction else len(self.p), self.N), np.uint64)
        _l.check(self.L.luxb200_approx_switch_crt_basis(self.h, _l.ptr(x), _l.ptr(out), cnt, direction))
        return out

    def ApproxSwitchCRTBasisStatic(self, x, direction=0):
        """luxb200_approx_switch_crt_basis_static: the same result with no cudaMalloc (device staging in the library's static
        16 MiB array; larger batches are refused).  No counterpart in the reference."""
        x = np.ascontiguousarray(x, np.uint64)
        cnt = x.shape[0]
        out = np.empty((cnt, len(self.q) if direction else len(self.p), self.N), np.uint64)
        _l.check(self.L.luxb200_approx_switch_crt_basis_static(self.h, _l.ptr(x), _l.ptr(out), cnt, direction))
        return out

    def ApproxModUp(self, x):
        x = np.ascontiguousarray(x, np.uint64)
        out = np.empty((x.shape[0], len(self.q) + len(self.p), self.N), np.uint64)
        _l.check(self.L.luxb200_approx_mod_switch(self.h, _l.ptr(x), _l.ptr(out), x.shape[0], 0))
        return out

    def KeySwitchInner(self, digits, evk_b, evk_a, sizeQl=None):
        """KeySwitchHYBRID::EvalFastKeySwitchCoreExt (keyswitch-hybrid.cpp:438-481): digits [count][nd][sizeQl + sizeP][N], the key's
        b and a vectors [nd][sizeQ + sizeP][N] -> [count][2][sizeQl + sizeP][N]."""
        digits = np.ascontiguousarray(digits, np.uint64)
        evk_b = np.ascontiguousarray(evk_b, np.uint64); evk_a = np.ascontiguousarray(evk_a, np.uint64)
        cnt, nd, L = digits.shape[0], digits.shape[1], digits.shape[2]
        sizeQl = L - len(self.p) if sizeQl is None else sizeQl
        out = np.empty((cnt, 2, L, self.N), np.uint64)
        _l.check(self.L.luxb200_rns_keyswitch_inner(self.h, _l.ptr(digits), _l.ptr(evk_b), _l.ptr(evk_a), _l.ptr(out), cnt, nd, sizeQl))
        return out

    def ApproxModDown(self, x, t=0):
        """t = 0: CKKS; t > 0: the BGV branch with plaintext modulus t (dcrtpoly-impl.h:1003-1005, :1021-1023)."""
        x = np.ascontiguousarray(x, np.uint64)
        out = np.empty((x.shape[0], len(self.q), self.N), np.uint64)
        _l.check(self.L.luxb200_approx_mod_switch_bgv(self.h, _l.ptr(x), _l.ptr(out), x.shape[0], 1, int(t)))
        return out

    def close(self):
        if self.h:
            self.L.luxb200_rns_plan_destroy(self.h)
            self.h = None


class HybridKeySwitchPlan:
    """KeySwitchHYBRID::EvalKeySwitchPrecomputeCore + EvalFastKeySwitchCore (src/pke/lib/keyswitch/keyswitch-hybrid.cpp:328-436) for
    one polynomial over Ql = q[:sizeQl]: c [count][sizeQl][N] EVALUATION, the key's b / a vectors [num_digits][len(q) + len(p)][N] ->
    [count][2][sizeQl][N]; t = 0 CKKS, t > 0 BGV."""

    def __init__(self, N, q, p, alpha, sizeQl=None, device=0):
        self.L = _l.load()
        self.q = np.ascontiguousarray(q, np.uint64)
        self.p = np.ascontiguousarray(p, np.uint64)
        self.N, self.sizeQl = N, len(self.q) if sizeQl is None else sizeQl
        h = _l.VP()
        _l.check(self.L.luxb200_hks_plan_create(N, len(self.q), _l.ptr(self.q), len(self.p), _l.ptr(self.p), alpha, self.sizeQl, device,
                                                C.byref(h)))
        self.h = h
        self.num_digits = int(self.L.luxb200_hks_num_digits(self.h))

    def KeySwitch(self, c, evk_b, evk_a, t=0):
        c = np.ascontiguousarray(c, np.uint64)
        evk_b = np.ascontiguousarray(evk_b, np.uint64); evk_a = np.ascontiguousarray(evk_a, np.uint64)
        out = np.empty((c.shape[0], 2, self.sizeQl, self.N), np.uint64)
        _l.check(self.L.luxb200_hks_keyswitch(self.h, _l.ptr(c), _l.ptr(evk_b), _l.ptr(evk_a), _l.ptr(out), c.shape[0], int(t)))
        return out

    def KeySwitchDev(self, c_ptr, evk_b_ptr, evk_a_ptr, out_ptr, count, t=0):
        """device pointers (ints) on the plan's device; returns when the result is complete"""
        _l.check(self.L.luxb200_hks_keyswitch_dev(self.h, _l.ptr(c_ptr), _l.ptr(evk_b_ptr), _l.ptr(evk_a_ptr), _l.ptr(out_ptr), int(count), int(t)))

    def close(self):
        if self.h:
            self.L.luxb200_hks_plan_destroy(self.h)
            self.h = None
