"""ctypes loaders for the two CPU checkers (TEST INFRASTRUCTURE):

  * oracle/_ref/liboracle.so   — plain-C restatement (oracle/oracle.c), always buildable
  * oracle/_ref/libref_shim.so — flat C ABI over the real reference built from /root/reference

Nothing in the product imports this module.
"""
import ctypes as C
import os
import subprocess

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
ORC_DIR = os.path.join(ROOT, "oracle")
OUT = os.path.join(ORC_DIR, "_ref")

u64p = C.c_void_p
VP = C.c_void_p
U64 = C.c_uint64
U32 = C.c_uint32
I64 = C.c_int64
INT = C.c_int

# BINFHE_PARAMSET / BINFHE_METHOD / BINGATE values (binfhe-constants.h:49-123)
PARAMSET = {"TOY": 0, "MEDIUM": 1, "STD128_AP": 2, "STD128": 3, "STD128_LMKCDEY": 21, "STD256": 15, "STD192": 9}
METHOD = {"AP": 1, "GINX": 2, "LMKCDEY": 3}
GATE = {"OR": 0, "AND": 1, "NOR": 2, "NAND": 3, "XOR": 4, "XNOR": 5, "MAJORITY": 6, "AND3": 7, "OR3": 8,
        "AND4": 9, "OR4": 10, "XOR_FAST": 11, "XNOR_FAST": 12, "CMUX": 13}


def P(x):
    assert x.flags["C_CONTIGUOUS"]
    return x.ctypes.data_as(C.c_void_p)


def _sig(fn, res, args):
    fn.restype = res
    fn.argtypes = args


_oracle = None
_ref = None


def oracle():
    """liboracle.so (built on demand; gcc only)."""
    global _oracle
    if _oracle is None:
        so = os.path.join(OUT, "liboracle.so")
        src = os.path.join(ORC_DIR, "oracle.c")
        if not os.path.exists(so) or os.path.getmtime(so) < os.path.getmtime(src):
            subprocess.check_call(["make", "-C", ORC_DIR, "port"], stdout=subprocess.DEVNULL)
        L = C.CDLL(so)
        _sig(L.orc_last_prime, U64, [U32, U64])
        _sig(L.orc_previous_prime, U64, [U64, U64])
        _sig(L.orc_root_of_unity, U64, [U64, U64])
        _sig(L.orc_dcrt_chain, INT, [U32, U32, U32, VP, VP])
        _sig(L.orc_ntt, INT, [U64, U64, U32, VP, U64, INT, INT])
        _sig(L.orc_ctx_create, VP, [U32, U32, U64, U64, U64, U32, U32, U32, U32, INT])
        _sig(L.orc_ctx_destroy, None, [VP])
        _sig(L.orc_ctx_params, INT, [VP, VP])
        _sig(L.orc_bsk_len, U64, [VP])
        _sig(L.orc_ksk_len, U64, [VP])
        _sig(L.orc_load_bsk, INT, [VP, VP])
        _sig(L.orc_load_ksk, INT, [VP, VP, VP])
        _sig(L.orc_keygen, INT, [VP, U64, INT, VP, VP, VP, VP])
        _sig(L.orc_encrypt, INT, [VP, VP, I64, U64, U64, U64, VP, VP])
        _sig(L.orc_decrypt, I64, [VP, VP, VP, U64, U32, U64, U64])
        _sig(L.orc_decompose, INT, [VP, VP, VP])
        _sig(L.orc_evalacc, INT, [VP, VP, U64, VP])
        _sig(L.orc_modswitch, INT, [U64, U64, VP, VP, U64])
        _sig(L.orc_keyswitch, INT, [VP, VP, U64, VP, VP])
        _sig(L.orc_bootstrap_core, INT, [VP, VP, U64, U64, VP, VP])
        _sig(L.orc_evalbingate, INT, [VP, INT, VP, VP, VP, VP, VP, VP, U64, INT, INT])
        _sig(L.orc_evalbingate_multi, INT, [VP, INT, U32, VP, VP, VP, VP, U64, U64, INT])
        _sig(L.orc_bootstrap, INT, [VP, VP, VP, VP, VP, U64, U64, INT, INT])
        _sig(L.orc_evalfunc, INT, [VP, VP, VP, VP, U64, U64, VP, VP, U64, U64, INT])
        _sig(L.orc_ctx_baseG, U64, [VP])
        _sig(L.orc_evalfloor, INT, [VP, VP, VP, VP, VP, U64, U64, U64, U32, INT])
        _sig(L.orc_evaldecomp_count, U32, [U64, U64, U64])
        _sig(L.orc_evalsign_decomp, INT, [VP, INT, VP, VP, VP, VP, VP, U64, U64, U64, INT, INT])
        for fn in (L.orc_approx_switch_crt_basis, L.orc_approx_mod_up, L.orc_approx_mod_down):
            _sig(fn, INT, [U32, U32, VP, U32, VP, VP, VP, U64])
        _sig(L.orc_approx_mod_down_t, INT, [U32, U32, VP, U32, VP, VP, VP, U64, U64])
        _sig(L.orc_keyswitch_inner, INT, [U32, U32, VP, U32, VP, U32, U32, VP, VP, VP, VP, U64])
        _sig(L.orc_hybrid_keyswitch, INT, [U32, U32, VP, U32, VP, U32, U32, VP, VP, VP, VP, U64, U64])
        _sig(L.orc_time_bingate, C.c_double, [VP, INT, VP, VP, VP, VP, VP, VP, U64, INT])
        _sig(L.orc_max_threads, INT, [])
        _oracle = L
    return _oracle


def ref_available():
    return os.path.exists(os.path.join(OUT, "libref_shim.so"))


def ref():
    """libref_shim.so — the real reference (prebuilt; built here by `make -C oracle ref`)."""
    global _ref
    if _ref is None:
        L = C.CDLL(os.path.join(OUT, "libref_shim.so"))
        _sig(L.ref_ctx_create, VP, [INT, INT])
        _sig(L.ref_ctx_destroy, None, [VP])
        _sig(L.ref_ctx_create_custom, VP, [U32, U32, U32, U32, U32, U32, U32, U32, U32, INT, INT])
        _sig(L.ref_ctx_params, INT, [VP, VP])
        _sig(L.ref_keygen, INT, [VP])
        _sig(L.ref_get_sk, INT, [VP, VP])
        _sig(L.ref_bsk_dims, INT, [VP, VP])
        _sig(L.ref_bsk_entry, INT, [VP, U32, U32, U32, VP])
        _sig(L.ref_ksk_dump, INT, [VP, VP, VP])
        _sig(L.ref_encrypt, INT, [VP, I64, U64, U64, VP, VP])
        _sig(L.ref_decrypt, I64, [VP, VP, U64, U32, U64, U64])
        _sig(L.ref_evalbingate, INT, [VP, INT, VP, VP, VP, VP, VP, VP, U64, INT, INT])
        _sig(L.ref_evalbingate_multi, INT, [VP, INT, U32, VP, VP, VP, VP, U64, U64, INT])
        _sig(L.ref_bootstrap, INT, [VP, VP, VP, VP, VP, U64, U64, INT, INT])
        _sig(L.ref_evalfunc, INT, [VP, VP, VP, VP, U64, VP, VP, U64, U64, INT])
        _sig(L.ref_get_beta, U64, [VP])
        _sig(L.ref_evalacc, INT, [VP, VP, U64, VP])
        _sig(L.ref_decompose, INT, [VP, VP, VP])
        _sig(L.ref_modswitch, INT, [VP, U64, U64, VP, VP, VP, VP, U32, U64])
        _sig(L.ref_keyswitch, INT, [VP, VP, VP, VP, VP, U64])
        _sig(L.ref_root_of_unity, U64, [U64, U64])
        _sig(L.ref_last_prime, U64, [U32, U64])
        _sig(L.ref_previous_prime, U64, [U64, U64])
        _sig(L.ref_ntt, INT, [U64, U64, U32, VP, U64, INT, INT])
        _sig(L.ref_dcrt_chain, INT, [U32, U32, U32, VP, VP])
        _sig(L.ref_dcrt_switchformat, INT, [U32, U32, U32, VP, U64, INT])
        _sig(L.ref_ctx_create_largeprec, VP, [INT, INT, U32, U32, INT])
        _sig(L.ref_btkey_bases, INT, [VP, VP, INT])
        _sig(L.ref_bsk_entry_base, INT, [VP, U32, U32, U32, U32, VP])
        _sig(L.ref_ksk_dump_base, INT, [VP, U32, VP, VP])
        _sig(L.ref_current_baseG, U32, [VP])
        _sig(L.ref_evalfloor, INT, [VP, VP, VP, VP, VP, U64, U64, U32, INT])
        _sig(L.ref_evalsign, INT, [VP, VP, VP, VP, VP, U64, U64, INT])
        _sig(L.ref_evaldecomp, INT, [VP, VP, VP, VP, VP, VP, U32, U64, U64])
        _sig(L.ref_rns_op, INT, [U32, U32, VP, U32, VP, VP, VP, U64, INT])
        _sig(L.ref_rns_moddown_t, INT, [U32, U32, VP, U32, VP, VP, VP, U64, U64])
        _sig(L.ref_keyswitch_inner, INT, [U32, U32, VP, U32, VP, U32, U32, VP, VP, VP, VP, U64])
        _sig(L.ref_hybrid_keyswitch, INT, [U32, U32, VP, U32, VP, U32, U32, VP, VP, VP, VP, U64, U64])
        _sig(L.ref_time_bingate, C.c_double, [VP, INT, VP, VP, VP, VP, VP, VP, U64, INT])
        _sig(L.ref_time_ntt, C.c_double, [U64, U64, U32, VP, U64, INT, INT])
        _sig(L.ref_max_threads, INT, [])
        _ref = L
    return _ref


class Params:
    """n N q Q qKS baseKS baseG baseR digitsG numAutoKeys keyDist psi digitsKS method"""
    FIELDS = ["n", "N", "q", "Q", "qKS", "baseKS", "baseG", "baseR", "digitsG", "numAutoKeys", "keyDist", "psi",
              "digitsKS", "method", "digitsR"]

    def __init__(self, arr):
        for k, v in zip(self.FIELDS, arr):
            setattr(self, k, int(v))
        self.rows = 2 * (self.digitsG - 1)

    def bsk_len(self):
        rg = self.rows * 2 * self.N
        if self.method == 2:
            return self.n * 2 * rg
        if self.method == 1:
            return self.n * self.baseR * self.digitsR * rg
        return self.n * rg + (self.numAutoKeys + 1) * (self.digitsG - 1) * 2 * self.N

    def ksk_rows(self):
        return self.N * self.baseKS * self.digitsKS


class RefContext:
    """The reference's BinFHEContext with keys, plus flat dumps of its keys in the canonical layout."""

    def __init__(self, paramset, method, custom=None, largeprec=None):
        """custom = (bits, cyclOrder, n, q, modKS, baseKS, baseG, baseR, numAutoKeys, keyDist) for a hand-made set;
        largeprec = (arbFunc, logQ, N, timeOptimization): GenerateBinFHEContext(set, arbFunc, logQ, N, GINX, timeOpt),
        binfhecontext.cpp:53-104 (set TOY or STD128)."""
        self.L = ref()
        if largeprec is not None:
            self.h = self.L.ref_ctx_create_largeprec(PARAMSET[paramset], *[int(x) for x in largeprec])
        elif custom is not None:
            self.h = self.L.ref_ctx_create_custom(*custom, METHOD[method])
        else:
            self.h = self.L.ref_ctx_create(PARAMSET[paramset], METHOD[method])
        assert self.h, "reference rejected the parameter set / method"
        arr = np.zeros(16, np.uint64)
        self.L.ref_ctx_params(self.h, P(arr))
        self.p = Params(arr)
        # digitsR as the reference computes it (rgsw-cryptoparameters.cpp:39-40)
        import math
        self.p.digitsR = int(math.ceil(math.log(float(self.p.q)) / math.log(float(self.p.baseR))))
        self.sk = None

    def keygen(self):
        assert self.L.ref_keygen(self.h) == 0
        self.sk = np.zeros(self.p.n, np.int64)
        self.L.ref_get_sk(self.h, P(self.sk))

    def dump_bsk(self):
        p = self.p
        out = np.zeros(p.bsk_len(), np.uint64)
        rg = p.rows * 2 * p.N
        if p.method == 2:      # GINX: ek[0][s][i]
            for i in range(p.n):
                for s in range(2):
                    off = (i * 2 + s) * rg
                    assert self.L.ref_bsk_entry(self.h, 0, s, i, P(out[off:off + rg])) == p.rows
        elif p.method == 1:    # AP: ek[i][j][k], j = 0 empty
            for i in range(p.n):
                for j in range(1, p.baseR):
                    for k in range(p.digitsR):
                        off = ((i * p.baseR + j) * p.digitsR + k) * rg
                        assert self.L.ref_bsk_entry(self.h, i, j, k, P(out[off:off + rg])) == p.rows
        else:                  # LMKCDEY: ek[0][0][i]; auto keys ek[0][1][0..w]
            for i in range(p.n):
                assert self.L.ref_bsk_entry(self.h, 0, 0, i, P(out[i * rg:(i + 1) * rg])) == p.rows
            ag = (p.digitsG - 1) * 2 * p.N
            for t in range(p.numAutoKeys + 1):
                off = p.n * rg + t * ag
                assert self.L.ref_bsk_entry(self.h, 0, 1, t, P(out[off:off + ag])) == p.digitsG - 1
        return out

    def dump_ksk(self):
        p = self.p
        A = np.zeros(p.ksk_rows() * p.n, np.uint64)
        B = np.zeros(p.ksk_rows(), np.uint64)
        self.L.ref_ksk_dump(self.h, P(A), P(B))
        return A, B

    def bases(self):
        """gadget bases of the key map (one after a plain BTKeyGen, three with timeOptimization)"""
        out = np.zeros(4, np.uint32)
        k = self.L.ref_btkey_bases(self.h, P(out), 4)
        return [int(x) for x in out[:k]]

    def dump_keys_of_base(self, baseG):
        """(bsk, kskA, kskB, Params) of the key set generated for gadget base `baseG` (GINX layout)"""
        import copy, math
        p = copy.copy(self.p)
        p.baseG = baseG
        p.digitsG = int(math.ceil(math.log(float(p.Q)) / math.log(float(baseG))))
        p.rows = 2 * (p.digitsG - 1)
        rg = p.rows * 2 * p.N
        bsk = np.zeros(p.bsk_len(), np.uint64)
        for i in range(p.n):
            for s in range(2):
                off = (i * 2 + s) * rg
                assert self.L.ref_bsk_entry_base(self.h, baseG, 0, s, i, P(bsk[off:off + rg])) == p.rows
        A = np.zeros(p.ksk_rows() * p.n, np.uint64)
        B = np.zeros(p.ksk_rows(), np.uint64)
        self.L.ref_ksk_dump_base(self.h, baseG, P(A), P(B))
        return bsk, A, B, p

    def encrypt(self, m, pmod=4, mod=None):
        p = self.p
        a = np.zeros(p.n, np.uint64)
        b = np.zeros(1, np.uint64)
        self.L.ref_encrypt(self.h, int(m), pmod, mod or p.q, P(a), P(b))
        return a, int(b[0])

    def decrypt(self, a, b, pmod=4, mod=None):
        a = np.ascontiguousarray(a, np.uint64)
        return int(self.L.ref_decrypt(self.h, P(a), int(b), len(a), mod or self.p.q, pmod))

    def close(self):
        if self.h:
            self.L.ref_ctx_destroy(self.h)
            self.h = None


class OrcContext:
    def __init__(self, p):
        """p: Params (from the reference or hand-made)."""
        self.L = oracle()
        self.h = self.L.orc_ctx_create(p.n, p.N, p.q, p.Q, p.qKS, p.baseKS, p.baseG, p.baseR, p.numAutoKeys, p.method)
        arr = np.zeros(16, np.uint64)
        self.L.orc_ctx_params(self.h, P(arr))
        self.p = Params(arr)
        self.p.keyDist = getattr(p, "keyDist", 0)

    def load(self, bsk, kskA, kskB):
        assert len(bsk) == self.L.orc_bsk_len(self.h), (len(bsk), self.L.orc_bsk_len(self.h))
        self.L.orc_load_bsk(self.h, P(bsk))
        self.L.orc_load_ksk(self.h, P(kskA), P(kskB))

    def keygen(self, seed=1, key_dist=0):
        p = self.p
        sk = np.zeros(p.n, np.int64)
        bsk = np.zeros(p.bsk_len(), np.uint64)
        A = np.zeros(p.ksk_rows() * p.n, np.uint64)
        B = np.zeros(p.ksk_rows(), np.uint64)
        self.L.orc_keygen(self.h, seed, key_dist, P(sk), P(bsk), P(A), P(B))
        self.load(bsk, A, B)
        self.sk = sk
        return sk, bsk, A, B

    def encrypt(self, sk, m, seed, pmod=4, mod=None):
        a = np.zeros(self.p.n, np.uint64)
        b = np.zeros(1, np.uint64)
        self.L.orc_encrypt(self.h, P(sk), int(m), pmod, mod or self.p.q, seed, P(a), P(b))
        return a, int(b[0])

    def decrypt(self, sk, a, b, pmod=4, mod=None):
        a = np.ascontiguousarray(a, np.uint64)
        return int(self.L.orc_decrypt(self.h, P(sk), P(a), int(b), len(a), mod or self.p.q, pmod))

    def evalbingate(self, gate, a1, b1, a2, b2, extended=False, threads=8):
        p = self.p
        cnt = len(b1)
        oa = np.zeros((cnt, p.N if extended else p.n), np.uint64)
        ob = np.zeros(cnt, np.uint64)
        self.L.orc_evalbingate(self.h, GATE[gate], P(a1), P(b1), P(a2), P(b2), P(oa), P(ob), cnt, int(extended),
                               threads)
        return oa, ob

    def close(self):
        if self.h:
            self.L.orc_ctx_destroy(self.h)
            self.h = None


def hand_params(n, N, q, Q, qKS, baseKS, baseG, baseR, numAutoKeys, method):
    p = Params([n, N, q, Q, qKS, baseKS, baseG, baseR, 0, numAutoKeys, 0, 0, 0, method, 0])
    return p
