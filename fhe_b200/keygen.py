"""Synthetic key material for benchmarks and self-tests (host numpy + the GPU NTT through the C ABI).

In a deployment the keys come from the reference's BinFHEContext::BTKeyGen (binfhecontext.cpp:285-307) and are
handed over in the canonical flat layout (include/luxb200.h).  This module produces keys with the same structure
and distributions so that bench.py and smoke() are self-contained:
  KSK  LWEEncryptionScheme::KeySwitchGen   lwe-pke.cpp:264-344
  BSK  RingGSWAccumulatorCGGI::KeyGenAcc   rgsw-acc-cggi.cpp:39-96   (GINX)
       RingGSWAccumulatorDM::KeyGenAcc     rgsw-acc-dm.cpp:39-114    (AP)
       RingGSWAccumulatorLMKCDEY::KeyGenAcc rgsw-acc-lmkcdey.cpp:39-223
It is NOT bit-compatible with the reference's Blake2 PRNG (nothing on the bootstrap path depends on that).
"""
import numpy as np

from .context import NTTPlan

STD_DEV = 3.19


def _gauss(rng, shape, mod):
    e = np.rint(rng.normal(0.0, STD_DEV, size=shape)).astype(np.int64)
    return np.mod(e, mod).astype(np.uint64)


def _mulmod(a, b, Q):
    # Q < 2^31: products fit in uint64
    return (a * b) % np.uint64(Q)


class KeySet:
    def __init__(self, sk, skN, bsk, kskA, kskB):
        self.sk, self.skN, self.bsk, self.kskA, self.kskB = sk, skN, bsk, kskA, kskB


def keygen(ctx, seed=1, key_dist=None):
    """ctx: fhe_b200.context.BinFHEContext with parameters generated. Returns KeySet (canonical flat uint64)."""
    rng = np.random.default_rng(seed)
    n, N, Q, qKS, q = ctx.n, ctx.N, ctx.Q, ctx.qKS, ctx.q
    kd = getattr(ctx, "keyDist", 0) if key_dist is None else key_dist
    if kd == 1:
        sk = np.rint(rng.normal(0.0, STD_DEV, size=n)).astype(np.int64)
    else:
        sk = rng.integers(-1, 2, size=n).astype(np.int64)
    skN = rng.integers(-1, 2, size=N).astype(np.int64)

    # ---- KSK
    rowsK = ctx.ksk_rows()
    A = rng.integers(0, qKS, size=(rowsK, n), dtype=np.uint64)
    inner = (A.astype(np.int64) @ np.mod(sk, qKS).astype(np.int64)) % qKS
    i_idx = np.arange(N).repeat(ctx.baseKS * ctx.digitsKS)
    j_idx = np.tile(np.arange(ctx.baseKS).repeat(ctx.digitsKS), N)
    k_idx = np.tile(np.arange(ctx.digitsKS), N * ctx.baseKS)
    msg = (np.mod(skN, qKS)[i_idx].astype(object) * j_idx.astype(object) * (ctx.baseKS ** k_idx.astype(object))) % qKS
    B = (_gauss(rng, rowsK, qKS).astype(np.int64) + inner + msg.astype(np.int64)) % qKS
    kskA = A.reshape(-1)
    kskB = B.astype(np.uint64)

    # ---- BSK
    plan = NTTPlan(N, [Q], [ctx.psi], device=ctx.device)
    skNTT = plan.SwitchFormat(np.mod(skN, Q).astype(np.uint64).reshape(1, N), 0)[0]
    rows = ctx.rows
    Gpow = [pow(ctx.baseG, i, Q) for i in range(ctx.digitsG + 1)]

    def rgsw_batch(msgs):
        """msgs: list of (has_msg, mm, neg) -> array [len][rows][2][N] in EVALUATION format."""
        m = len(msgs)
        a = rng.integers(0, Q, size=(m, rows, N), dtype=np.uint64)
        e = _gauss(rng, (m, rows, N), Q)
        r0 = a.copy()
        r1 = e
        for t, (has, mm, neg) in enumerate(msgs):
            if not has:
                continue
            for i in range(rows):
                g = Gpow[(i >> 1) + 1]
                tgt = r1 if (i & 1) else r0
                tgt[t, i, mm] = (int(tgt[t, i, mm]) + (Q - g if neg else g)) % Q
        aN = plan.SwitchFormat(a.reshape(-1, N), 0).reshape(m, rows, N)
        r0 = plan.SwitchFormat(r0.reshape(-1, N), 0).reshape(m, rows, N)
        r1 = plan.SwitchFormat(r1.reshape(-1, N), 0).reshape(m, rows, N)
        r1 = (r1 + _mulmod(aN, skNTT[None, None, :], Q)) % np.uint64(Q)
        return np.stack([r0, r1], axis=2)

    def exponent(mv):
        mm = (((mv % q) + q) % q) * (2 * N // q)
        return (True, mm - N, True) if mm >= N else (True, mm, False)

    method = ctx.method
    if method == 2:      # GINX
        msgs = []
        for s in sk:
            msgs.append((s == 1, 0, False))
            msgs.append((s == -1, 0, False))
        bsk = rgsw_batch(msgs).reshape(-1)
    elif method == 1:    # AP
        out = np.zeros((n, ctx.baseR, ctx.digitsR, rows, 2, N), np.uint64)
        for i in range(n):
            msgs = [exponent(int(sk[i]) * j * (ctx.baseR ** k)) for j in range(1, ctx.baseR) for k in range(ctx.digitsR)]
            out[i, 1:] = rgsw_batch(msgs).reshape(ctx.baseR - 1, ctx.digitsR, rows, 2, N)
        bsk = out.reshape(-1)
    else:                # LMKCDEY
        main = rgsw_batch([exponent(int(s)) for s in sk]).reshape(-1)
        dg, M, w = ctx.digitsG - 1, 2 * N, ctx.numAutoKeys
        logN = N.bit_length() - 1
        brv = np.array([int(format(j, f"0{logN}b")[::-1], 2) for j in range(N)])
        autos = np.zeros((w + 1, dg, 2, N), np.uint64)
        for t in range(w + 1):
            k = M - 5 if t == 0 else pow(5, t, M)
            j = np.arange(N)
            idx = (((2 * j + 1) * k) % M) >> 1
            skA = np.zeros(N, np.uint64)
            skA[brv[j]] = skNTT[brv[idx]]           # poly-impl.h:345-356
            for d in range(dg):
                r0 = rng.integers(0, Q, size=N, dtype=np.uint64)
                e = plan.SwitchFormat(_gauss(rng, (1, N), Q), 0)[0]
                r1 = (e + np.uint64(Q) - _mulmod(skA, np.uint64(Gpow[d + 1]), Q) + _mulmod(r0, skNTT, Q)) % np.uint64(Q)
                autos[t, d, 0], autos[t, d, 1] = r0, r1
        bsk = np.concatenate([main, autos.reshape(-1)])
    plan.close()
    return KeySet(sk, skN, np.ascontiguousarray(bsk, np.uint64), np.ascontiguousarray(kskA), kskB)


def encrypt(ctx, sk, m, rng, p=4, mod=None):
    """LWEEncryptionScheme::Encrypt (lwe-pke.cpp:104-130) for a batch of messages m (array)."""
    mod = mod or ctx.q
    m = np.asarray(m, np.int64)
    a = rng.integers(0, mod, size=(len(m), ctx.n), dtype=np.uint64)
    e = np.rint(rng.normal(0.0, STD_DEV, size=len(m))).astype(np.int64)
    b = (np.mod(m, p) * (mod // p) + e + a.astype(np.int64) @ np.mod(sk, mod)) % mod
    return a, b.astype(np.uint64)


def decrypt(ctx, sk, a, b, p=4, mod=None):
    """LWEEncryptionScheme::Decrypt (lwe-pke.cpp:183-223)."""
    mod = mod or ctx.q
    a = np.asarray(a).astype(np.int64)
    r = (np.asarray(b).astype(np.int64) - a @ np.mod(sk, mod)[: a.shape[1]]) % mod
    r = (r + mod // (2 * p)) % mod
    return (p * r) // mod
