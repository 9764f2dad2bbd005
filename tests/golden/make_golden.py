"""Generate the committed golden vectors from the REAL reference (oracle/_ref/libref_shim.so, built from
/root/reference by `make -C oracle ref`).  Run in the build container:  python tests/golden/make_golden.py

The reference's own tests hold no coefficient-level known answers for this path, so these
vectors — outputs of the unmodified reference on its own keys — are the KATs that pin the oracle and the CUDA path
wherever /root/reference is absent.  Small hand-made parameter sets keep the key material at a few MB.
"""
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(HERE))
from orclib import GATE, P, RefContext, ref  # noqa: E402

# (bits, cyclOrder, n, q, modKS, baseKS, baseG, baseR, numAutoKeys, keyDist[1 = UNIFORM_TERNARY])
GOLD_SETS = {
    "GINX": (27, 1024, 16, 512, 1024, 32, 512, 32, 4, 1),
    "LMKCDEY": (27, 1024, 16, 512, 1024, 32, 512, 32, 4, 1),
    "AP": (27, 1024, 6, 512, 1024, 32, 512, 8, 4, 1),
}


def splitmix(seed, n, mod):
    out = np.zeros(n, np.uint64)
    s = seed
    M = (1 << 64) - 1
    for i in range(n):
        s = (s + 0x9E3779B97F4A7C15) & M
        z = s
        z = ((z ^ (z >> 30)) * 0xBF58476D1CE4E5B9) & M
        z = ((z ^ (z >> 27)) * 0x94D049BB133111EB) & M
        z ^= z >> 31
        out[i] = z % mod
    return out


def make_ntt():
    R = ref()
    d = {}
    for bits, N, cnt in [(27, 1024, 4), (27, 512, 2), (28, 1024, 2), (29, 2048, 2), (37, 2048, 2), (60, 4096, 1),
                         (60, 16384, 1)]:
        Q = R.ref_last_prime(bits, 2 * N)
        psi = R.ref_root_of_unity(2 * N, Q)
        x = splitmix(0x4C555846 + N + bits, cnt * N, Q).reshape(cnt, N)
        X = x.copy()
        R.ref_ntt(Q, psi, N, P(X), cnt, 0, 1)
        back = X.copy()
        R.ref_ntt(Q, psi, N, P(back), cnt, 1, 1)
        assert (back == x).all()
        d[f"q_{bits}_{N}"] = np.array([Q, psi], np.uint64)
        d[f"x_{bits}_{N}"] = x
        d[f"X_{bits}_{N}"] = X
    np.savez_compressed(os.path.join(HERE, "ntt.npz"), **d)
    # moduli chains / roots
    q = np.zeros(32, np.uint64); psi = np.zeros(32, np.uint64)
    R.ref_dcrt_chain(65536, 32, 60, P(q), P(psi))
    prim = []
    for bits, m in [(27, 2048), (28, 2048), (27, 1024), (29, 4096), (37, 4096), (25, 2048), (50, 4096)]:
        Q = R.ref_last_prime(bits, m)
        prim.append([bits, m, Q, R.ref_root_of_unity(m, Q), R.ref_previous_prime(Q, m)])
    np.savez_compressed(os.path.join(HERE, "chain.npz"), dcrt_q=q, dcrt_psi=psi, primes=np.array(prim, np.uint64))


def make_binfhe(method):
    R = ref()
    rc = RefContext(None, method, custom=GOLD_SETS[method])
    rc.keygen()
    p = rc.p
    bsk = rc.dump_bsk()
    A, B = rc.dump_ksk()
    cnt = 8
    a1 = np.zeros((cnt, p.n), np.uint64); b1 = np.zeros(cnt, np.uint64); a2 = a1.copy(); b2 = b1.copy()
    for i in range(cnt):
        a1[i], b1[i] = rc.encrypt(i & 1)
        a2[i], b2[i] = rc.encrypt((i >> 1) & 1)
    d = dict(params=np.array(GOLD_SETS[method], np.uint64), sk=rc.sk, bsk=bsk.astype(np.uint32), kskA=A.astype(np.uint16),
             kskB=B.astype(np.uint16), a1=a1.astype(np.uint16), b1=b1, a2=a2.astype(np.uint16), b2=b2)
    for g in ["AND", "OR", "NAND", "NOR", "XOR", "XNOR"]:
        for ext in (0, 1):
            oa = np.zeros((cnt, p.N if ext else p.n), np.uint64); ob = np.zeros(cnt, np.uint64)
            assert R.ref_evalbingate(rc.h, GATE[g], P(a1), P(b1), P(a2), P(b2), P(oa), P(ob), cnt, ext, 8) == 0
            d[f"{g}_{ext}_a"] = oa.astype(np.uint32); d[f"{g}_{ext}_b"] = ob
            if not ext:
                dec = [rc.decrypt(oa[i], ob[i]) for i in range(cnt)]
                d[f"{g}_dec"] = np.array(dec)
    oa = np.zeros((cnt, p.n), np.uint64); ob = np.zeros(cnt, np.uint64)
    R.ref_bootstrap(rc.h, P(a1), P(b1), P(oa), P(ob), cnt, 4, 0, 8)
    d["BOOT_a"] = oa.astype(np.uint32); d["BOOT_b"] = ob
    # multi-input gates
    for g, k, pm in [("AND3", 3, 6), ("OR3", 3, 6), ("MAJORITY", 3, 4), ("AND4", 4, 8), ("OR4", 4, 8), ("CMUX", 3, 4)]:
        am = np.zeros((cnt, k, p.n), np.uint64); bm = np.zeros((cnt, k), np.uint64)
        for i in range(cnt):
            for j in range(k):
                am[i, j], bm[i, j] = rc.encrypt((i >> j) & 1, pm)
        oa = np.zeros((cnt, p.n), np.uint64); ob = np.zeros(cnt, np.uint64)
        assert R.ref_evalbingate_multi(rc.h, GATE[g], k, P(am), P(bm), P(oa), P(ob), cnt, pm, 8) == 0
        d[f"{g}_in_a"] = am.astype(np.uint16); d[f"{g}_in_b"] = bm
        d[f"{g}_a"] = oa.astype(np.uint32); d[f"{g}_b"] = ob
    # EvalFunc: negacyclic / periodic / arbitrary LUTs over Z_q with p = q / (2 beta)
    q = p.q
    beta = R.ref_get_beta(rc.h)
    pp = max(2, q // (2 * beta))
    step = q // pp
    luts = {
        "neg": np.array([(q // pp) * ((x // step) if (x // step) < pp // 2 else (pp - ((x // step) - pp // 2)) % pp)
                         for x in range(q)], np.uint64),
        "per": np.array([(q // pp) * ((x // step) % (pp // 2)) for x in range(q)], np.uint64),
        "arb": np.array([(q // pp) * (((x // step) ** 3) % pp) for x in range(q)], np.uint64),
    }
    fa = np.zeros((cnt, p.n), np.uint64); fb = np.zeros(cnt, np.uint64)
    for i in range(cnt):
        fa[i], fb[i] = rc.encrypt(i % pp, pp)
    d["F_in_a"] = fa.astype(np.uint16); d["F_in_b"] = fb
    for name, lut in luts.items():
        oa = np.zeros((cnt, p.n), np.uint64); ob = np.zeros(cnt, np.uint64)
        rcode = R.ref_evalfunc(rc.h, P(fa), P(fb), P(lut), q, P(oa), P(ob), cnt, q, 8)
        d[f"F_{name}_lut"] = lut; d[f"F_{name}_rc"] = np.array([rcode])
        d[f"F_{name}_a"] = oa.astype(np.uint32); d[f"F_{name}_b"] = ob
    # raw EvalAcc, decompose, keyswitch, modswitch
    acc = splitmix(77, 2 * p.N, p.Q)
    acc_out = acc.copy()
    R.ref_evalacc(rc.h, P(a1[0].copy()), q, P(acc_out))
    d["ACC_in"] = acc.astype(np.uint32); d["ACC_out"] = acc_out.astype(np.uint32)
    dec = np.zeros(p.rows * p.N, np.uint64)
    R.ref_decompose(rc.h, P(acc), P(dec))
    d["DEC_out"] = dec.astype(np.uint32)
    ksa = splitmix(78, 2 * p.N, p.qKS).reshape(2, p.N); ksb = splitmix(79, 2, p.qKS)
    koa = np.zeros((2, p.n), np.uint64); kob = np.zeros(2, np.uint64)
    R.ref_keyswitch(rc.h, P(ksa), P(ksb), P(koa), P(kob), 2)
    d["KS_in_a"] = ksa; d["KS_in_b"] = ksb; d["KS_a"] = koa; d["KS_b"] = kob
    msa = splitmix(80, p.N, p.Q).reshape(1, p.N); msb = splitmix(81, 1, p.Q)
    moa = np.zeros((1, p.N), np.uint64); mob = np.zeros(1, np.uint64)
    R.ref_modswitch(rc.h, p.qKS, p.Q, P(msa), P(msb), P(moa), P(mob), p.N, 1)
    d["MS_in"] = msa; d["MS_out"] = moa
    np.savez_compressed(os.path.join(HERE, f"binfhe_{method.lower()}.npz"), **d)
    rc.close()


if __name__ == "__main__":
    make_ntt()
    for m in GOLD_SETS:
        make_binfhe(m)
    for f in sorted(os.listdir(HERE)):
        if f.endswith(".npz"):
            print(f, os.path.getsize(os.path.join(HERE, f)) // 1024, "KiB")
