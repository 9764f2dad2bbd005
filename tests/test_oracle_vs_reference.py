"""CPU: pin the plain-C oracle against the REAL reference (oracle/_ref/libref_shim.so) on the reference's own
fresh keys.  Skipped where the prebuilt reference library is absent."""
import numpy as np
import pytest

import orclib
from orclib import GATE, OrcContext, P, RefContext

pytestmark = pytest.mark.skipif(not orclib.ref_available(), reason="oracle/_ref/libref_shim.so not built")


def test_primes_roots_ntt():
    R, O = orclib.ref(), orclib.oracle()
    rng = np.random.default_rng(1)
    for bits, N in [(27, 1024), (29, 2048), (37, 2048), (60, 4096)]:
        Q = R.ref_last_prime(bits, 2 * N)
        assert Q == O.orc_last_prime(bits, 2 * N)
        psi = R.ref_root_of_unity(2 * N, Q)
        assert psi == O.orc_root_of_unity(2 * N, Q)
        x = rng.integers(0, Q, size=(3, N), dtype=np.uint64)
        a, b = x.copy(), x.copy()
        R.ref_ntt(Q, psi, N, P(a), 3, 0, 1); O.orc_ntt(Q, psi, N, P(b), 3, 0, 1)
        assert (a == b).all()
        R.ref_ntt(Q, psi, N, P(a), 3, 1, 1); O.orc_ntt(Q, psi, N, P(b), 3, 1, 1)
        assert (a == b).all() and (a == x).all()


@pytest.mark.parametrize("method", ["GINX", "AP", "LMKCDEY"])
def test_toy_gates_bootstrap_func(method):
    R, O = orclib.ref(), orclib.oracle()
    rc = RefContext("TOY", method)
    rc.keygen()
    oc = OrcContext(rc.p)
    assert (oc.p.digitsG, oc.p.digitsKS, oc.p.psi) == (rc.p.digitsG, rc.p.digitsKS, rc.p.psi)
    oc.load(rc.dump_bsk(), *rc.dump_ksk())
    p, cnt = rc.p, 4
    a1 = np.zeros((cnt, p.n), np.uint64); b1 = np.zeros(cnt, np.uint64); a2 = a1.copy(); b2 = b1.copy()
    for i in range(cnt):
        a1[i], b1[i] = rc.encrypt(i & 1); a2[i], b2[i] = rc.encrypt((i >> 1) & 1)
    truth = {"AND": lambda x, y: x & y, "OR": lambda x, y: x | y, "XOR": lambda x, y: x ^ y,
             "NAND": lambda x, y: 1 - (x & y)}
    for g in truth:
        for ext in (0, 1):
            oa = np.zeros((cnt, p.N if ext else p.n), np.uint64); ob = np.zeros(cnt, np.uint64)
            assert R.ref_evalbingate(rc.h, GATE[g], P(a1), P(b1), P(a2), P(b2), P(oa), P(ob), cnt, ext, 4) == 0
            xa, xb = oc.evalbingate(g, a1, b1, a2, b2, bool(ext), threads=4)
            assert (oa == xa).all() and (ob == xb).all(), (method, g, ext)
            if not ext:  # the reference's own test tier: decrypt == truth table (UnitTestFHEW.cpp:380-420)
                assert [rc.decrypt(xa[i], xb[i]) for i in range(cnt)] == [truth[g](i & 1, (i >> 1) & 1) for i in range(cnt)]
    oa = np.zeros((cnt, p.n), np.uint64); ob = np.zeros(cnt, np.uint64); xa = oa.copy(); xb = ob.copy()
    R.ref_bootstrap(rc.h, P(a1), P(b1), P(oa), P(ob), cnt, 4, 0, 4)
    O.orc_bootstrap(oc.h, P(a1), P(b1), P(xa), P(xb), cnt, 4, 0, 4)
    assert (oa == xa).all() and (ob == xb).all()
    # a ciphertext that claims plaintext modulus 8: the table is still built for 4 (the copy at binfhe-base-scheme.cpp:200
    # drops m_p), the offset uses 8 (:214)
    R.ref_bootstrap(rc.h, P(a1), P(b1), P(oa), P(ob), cnt, 8, 0, 4)
    O.orc_bootstrap(oc.h, P(a1), P(b1), P(xa), P(xb), cnt, 8, 0, 4)
    assert (oa == xa).all() and (ob == xb).all()
    q, beta = p.q, R.ref_get_beta(rc.h)
    pp = max(2, q // (2 * beta)); step = q // pp
    for f in (lambda m: m % (pp // 2), lambda m: (m * m * m) % pp):
        lut = np.array([(q // pp) * f(x // step) for x in range(q)], np.uint64)
        r1 = R.ref_evalfunc(rc.h, P(a1), P(b1), P(lut), q, P(oa), P(ob), cnt, q, 4)
        r2 = O.orc_evalfunc(oc.h, P(a1), P(b1), P(lut), q, beta, P(xa), P(xb), cnt, q, 4)
        assert r1 == r2 and (r1 != 0 or ((oa == xa).all() and (ob == xb).all()))
    rc.close(); oc.close()


def test_dcrt_switchformat_is_per_limb_ntt():
    """DCRTPolyImpl::SwitchFormat (dcrtpoly-impl.h:1953-1959) of the real reference == the oracle's NTT applied limb by limb
    with the ILDCRTParams chain (ildcrtparams.h:100-118) — the shape luxb200_ntt_batch is tested against on the GPU."""
    R, O = orclib.ref(), orclib.oracle()
    for N, L, bits, cnt in [(4096, 3, 60, 2), (65536, 4, 60, 1), (2048, 2, 50, 2)]:
        q = np.zeros(L, np.uint64); psi = np.zeros(L, np.uint64); q2 = q.copy(); psi2 = psi.copy()
        R.ref_dcrt_chain(N, L, bits, P(q), P(psi)); O.orc_dcrt_chain(N, L, bits, P(q2), P(psi2))
        assert (q == q2).all() and (psi == psi2).all()
        rng = np.random.default_rng(N + L)
        x = np.stack([np.stack([rng.integers(0, int(q[l]), size=N, dtype=np.uint64) for l in range(L)]) for _ in range(cnt)])
        a, b = x.copy(), x.copy()
        assert R.ref_dcrt_switchformat(N, L, bits, P(a), cnt, 0) == 0
        for i in range(cnt):
            for l in range(L):
                O.orc_ntt(int(q[l]), int(psi[l]), N, P(b[i, l]), 1, 0, 1)
        assert (a == b).all()
        assert R.ref_dcrt_switchformat(N, L, bits, P(a), cnt, 1) == 0
        assert (a == x).all()


@pytest.mark.parametrize("method", ["GINX", "AP", "LMKCDEY"])
@pytest.mark.parametrize("bits,cyc,baseG", [(54, 4096, 1 << 14), (37, 4096, 1 << 19), (29, 4096, 64), (27, 2048, 32)])
def test_wide_regime_pins_the_oracle(method, bits, cyc, baseG):
    """The regimes the shared-memory GPU kernel serves — N = 2048, Q up to 54 bits, 1 to 5 gadget digits — pinned on the
    CPU: the real reference with a hand-made parameter set (short LWE key, so it runs in seconds) against the oracle on
    the reference's own keys: gates in both output formats and Bootstrap."""
    R, O = orclib.ref(), orclib.oracle()
    N = cyc // 2
    q = 2 * N if method == "LMKCDEY" else N
    rc = RefContext(None, method, custom=(bits, cyc, 12, q, 1 << 20, 32, baseG, 8, 6, 0))
    rc.keygen()
    oc = OrcContext(rc.p)
    assert (oc.p.digitsG, oc.p.digitsKS, oc.p.psi, oc.p.Q) == (rc.p.digitsG, rc.p.digitsKS, rc.p.psi, rc.p.Q)
    oc.load(rc.dump_bsk(), *rc.dump_ksk())
    p, cnt = rc.p, 3
    a1 = np.zeros((cnt, p.n), np.uint64); b1 = np.zeros(cnt, np.uint64); a2 = a1.copy(); b2 = b1.copy()
    for i in range(cnt):
        a1[i], b1[i] = rc.encrypt(i & 1); a2[i], b2[i] = rc.encrypt((i >> 1) & 1)
    for g in ("NAND", "XOR"):
        for ext in (0, 1):
            oa = np.zeros((cnt, p.N if ext else p.n), np.uint64); ob = np.zeros(cnt, np.uint64)
            assert R.ref_evalbingate(rc.h, GATE[g], P(a1), P(b1), P(a2), P(b2), P(oa), P(ob), cnt, ext, 4) == 0
            xa, xb = oc.evalbingate(g, a1, b1, a2, b2, bool(ext), threads=4)
            assert (oa == xa).all() and (ob == xb).all(), (method, g, ext)
    oa = np.zeros((cnt, p.n), np.uint64); ob = np.zeros(cnt, np.uint64); xa = oa.copy(); xb = ob.copy()
    R.ref_bootstrap(rc.h, P(a1), P(b1), P(oa), P(ob), cnt, 4, 0, 4)
    O.orc_bootstrap(oc.h, P(a1), P(b1), P(xa), P(xb), cnt, 4, 0, 4)
    assert (oa == xa).all() and (ob == xb).all()
    rc.close(); oc.close()


def _largeprec_pair(logQ, time_opt, arb=False):
    """reference large-precision context (TOY lattice, binfhecontext.cpp:53-104) + one oracle context per gadget base of
    its key map, loaded with the reference's own keys; contexts ordered current base first"""
    rc = RefContext("TOY", "GINX", largeprec=(arb, logQ, 0, time_opt))
    rc.keygen()
    cur = orclib.ref().ref_current_baseG(rc.h)
    bases = rc.bases()
    assert cur in bases and len(bases) == (3 if time_opt and logQ != 11 else 1)
    ocs = []
    for bg in [cur] + [x for x in bases if x != cur]:
        bsk, A, B, p = rc.dump_keys_of_base(bg)
        oc = OrcContext(p)
        assert oc.p.digitsG == p.digitsG and oc.p.Q == rc.p.Q and oc.p.psi == rc.p.psi
        oc.load(bsk, A, B)
        ocs.append(oc)
    return rc, ocs


@pytest.mark.parametrize("logQ,time_opt", [(17, True), (23, False), (27, True), (13, False)])
def test_largeprec_floor_sign_decomp_pin_the_oracle(logQ, time_opt):
    """EvalFloor / EvalSign / EvalDecomp (binfhe-base-scheme.cpp:335-521; the cases of UnitTestFunc.cpp:45-266: messages
    around p/2 at Q = 2^logQ) — the reference's ciphertexts against the oracle's, coefficient for coefficient, with one
    key and with the three-key dynamic mode."""
    import ctypes as C
    R, O = orclib.ref(), orclib.oracle()
    rc, ocs = _largeprec_pair(logQ, time_opt)
    p = rc.p
    Qbig, q, beta = 1 << logQ, p.q, R.ref_get_beta(rc.h)
    pl = (q // (2 * beta)) * (Qbig // q)      # GetMaxPlaintextSpace() * factor
    cnt = 6
    a = np.zeros((cnt, p.n), np.uint64); b = np.zeros(cnt, np.uint64)
    for i in range(cnt):
        a[i], b[i] = rc.encrypt(pl // 2 + i - 3, pl, Qbig)
    arr = (C.c_void_p * len(ocs))(*[oc.h for oc in ocs])
    # floor (current base only)
    oa = np.zeros_like(a); ob = np.zeros_like(b); xa = oa.copy(); xb = ob.copy()
    assert R.ref_evalfloor(rc.h, P(a), P(b), P(oa), P(ob), cnt, Qbig, 0, 4) == 0
    assert O.orc_evalfloor(ocs[0].h, P(a), P(b), P(xa), P(xb), cnt, Qbig, beta, 0, 4) == 0
    assert (oa == xa).all() and (ob == xb).all()
    # sign, both encodings
    for ss in (0, 1):
        assert R.ref_evalsign(rc.h, P(a), P(b), P(oa), P(ob), cnt, Qbig, ss) == 0
        assert O.orc_evalsign_decomp(arr, len(ocs), P(a), P(b), P(xa), P(xb), None, cnt, Qbig, beta, ss, 4) == 1
        assert (oa == xa).all() and (ob == xb).all(), ss
    assert R.ref_evalsign(rc.h, P(a), P(b), P(oa), P(ob), cnt, Qbig, 0) == 0
    assert [rc.decrypt(oa[i], ob[i], 2, q) for i in range(cnt)] == [int(i >= 3) for i in range(cnt)]
    # digit decomposition
    k = O.orc_evaldecomp_count(q, Qbig, beta)
    da = np.zeros((k, cnt, p.n), np.uint64); db = np.zeros((k, cnt), np.uint64); dm = np.zeros(k, np.uint64)
    ya = da.copy(); yb = db.copy(); ym = dm.copy()
    assert R.ref_evaldecomp(rc.h, P(a), P(b), P(da), P(db), P(dm), k, cnt, Qbig) == k
    assert O.orc_evalsign_decomp(arr, len(ocs), P(a), P(b), P(ya), P(yb), P(ym), cnt, Qbig, beta, 2, 4) == k
    assert (dm == ym).all() and (da == ya).all() and (db == yb).all()
    rc.close()
    for oc in ocs:
        oc.close()


@pytest.mark.parametrize("N,sizeQ,sizeP,bits", [(1024, 3, 2, 50), (4096, 4, 5, 60), (2048, 1, 3, 45)])
def test_rns_basis_conversion_pins_the_oracle(N, sizeQ, sizeP, bits):
    """DCRTPolyImpl::ApproxSwitchCRTBasis / ApproxModUp / ApproxModDown (dcrtpoly-impl.h:910-1030) of the real reference, fed
    the tables of CryptoParametersRNS::PrecomputeCRTTables, against the oracle's restatement, every residue."""
    R, O = orclib.ref(), orclib.oracle()
    chain = np.zeros(sizeQ + sizeP, np.uint64); psi = chain.copy()
    O.orc_dcrt_chain(N, sizeQ + sizeP, bits, P(chain), P(psi))
    q, p = np.ascontiguousarray(chain[:sizeQ]), np.ascontiguousarray(chain[sizeQ:])
    rng = np.random.default_rng(N + sizeQ)
    cnt = 2
    x = np.stack([np.stack([rng.integers(0, int(m), size=N, dtype=np.uint64) for m in q]) for _ in range(cnt)])
    x[0, 0, :] = q[0] - np.uint64(1)
    xqp = np.stack([np.stack([rng.integers(0, int(m), size=N, dtype=np.uint64) for m in chain]) for _ in range(cnt)])
    for what, (inp, lout, fn) in enumerate([(x, sizeP, O.orc_approx_switch_crt_basis), (x, sizeQ + sizeP, O.orc_approx_mod_up),
                                            (xqp, sizeQ, O.orc_approx_mod_down)]):
        a = np.zeros((cnt, lout, N), np.uint64); b = a.copy()
        assert R.ref_rns_op(N, sizeQ, P(q), sizeP, P(p), P(inp), P(a), cnt, what) == 0
        assert fn(N, sizeQ, P(q), sizeP, P(p), P(inp), P(b), cnt) == 0
        assert (a == b).all(), what
    # the inner product of EvalFastKeySwitchCoreExt (keyswitch-hybrid.cpp:438-481) on a truncated basis Ql u P
    nd, sizeQl = 3, max(1, sizeQ - 1)
    L, LK = sizeQl + sizeP, sizeQ + sizeP
    mods_l = list(q[:sizeQl]) + list(p)
    dg = np.stack([np.stack([np.stack([rng.integers(0, int(m), size=N, dtype=np.uint64) for m in mods_l]) for _ in range(nd)]) for _ in range(cnt)])
    dg[0, 0, 0, :] = q[0] - np.uint64(1)
    kb = np.stack([np.stack([rng.integers(0, int(m), size=N, dtype=np.uint64) for m in chain]) for _ in range(nd)])
    ka = np.stack([np.stack([rng.integers(0, int(m), size=N, dtype=np.uint64) for m in chain]) for _ in range(nd)])
    kb[0, 0, :] = q[0] - np.uint64(1)
    a = np.zeros((cnt, 2, L, N), np.uint64); b = a.copy()
    assert R.ref_keyswitch_inner(N, sizeQ, P(q), sizeP, P(p), sizeQl, nd, P(dg), P(kb), P(ka), P(a), cnt) == 0
    assert O.orc_keyswitch_inner(N, sizeQ, P(q), sizeP, P(p), sizeQl, nd, P(dg), P(kb), P(ka), P(b), cnt) == 0
    assert (a == b).all()
    # the whole hybrid key switch of one polynomial (EvalKeySwitchPrecomputeCore + EvalFastKeySwitchCore), CKKS and BGV, on the full
    # chain and on a truncated one, with part sizes that do and do not divide the limb count
    for alpha, sizeQl2, t in ((1, sizeQ, 0), (2, sizeQ, 65537), (2, max(1, sizeQ - 1), 0), (sizeQ, sizeQ, 0)):
        nd2 = min((sizeQl2 + alpha - 1) // alpha, (sizeQ + alpha - 1) // alpha)
        cin = np.stack([np.stack([rng.integers(0, int(m), size=N, dtype=np.uint64) for m in q[:sizeQl2]]) for _ in range(cnt)])
        kb2 = np.stack([np.stack([rng.integers(0, int(m), size=N, dtype=np.uint64) for m in chain]) for _ in range(nd2)])
        ka2 = np.stack([np.stack([rng.integers(0, int(m), size=N, dtype=np.uint64) for m in chain]) for _ in range(nd2)])
        a = np.zeros((cnt, 2, sizeQl2, N), np.uint64); b = a.copy()
        assert R.ref_hybrid_keyswitch(N, sizeQ, P(q), sizeP, P(p), alpha, sizeQl2, P(cin), P(kb2), P(ka2), P(a), cnt, t) == 0
        assert O.orc_hybrid_keyswitch(N, sizeQ, P(q), sizeP, P(p), alpha, sizeQl2, P(cin), P(kb2), P(ka2), P(b), cnt, t) == 0
        assert (a == b).all(), (alpha, sizeQl2, t)
    # the BGV branch of ApproxModDown (plaintext modulus t > 0; tables of bgvrns-cryptoparameters.cpp:78-95)
    for t in (65537, 2, (1 << 40) + 15):
        a = np.zeros((cnt, sizeQ, N), np.uint64); b = a.copy()
        assert R.ref_rns_moddown_t(N, sizeQ, P(q), sizeP, P(p), P(xqp), P(a), cnt, t) == 0
        assert O.orc_approx_mod_down_t(N, sizeQ, P(q), sizeP, P(p), P(xqp), P(b), cnt, t) == 0
        assert (a == b).all(), t
