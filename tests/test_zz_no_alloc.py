"""luxb200_approx_switch_crt_basis_static — the allocation-free variant of luxb200_approx_switch_crt_basis (device staging in
a static __device__ array of LUXB200_STATIC_SCRATCH_BYTES instead of cudaMalloc'ed buffers) — through the C ABI, against the
existing entry point and the oracle, bit-exact (integer path), on two calls in a row that share the array.  Named test_zz_* so
that pytest collects it last: a failure here must not hide the other tests."""
import ctypes as C

import numpy as np
import pytest

from orclib import P, oracle

pytestmark = pytest.mark.gpu

SCRATCH_BYTES = 16 << 20   # LUXB200_STATIC_SCRATCH_BYTES, include/luxb200.h


def _plan(L, N, q, p, device=0):
    from fhe_b200 import lib
    h = lib.VP()
    lib.check(L.luxb200_rns_plan_create(N, len(q), lib.ptr(q), len(p), lib.ptr(p), device, C.byref(h)))
    return h


def test_static_scratch_basis_switch_matches_allocating_entry_point():
    from fhe_b200 import lib
    L = lib.load()
    O = oracle()
    N, sizeQ, sizeP, bits = 4096, 4, 5, 60
    chain = np.zeros(sizeQ + sizeP, np.uint64); psi = chain.copy()
    O.orc_dcrt_chain(N, sizeQ + sizeP, bits, P(chain), P(psi))
    q, p = np.ascontiguousarray(chain[:sizeQ]), np.ascontiguousarray(chain[sizeQ:])
    h = _plan(L, N, q, p)
    rng = np.random.default_rng(2024)

    def draw(mods, cnt):
        return np.stack([np.stack([rng.integers(0, int(m), size=N, dtype=np.uint64) for m in mods]) for _ in range(cnt)])

    # two calls in a row on the shared array: a larger Q -> P batch, then a smaller P -> Q one whose staging lies inside the
    # region the first call left its data in — stale contents must not leak into the result
    for cnt, direction in ((7, 0), (3, 1), (5, 0)):
        src, dst = (q, p) if direction == 0 else (p, q)
        x = draw(src, cnt)
        x[0, 0, :] = src[0] - np.uint64(1)
        x[-1, -1, :] = 0
        want = np.zeros((cnt, len(dst), N), np.uint64)
        lib.check(L.luxb200_approx_switch_crt_basis(h, lib.ptr(x), lib.ptr(want), cnt, direction))
        got = np.full((cnt, len(dst), N), 0xDEADBEEF, np.uint64)
        lib.check(L.luxb200_approx_switch_crt_basis_static(h, lib.ptr(x), lib.ptr(got), cnt, direction))
        assert (got == want).all(), f"static-scratch result differs from luxb200_approx_switch_crt_basis (count {cnt}, direction {direction})"
        ref = np.zeros_like(want)
        if direction == 0:
            O.orc_approx_switch_crt_basis(N, sizeQ, P(q), sizeP, P(p), P(x), P(ref), cnt)
        else:
            O.orc_approx_switch_crt_basis(N, sizeP, P(p), sizeQ, P(q), P(x), P(ref), cnt)
        assert (got == ref).all(), "static-scratch result differs from the oracle"
    # the same call twice gives the same words
    again = np.zeros_like(got)
    lib.check(L.luxb200_approx_switch_crt_basis_static(h, lib.ptr(x), lib.ptr(again), cnt, direction))
    assert (again == got).all()
    L.luxb200_rns_plan_destroy(h)


def test_static_scratch_limit_is_enforced():
    from fhe_b200 import lib
    L = lib.load()
    O = oracle()
    N, sizeQ, sizeP = 4096, 4, 5
    chain = np.zeros(sizeQ + sizeP, np.uint64); psi = chain.copy()
    O.orc_dcrt_chain(N, sizeQ + sizeP, 60, P(chain), P(psi))
    q, p = np.ascontiguousarray(chain[:sizeQ]), np.ascontiguousarray(chain[sizeQ:])
    h = _plan(L, N, q, p)
    per = (sizeQ + sizeP) * N * 8
    fit = SCRATCH_BYTES // per            # the largest batch that fits (per-ciphertext staging is a multiple of 256 bytes here)
    rng = np.random.default_rng(7)
    x = rng.integers(0, int(q.min()), size=(fit + 1, sizeQ, N), dtype=np.uint64)
    out = np.zeros((fit + 1, sizeP, N), np.uint64)
    want = np.zeros_like(out)
    # exactly at the limit: accepted and equal to the allocating entry point
    lib.check(L.luxb200_approx_switch_crt_basis_static(h, lib.ptr(x), lib.ptr(out), fit, 0))
    lib.check(L.luxb200_approx_switch_crt_basis(h, lib.ptr(x), lib.ptr(want), fit, 0))
    assert (out[:fit] == want[:fit]).all()
    # one more ciphertext: refused with a clear error, output untouched
    out[:] = 5
    rc = L.luxb200_approx_switch_crt_basis_static(h, lib.ptr(x), lib.ptr(out), fit + 1, 0)
    assert rc == -1, rc   # LUXB200_ERR_INVALID_ARG
    msg = L.luxb200_last_error().decode()
    assert "static scratch" in msg and str(SCRATCH_BYTES) in msg, msg
    assert (out == 5).all()
    # a count whose byte size overflows 64 bits is refused too
    assert L.luxb200_approx_switch_crt_basis_static(h, lib.ptr(x), lib.ptr(out), (1 << 63) + 1, 0) == -1
    L.luxb200_rns_plan_destroy(h)
