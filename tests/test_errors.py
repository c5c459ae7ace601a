"""Error behaviour of the C ABI (host logic in csrc/api.cpp, exercised through tests/emu's build of
the same file). The reference signals nothing (unchecked CUDA calls, `int` returns always 0); the
C ABI returns negative codes and a message, and never launches on bad arguments."""
import ctypes as C

import numpy as np
import pytest

from fluid_simulation_b200 import _lib
from tests import helpers as H

EINVAL = -1


@pytest.fixture(scope="module")
def lib():
    return _lib.load(H.build_emu())


def _slab(X=40, Y=50, pitch=64, row0=0, rows=None, dtype=0, hx=0.1, hy=0.1):
    return _lib.Slab(dtype, X, Y, pitch, row0, X if rows is None else rows, hx, hy)


def _buf(n):
    a = np.zeros(n + 8, np.float32)
    off = (-a.ctypes.data) % 16 // 4
    return a, a.ctypes.data + off * 4


def _msg(lib):
    return lib.fluidb200_last_error().decode()


def test_poisson_argument_checks(lib):
    g = _slab()
    edges = _lib.Edges()
    assert lib.fluidb200_cavity_edges(b"p", 1.0, C.byref(edges)) == 0
    keep, p = _buf(40 * 64)
    keep2, q = _buf(40 * 64)
    keep3, r = _buf(40 * 64)
    ok = lambda **kw: lib.fluidb200_poisson_sweeps(C.byref(kw.get("g", g)), kw.get("p", p), kw.get("q", q), r,
                                                   C.byref(edges), kw.get("mask"), kw.get("k", 2), kw.get("lo", 0),
                                                   kw.get("hi", 40), None)
    assert ok() == 0
    assert ok(q=p) == EINVAL and "alias" in _msg(lib)
    assert ok(k=0) == EINVAL and ok(k=9) == EINVAL and "sweeps" in _msg(lib)
    assert ok(p=p + 4) == EINVAL and "aligned" in _msg(lib)
    assert ok(lo=10, hi=10) == EINVAL
    assert ok(lo=0, hi=1) == EINVAL and "wall" in _msg(lib)
    assert ok(lo=39, hi=40) == EINVAL
    assert ok(g=_slab(pitch=50)) == EINVAL and "pitch" in _msg(lib)
    assert ok(g=_slab(X=2)) == EINVAL and "3x3" in _msg(lib)
    assert ok(g=_slab(dtype=7)) == EINVAL
    assert ok(g=_slab(hx=0.0)) == EINVAL
    # a slab that does not hold the halo rows the range needs
    assert ok(g=_slab(row0=10, rows=20), lo=10, hi=30) == EINVAL and "needs rows" in _msg(lib)
    assert ok(g=_slab(row0=10, rows=20), lo=12, hi=28, k=2) == 0
    mask = _lib.Mask(r, r)
    assert ok(mask=C.byref(mask), k=2) == 0 and ok(mask=C.byref(mask), k=1) == 0
    assert ok(mask=C.byref(_lib.Mask(r, None)), k=1) == EINVAL and "mask" in _msg(lib)
    bad = _lib.Edges()
    bad.mode[0] = 2     # ARRAY without arrays
    assert lib.fluidb200_poisson_sweeps(C.byref(g), p, q, r, C.byref(bad), None, 1, 0, 40, None) == EINVAL


def test_solver_handle_argument_checks(lib):
    h = C.c_void_p()
    assert lib.fluidb200_create(C.byref(h), 0, 2, 40, 1.0, 1.0) == EINVAL
    assert lib.fluidb200_create(C.byref(h), 5, 40, 40, 1.0, 1.0) == EINVAL
    assert lib.fluidb200_create(None, 0, 40, 40, 1.0, 1.0) == EINVAL
    assert lib.fluidb200_create(C.byref(h), 0, 40, 40, 1.0, 1.0) == 0
    assert lib.fluidb200_set_params(h, 0.001, 0.1, 1.0, -1, 1) == EINVAL
    assert lib.fluidb200_set_sweeps_per_launch(h, 0) == EINVAL
    assert lib.fluidb200_set_sweeps_per_launch(h, 9) == EINVAL
    out = np.zeros((40, 40), np.float32)
    assert lib.fluidb200_download(h, 99, out.ctypes.data) == EINVAL
    assert lib.fluidb200_download(h, 0, None) == EINVAL
    assert lib.fluidb200_solve(None) == EINVAL
    assert lib.fluidb200_destroy(h) == 0
    assert lib.fluidb200_destroy(None) == 0


def test_predictor_and_corrector_argument_checks(lib):
    g = _slab()
    eu, ev = _lib.Edges(), _lib.Edges()
    lib.fluidb200_cavity_edges(b"u", 1.0, C.byref(eu)); lib.fluidb200_cavity_edges(b"v", 1.0, C.byref(ev))
    bufs = [_buf(40 * 64) for _ in range(6)]
    u, v, us, vs, rhs, p = [b[1] for b in bufs]
    assert lib.fluidb200_predictor(C.byref(g), u, v, us, vs, rhs, 1e-3, 0.1, 1.0, 0, 40, 0, 40, None) == 0
    assert lib.fluidb200_predictor(C.byref(g), u, v, us, vs, rhs, 0.0, 0.1, 1.0, 0, 40, 0, 40, None) == EINVAL
    assert lib.fluidb200_predictor(C.byref(g), None, v, us, vs, rhs, 1e-3, 0.1, 1.0, 0, 40, 0, 40, None) == EINVAL
    assert lib.fluidb200_corrector(C.byref(g), p, us, vs, u, v, C.byref(eu), C.byref(ev), None, None, 1e-3, 1.0, 0, 40, None) == 0
    assert lib.fluidb200_corrector(C.byref(g), p, us, vs, u, v, C.byref(eu), C.byref(ev), None, None, 1e-3, 0.0, 0, 40, None) == EINVAL
    assert lib.fluidb200_corrector(C.byref(g), p, us, vs, u, v, None, C.byref(ev), None, None, 1e-3, 1.0, 0, 40, None) == EINVAL
    assert lib.fluidb200_cavity_edges(b"x", 1.0, C.byref(eu)) == EINVAL


def test_migrate_solve_argument_checks_and_plain_sequence(lib):
    """NULL arguments are refused; with interior flags or several timesteps the call is the plain
    upload + solve + download (same bits as those three calls)."""
    from oracle import Oracle
    h = C.c_void_p()
    assert lib.fluidb200_create(C.byref(h), 0, 30, 20, 1.0, 1.0) == 0
    bad = (C.c_void_p * 3)()
    assert lib.fluidb200_migrate_solve(h, None, bad) == EINVAL
    assert lib.fluidb200_migrate_solve(h, bad, bad) == EINVAL and "host_in[0]" in _msg(lib)
    assert lib.fluidb200_migrate_solve(None, bad, bad) == EINVAL
    rng = np.random.default_rng(0)
    st = [(0.01 * rng.standard_normal((30, 20))).astype(np.float32) for _ in range(3)]
    out = [np.empty((30, 20), np.float32) for _ in range(3)]
    assert lib.fluidb200_upload_cavity(h, 1.0) == 0
    assert lib.fluidb200_set_params(h, 1e-4, 0.1, 1.0, 5, 2) == 0      # two timesteps: no chunk pipeline
    pi = (C.c_void_p * 3)(*[a.ctypes.data for a in st])
    po = (C.c_void_p * 3)(*[a.ctypes.data for a in out])
    assert lib.fluidb200_migrate_solve(h, pi, po) == 0
    o = Oracle(30, 20).set_cavity(1.0)
    for n, a in zip(("u", "v", "p"), st):
        getattr(o, n)[:] = a
    o.p_next[:] = o.p
    o.time_step, o.num_iterations, o.num_poisson_iterations = 1e-4, 2, 5
    o.solve()
    for n, a in zip(("u", "v", "p"), out):
        H.assert_bit_equal(a, getattr(o, n), f"migrate_solve/plain:{n}")
    assert lib.fluidb200_destroy(h) == 0


def test_multi_handle_argument_checks(lib):
    m = C.c_void_p()
    assert lib.fluidb200_multi_create(C.byref(m), 7, 40, 40, 1.0, 1.0, 1, None) == EINVAL
    assert lib.fluidb200_multi_create(C.byref(m), 0, 2, 40, 1.0, 1.0, 1, None) == EINVAL
    assert lib.fluidb200_multi_create(C.byref(m), 0, 40, 40, 1.0, 1.0, 0, None) == EINVAL
    assert lib.fluidb200_multi_create(C.byref(m), 0, 40, 40, 1.0, 1.0, 8, None) == EINVAL and "too thin" in _msg(lib)
    assert lib.fluidb200_multi_create(None, 0, 40, 40, 1.0, 1.0, 1, None) == EINVAL
    assert lib.fluidb200_multi_create(C.byref(m), 0, 40, 40, 1.0, 1.0, 1, None) == 0
    assert lib.fluidb200_multi_num_slabs(m) == 1
    assert lib.fluidb200_multi_set_params(m, 0.001, 0.1, 1.0, -1, 1) == EINVAL
    assert lib.fluidb200_multi_set_params(m, 0.0, 0.1, 1.0, 5, 1) == 0 and lib.fluidb200_multi_solve(m) == EINVAL
    assert lib.fluidb200_multi_download(m, 9, C.c_void_p(1)) == EINVAL
    assert lib.fluidb200_multi_solve(None) == EINVAL and lib.fluidb200_multi_destroy(None) == 0
    assert lib.fluidb200_multi_destroy(m) == 0
