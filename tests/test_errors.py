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
