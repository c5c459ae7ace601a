"""CPU-only: the fp32 kernels replace the reference's divisions by a division-free fast path
with an exact fallback (Ar<float>::cdiff / lap_tail in csrc/warp_kernels.cuh). This hammers them
against the reference's own formulas evaluated in NumPy:

  (a - b) / 2. / h                      SerialNavierStokes.cpp:302-305  (float sub, double divs, one float rounding)
  (s - 4. * c) / hx / hy                SerialNavierStokes.cpp:309      (all double, one float rounding)

with random magnitudes over the whole float range, subnormals, zeros of both signs, exact ties
in the subnormal range, and spacings that are / are not powers of two."""
import ctypes as C

import numpy as np
import pytest

from tests import helpers as H


@pytest.fixture(scope="module")
def emu():
    lib = C.CDLL(H.build_emu())
    lib.emu_cdiff_f32.argtypes = [C.c_int, C.c_void_p, C.c_void_p, C.c_float, C.c_longlong, C.c_void_p]
    lib.emu_lap_tail_f32.argtypes = [C.c_int, C.c_void_p, C.c_void_p, C.c_float, C.c_float, C.c_longlong, C.c_void_p]
    lib.emu_metric_mode_f32.argtypes = [C.c_float, C.c_float]
    return lib


def _inputs(n, seed):
    rng = np.random.default_rng(seed)
    mant = rng.uniform(1.0, 2.0, n)
    expo = rng.integers(-149, 60, n)
    x = (mant * np.exp2(expo.astype(np.float64)) * rng.choice([-1.0, 1.0], n)).astype(np.float32)
    x[rng.integers(0, n, n // 50)] = 0.0
    x[rng.integers(0, n, n // 50)] = -0.0
    sub = rng.integers(1, 1 << 23, n // 20).astype(np.uint32).view(np.float32)   # true subnormals
    x[rng.integers(0, n, n // 20)] = sub
    return x


SPACINGS = [1.0, 0.5, 1.0 / 128, 1.0 / 99, 1.0 / 16383, 3.0, 0.1, 7.0 / 3, 1e-3, 123.456]


@pytest.mark.parametrize("h", SPACINGS)
def test_central_difference_matches_reference_formula(emu, h):
    n = 400_000
    a, b = _inputs(n, 1), _inputs(n, 2)
    # near-cancelling pairs and tiny differences (subnormal quotients, exact ties for h = 3)
    b[: n // 4] = a[: n // 4] * np.float32(1 + 2.0 ** -20)
    k = np.arange(1, 2001, dtype=np.float64)
    a[n // 2: n // 2 + 2000] = (k * 2.0 ** -149).astype(np.float32)
    b[n // 2: n // 2 + 2000] = 0.0
    h32 = np.float32(h)
    with np.errstate(all="ignore"):
        d = (a - b).astype(np.float32)
        want = (d.astype(np.float64) / 2.0 / np.float64(h32)).astype(np.float32)
    best = emu.emu_metric_mode_f32(h32, h32)     # the mode the product would pick for this spacing
    assert best == (2 if h in (1.0, 0.5, 1.0 / 128) else 1)
    for mode in range(best + 1):                 # ... and every more conservative one
        out = np.empty(n, np.float32)
        emu.emu_cdiff_f32(mode, a.ctypes.data, b.ctypes.data, h32, n, out.ctypes.data)
        ok = (out.view(np.uint32) == want.view(np.uint32)) | (np.isnan(out) & np.isnan(want))
        assert ok.all(), f"h={h} mode {mode}: {np.count_nonzero(~ok)} mismatches, first {a[~ok][:3]}, {b[~ok][:3]}"


def test_central_difference_over_many_spacings(emu):
    """The division-free central difference over 400 spacings drawn at random -- among them spacings one ulp
    away from a power of two and significands that are nearly all ones -- with numerators chosen so that the
    quotient lands on, or next to, the midpoint of two floats (d = RN32(m * 2h), m a midpoint): the cases a
    wrong last-bit correction would get wrong."""
    rng = np.random.default_rng(11)
    n = 60_000
    hs = list(np.exp2(rng.uniform(-20, 6, 380)).astype(np.float32))
    for e in (-7, 0, 3):
        one = np.float32(2.0 ** e)
        hs += [np.nextafter(one, np.float32(0)), np.nextafter(one, np.float32(4 * one)),
               np.nextafter(np.nextafter(one, np.float32(0)), np.float32(0))]
    hs += [np.float32(1.0 / 999), np.float32(1.0 / 40), np.float32(1.0 / 16383), np.float32(1.9999998), np.float32(1.9999996),
           np.float32(0.99999994 / 2), np.float32(3.0), np.float32(5.0), np.float32(0.3), np.float32(1.0 / 3), np.float32(1e-3)]
    bad = []
    for h32 in hs:
        h32 = np.float32(h32)
        if emu.emu_metric_mode_f32(h32, h32) != 1:
            continue
        a, b = _inputs(n, 100), np.zeros(n, np.float32)
        m = rng.uniform(1.0, 2.0, n // 2).astype(np.float32).astype(np.float64) + 2.0 ** -24      # midpoints of floats in [1, 2)
        scale = np.exp2(rng.integers(-30, 30, n // 2).astype(np.float64))
        a[: n // 2] = (m * scale * 2.0 * np.float64(h32)).astype(np.float32)
        a[n // 2: n // 2 + n // 4] *= np.float32(1e-20)                                            # small, still normal
        with np.errstate(all="ignore"):
            d = (a - b).astype(np.float32)
            want = (d.astype(np.float64) / 2.0 / np.float64(h32)).astype(np.float32)
        out = np.empty(n, np.float32)
        emu.emu_cdiff_f32(1, a.ctypes.data, b.ctypes.data, h32, n, out.ctypes.data)
        ok = (out.view(np.uint32) == want.view(np.uint32)) | (np.isnan(out) & np.isnan(want))
        if not ok.all():
            bad.append((float(h32), int(np.count_nonzero(~ok)), a[~ok][:2].tolist()))
    assert not bad, bad[:5]


def test_second_fallback_is_rare_and_ties_take_it(emu):
    """Flagged elements are first re-evaluated with checked products (DIV_FAST_CHECKED); only a quotient on or
    next to a rounding boundary goes on to the divisions. Random subnormal quotients never do; exact ties
    in the subnormal range (h = 3: k * 2^-149 / 6) always do."""
    rng = np.random.default_rng(12)
    n = 200_000
    h32 = np.float32(1.0 / 999)
    a = (rng.uniform(1, 2, n) * np.exp2(rng.integers(-149, -100, n).astype(np.float64))).astype(np.float32)
    b = np.zeros(n, np.float32)
    out = np.empty(n, np.float32)
    emu.emu_second_fallbacks.restype = C.c_longlong
    emu.emu_second_fallbacks()
    emu.emu_cdiff_f32(1, a.ctypes.data, b.ctypes.data, h32, n, out.ctypes.data)
    assert emu.emu_second_fallbacks() <= 2
    want = (a.astype(np.float64) / 2.0 / np.float64(h32)).astype(np.float32)
    assert np.array_equal(out.view(np.uint32), want.view(np.uint32))
    k = np.arange(3, 6000, 6, dtype=np.float64)                      # k / 6 = integer + 1/2
    a = (k * 2.0 ** -149).astype(np.float32)
    b = np.zeros(a.size, np.float32)
    out = np.empty(a.size, np.float32)
    emu.emu_cdiff_f32(1, a.ctypes.data, b.ctypes.data, np.float32(3.0), a.size, out.ctypes.data)
    assert emu.emu_second_fallbacks() == a.size
    want = (a.astype(np.float64) / 2.0 / 3.0).astype(np.float32)
    assert np.array_equal(out.view(np.uint32), want.view(np.uint32))


@pytest.mark.parametrize("hx,hy", [(1.0, 1.0), (1.0 / 99, 1.0 / 99), (1.0 / 16383, 1.0 / 16383), (0.1, 3.0),
                                   (1.0 / 128, 1.0 / 40), (7.0 / 3, 1e-3), (1.0 / 128, 1.0 / 256), (4.0, 0.25),
                                   (2.0 ** -30, 2.0 ** -30), (2.0 ** 20, 2.0 ** 20)])
def test_laplacian_tail_matches_reference_formula(emu, hx, hy):
    n = 400_000
    c = _inputs(n, 3)
    s = (4.0 * c.astype(np.float64) * (1 + np.random.default_rng(4).normal(0, 1e-3, n))).astype(np.float32)
    s[: n // 3] = _inputs(n // 3, 5)
    hx32, hy32 = np.float32(hx), np.float32(hy)
    with np.errstate(all="ignore"):
        want = ((s.astype(np.float64) - 4.0 * c.astype(np.float64)) / np.float64(hx32) / np.float64(hy32)).astype(np.float32)
    for mode in range(emu.emu_metric_mode_f32(hx32, hy32) + 1):
        out = np.empty(n, np.float32)
        emu.emu_lap_tail_f32(mode, s.ctypes.data, c.ctypes.data, hx32, hy32, n, out.ctypes.data)
        ok = (out.view(np.uint32) == want.view(np.uint32)) | (np.isnan(out) & np.isnan(want))
        assert ok.all(), f"mode {mode}: {np.count_nonzero(~ok)} mismatches, e.g. s={s[~ok][:3]} c={c[~ok][:3]}"


def test_laplacian_tail_ambiguous_cases_take_the_exact_path(emu):
    """Construct quotients that land (nearly) on a float rounding boundary: d = m * hx * hy with m
    a float midpoint, so only the exact two-division path rounds like the reference."""
    rng = np.random.default_rng(7)
    hx32, hy32 = np.float32(1.0 / 99), np.float32(1.0 / 57)
    base = rng.uniform(1.0, 2.0, 20000).astype(np.float32)
    mid = base.astype(np.float64) + 2.0 ** -24            # midpoints between adjacent floats in [1,2)
    d = mid * np.float64(hx32) * np.float64(hy32)         # quotient ~ mid (up to double rounding)
    c = np.zeros_like(base)
    s = d.astype(np.float32)                               # d must come from floats: use c = 0, s = RN32(d)
    out = np.empty(s.size, np.float32)
    emu.emu_lap_tail_f32(1, s.ctypes.data, c.ctypes.data, hx32, hy32, s.size, out.ctypes.data)
    want = (s.astype(np.float64) / np.float64(hx32) / np.float64(hy32)).astype(np.float32)
    assert np.array_equal(out.view(np.uint32), want.view(np.uint32))


@pytest.mark.parametrize("hx,hy", [(1.0, 1.0), (1.0 / 128, 1.0 / 256), (4.0, 0.25), (2.0 ** -30, 2.0 ** -30), (2.0 ** 20, 2.0 ** 20)])
def test_fp64_power_of_two_mode_is_the_division(emu, hx, hy):
    """fp64 with power-of-two spacings: every division is a multiplication by an exact reciprocal, i.e. the
    same real number rounded once -- for every input, subnormal results included, so there is nothing to flag."""
    emu.emu_cdiff_f64.argtypes = [C.c_int, C.c_void_p, C.c_void_p, C.c_double, C.c_longlong, C.c_void_p]
    emu.emu_lap_tail_f64.argtypes = [C.c_int, C.c_void_p, C.c_void_p, C.c_double, C.c_double, C.c_longlong, C.c_void_p]
    emu.emu_metric_mode_f64.argtypes = [C.c_double, C.c_double]
    assert emu.emu_metric_mode_f64(hx, hy) == 2 and emu.emu_metric_mode_f64(0.1, hy) == 1
    n = 200_000
    rng = np.random.default_rng(11)
    def inputs(seed):
        r = np.random.default_rng(seed)
        x = r.uniform(1.0, 2.0, n) * np.exp2(r.integers(-1074, 200, n).astype(np.float64)) * r.choice([-1.0, 1.0], n)
        x[r.integers(0, n, n // 50)] = 0.0
        x[r.integers(0, n, n // 50)] = -0.0
        x[r.integers(0, n, n // 20)] = r.integers(1, 1 << 52, n // 20).astype(np.uint64).view(np.float64)  # subnormals
        return x
    a, b, c = inputs(1), inputs(2), inputs(3)
    s = 4.0 * c * (1 + rng.normal(0, 1e-9, n))
    s[: n // 3] = inputs(5)[: n // 3]
    with np.errstate(all="ignore"):
        want_c = (a - b) / 2.0 / hx
        want_l = (s - 4.0 * c) / hx / hy     # numpy rounds s - 4c twice (4c is exact, so once): same as the fma
    for want, call in ((want_c, lambda m, o: emu.emu_cdiff_f64(m, a.ctypes.data, b.ctypes.data, hx, n, o.ctypes.data)),
                       (want_l, lambda m, o: emu.emu_lap_tail_f64(m, s.ctypes.data, c.ctypes.data, hx, hy, n, o.ctypes.data))):
        for mode in (0, 2):
            out = np.empty(n, np.float64)
            call(mode, out)
            ok = (out.view(np.uint64) == want.view(np.uint64)) | (np.isnan(out) & np.isnan(want))
            assert ok.all(), f"mode {mode}: {np.count_nonzero(~ok)} mismatches"


def test_fp64_quotients_by_fma_corrections_are_the_divisions(emu):
    """fp64, spacings that are not powers of two (mode 1): every division by a constant is a multiplication by the
    rounded reciprocal plus two remainder corrections by FMA. Against NumPy's IEEE divisions over 300 spacings:
    the whole double range (what underflows or overflows on the way is flagged and redone), and numerators built
    so that the quotient sits on or next to the midpoint of two doubles (via x87 extended precision), the inputs
    a wrong last-bit correction gets wrong."""
    emu.emu_cdiff_f64.argtypes = [C.c_int, C.c_void_p, C.c_void_p, C.c_double, C.c_longlong, C.c_void_p]
    emu.emu_lap_tail_f64.argtypes = [C.c_int, C.c_void_p, C.c_void_p, C.c_double, C.c_double, C.c_longlong, C.c_void_p]
    emu.emu_metric_mode_f64.argtypes = [C.c_double, C.c_double]
    emu.emu_second_fallbacks.restype = C.c_longlong
    assert np.finfo(np.longdouble).nmant >= 63, "needs x87 extended precision to build double midpoints"
    rng = np.random.default_rng(21)
    n = 40_000
    hs = list(np.exp2(rng.uniform(-20, 6, 280)))
    hs += [1.0 / 999, 1.0 / 40, 1.0 / 16383, 0.1, 0.3, 3.0, 1.0 / 3, np.nextafter(1.0, 0.0), np.nextafter(0.25, 1.0), np.nextafter(np.nextafter(2.0, 0.0), 0.0)]
    assert emu.emu_metric_mode_f64(np.nextafter(1.0, 0.0), 0.1) == 0      # significand all ones: IEEE divisions
    bad = []
    for k, hx in enumerate(hs):
        hx = float(hx)
        hy = float(hs[(k * 7 + 3) % len(hs)])
        if emu.emu_metric_mode_f64(hx, hy) != 1:
            continue
        r = np.random.default_rng(1000 + k)
        a = r.uniform(1.0, 2.0, n) * np.exp2(r.integers(-1074, 1000, n).astype(np.float64)) * r.choice([-1.0, 1.0], n)
        a[r.integers(0, n, n // 50)] = 0.0
        a[r.integers(0, n, n // 50)] = -0.0
        b = np.zeros(n)
        # quotients on the midpoints of doubles in [1, 2) times a power of two: d = RN64(m * 2hx), and for the
        # Laplacian tail D = RN64(RN80(m * hx) * hy), computed in extended precision
        m = r.uniform(1.0, 2.0, n // 2).astype(np.longdouble) + np.longdouble(2.0) ** -53
        sc = np.exp2(r.integers(-200, 200, n // 2).astype(np.float64)).astype(np.longdouble)
        a[: n // 2] = (m * sc * np.longdouble(2.0 * hx)).astype(np.float64)
        with np.errstate(all="ignore"):
            want = (a - b) / 2.0 / hx
        out = np.empty(n)
        emu.emu_second_fallbacks()
        emu.emu_cdiff_f64(1, a.ctypes.data, b.ctypes.data, hx, n, out.ctypes.data)
        redone = emu.emu_second_fallbacks()
        ok = (out.view(np.uint64) == want.view(np.uint64)) | (np.isnan(out) & np.isnan(want))
        if not ok.all():
            bad.append(("cdiff", hx, int(np.count_nonzero(~ok)), a[~ok][:2].tolist()))
        assert redone < n // 3, "the corrected quotient is meant to cover the working range"
        c = np.zeros(n)
        s_ = a.copy()
        s_[: n // 2] = ((m * sc * np.longdouble(hx)) * np.longdouble(hy)).astype(np.float64)
        with np.errstate(all="ignore"):
            want = (s_ - 4.0 * c) / hx / hy
        emu.emu_lap_tail_f64(1, s_.ctypes.data, c.ctypes.data, hx, hy, n, out.ctypes.data)
        ok = (out.view(np.uint64) == want.view(np.uint64)) | (np.isnan(out) & np.isnan(want))
        if not ok.all():
            bad.append(("lap", hx, hy, int(np.count_nonzero(~ok)), s_[~ok][:2].tolist()))
    assert not bad, bad[:5]
