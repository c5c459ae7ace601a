"""Parity of the product's C ABI against the reference, on two back ends:

* ``cuda`` (marked gpu): the real sm_100a library, fluid_simulation_b200/libfluidb200.so;
* ``emu``  (CPU): tests/emu's host build of the SAME host logic (csrc/api.cpp) and the SAME
  warp task bodies (csrc/warp_kernels.cuh), 32 threads per warp -- so the kernel and planner
  logic is checked here, where there is no GPU.

Expected values: the golden fields recorded from the reference (tests/golden/*.npz), and the
oracle (pinned to the reference in tests/test_oracle.py) on the same inputs. Comparison is
bit-exact on all seven fields (stricter than fluid_cpp/tests/test_main.cpp:77, which checks v
only, with ==).
"""
import os

import numpy as np
import pytest

from oracle import Oracle
from tests import helpers as H

BACKENDS = [pytest.param("emu"), pytest.param("cuda", marks=pytest.mark.gpu)]


@pytest.fixture(scope="module", params=BACKENDS)
def Solver(request):
    if request.param == "emu":
        return H.solver_factory(H.build_emu())
    import torch
    assert torch.cuda.is_available(), "gpu-marked test needs a CUDA device"
    return H.solver_factory(None)


@pytest.fixture
def env():
    saved = dict(os.environ)
    yield os.environ
    os.environ.clear()
    os.environ.update(saved)


GOLDEN_FIELD_CASES = [
    ("cavity41_S3_K7", None, 1), ("cavity41_S3_K7", None, 2), ("cavity41_S3_K7", None, 3),
    ("cavity41_S3_K7", None, 4), ("cavity41_S3_K7", None, 5), ("cavity41_S3_K7", None, 7),
    ("cavity41_S3_K7", None, 8), ("cavity41_S1_K1", None, 4), ("cavity41_S10_K50", None, 4),
    ("cavity41_S10_K50", None, 8), ("cavity67x45_S25_K30", 6, 4), ("cavity45x131_S12_K33", 5, 4),
    ("cavity45x131_S12_K33", 3, 8), ("cavity41_f64_S100_K50", 6, 4), ("cavity41_f64_S100_K50", 3, 7),
    ("general53x37_S20_K30", 5, 4), ("general53x37_f64_S20_K30", 4, 4),
]


@pytest.mark.parametrize("name,steps,kmax", GOLDEN_FIELD_CASES)
def test_golden_fields(Solver, golden, name, steps, kmax):
    """Full solve through migrateSolve(); S may be cut short for the emulator, in which case the
    oracle (pinned to the same golden files) provides the expected fields."""
    rec = dict(golden[name])
    full = steps is None or steps >= rec["S"]
    if not full:
        rec["S"] = steps
    s = H.golden_case_setup(rec, Solver)
    s.set_sweeps_per_launch(kmax)
    s.migrateSolve()
    want = H.load_fields(golden[name]) if full else H.golden_case_setup(rec, Oracle).solve()
    H.compare_with(s, want, tag=f"{name}/k{kmax}")
    s.close()


@pytest.mark.parametrize("dtype", ["float32", "float64"])
@pytest.mark.parametrize("kmax", [1, 4, 5, 8])
def test_per_cell_edge_bcs_inside_temporal_blocking(Solver, dtype, kmax):
    X, Y = 61, 400 if dtype == "float32" else 200   # interior strips (fast path) between the edge strips
    src = H.edges_only_bcs(X, Y, np.dtype(dtype))
    src.time_step, src.num_iterations, src.num_poisson_iterations = 0.0002, 3, 13
    s = H.copy_bcs(src, Solver(X, Y, 1.0, 1.0, np.dtype(dtype)))
    s.set_sweeps_per_launch(kmax)
    s.migrateSolve()
    src.solve()
    H.compare_with(s, src, tag=f"edges/{dtype}/k{kmax}")
    s.close()


@pytest.mark.parametrize("rpc", [2, 3, 5, 16])
def test_chunking_does_not_change_bits(Solver, env, golden, rpc):
    env["FLUIDB200_RPC"] = str(rpc)
    rec = dict(golden["cavity67x45_S25_K30"]); rec["S"] = 2; rec["K"] = 9
    s = H.golden_case_setup(rec, Solver)
    s.set_sweeps_per_launch(4)
    s.migrateSolve()
    H.compare_with(s, H.golden_case_setup(rec, Oracle).solve(), tag=f"rpc{rpc}")
    s.close()


def test_phases_one_at_a_time(Solver, golden):
    rec = dict(golden["cavity45x131_S12_K33"]); rec["S"] = 1
    s = H.golden_case_setup(rec, Solver)
    o = H.golden_case_setup(rec, Oracle)
    s.migrateHostToDevice()
    for step in range(2):
        s.unifiedApproximateTimeStepAndRightHand(); o.predictor()
        for n in ("u_next", "v_next", "rhs"):
            H.assert_bit_equal(s.device_field(n), getattr(o, n), f"step{step}:{n}")
        s.poissonSweeps(11); o.poisson_sweeps(11)
        for n in ("p", "p_next"):
            H.assert_bit_equal(s.device_field(n), getattr(o, n), f"step{step}:{n}")
        s.unifiedVelocityCorrection(); o.corrector()
        for n in ("u", "v"):
            H.assert_bit_equal(s.device_field(n), getattr(o, n), f"step{step}:{n}")
    s.close()


def test_solve_continues_from_device_state(Solver, golden):
    """Repeated solve() continues the simulation (SURVEY 8b 'Semantics to preserve')."""
    rec = dict(golden["cavity41_S10_K50"]); rec["K"] = 6
    a = H.golden_case_setup(rec, Solver); a.num_iterations = 4
    a.migrateHostToDevice(); a.solve(); a.solve(); a.migrateDeviceToHost()
    o = H.golden_case_setup(rec, Oracle); o.num_iterations = 8; o.solve()
    H.compare_with(a, o, tag="continue")
    a.close()


def test_setters_touch_host_only_until_migrate(Solver):
    s = Solver(41, 41)
    s.num_iterations, s.num_poisson_iterations = 1, 2
    s.migrateHostToDevice()          # pristine grid, like the reference ctor upload
    s.set_cavity(1.0)                # host only
    s.solve(); s.synchronize()
    assert not s.device_field("u").any(), "device must not see BCs before migrateHostToDevice"
    s.close()


def test_tiny_and_degenerate_boxes(Solver):
    from fluid_simulation_b200 import FluidB200Error
    for X, Y in ((3, 3), (3, 9), (9, 3), (4, 5)):
        src = Oracle(X, Y).set_cavity()
        src.num_iterations, src.num_poisson_iterations = 3, 5
        s = H.copy_bcs(src, Solver(X, Y)); s.set_sweeps_per_launch(4)
        s.migrateSolve(); src.solve()
        H.compare_with(s, src, tag=f"{X}x{Y}")
        s.close()
    with pytest.raises(FluidB200Error):
        Solver(2, 41)


def test_zero_iterations_are_noops(Solver):
    s = Solver(41, 41).set_cavity()
    s.num_iterations, s.num_poisson_iterations = 2, 0
    s.migrateSolve()
    o = Oracle(41, 41).set_cavity(); o.num_iterations, o.num_poisson_iterations = 2, 0; o.solve()
    H.compare_with(s, o, tag="K0")
    s.num_iterations = 0
    s.solve(); s.synchronize()
    s.close()


@pytest.mark.parametrize("X,Y,K,kmax,dtype,rpc", [
    (40, 300, 9, 4, "float32", 0), (40, 300, 17, 8, "float32", 0), (40, 300, 7, 3, "float32", 0),
    (50, 300, 5, 4, "float32", 7), (40, 170, 5, 2, "float64", 0), (23, 300, 3, 1, "float32", 0),
])
def test_interior_strip_fast_paths(Solver, env, X, Y, K, kmax, dtype, rpc):
    """Grids wide enough (>= 3 strips along y) that the register-pipelined fast paths of all three
    kernels run (interior strips, rows away from the walls), next to the general path."""
    if rpc:
        env["FLUIDB200_RPC"] = str(rpc)
    o = Oracle(X, Y, dtype=np.dtype(dtype)).set_cavity()
    o.time_step, o.num_iterations, o.num_poisson_iterations = 1e-5, 2, K
    s = Solver(X, Y, dtype=np.dtype(dtype)).set_cavity()
    s.time_step, s.num_iterations, s.num_poisson_iterations = 1e-5, 2, K
    s.set_sweeps_per_launch(kmax)
    s.migrateSolve()
    o.solve()
    H.compare_with(s, o, tag=f"fast/{X}x{Y}/k{kmax}")
    s.close()


@pytest.mark.parametrize("X,Y", [(40, 300), (33, 257)])
def test_subnormal_front_takes_exact_fallback(Solver, X, Y):
    """A tiny diffusion number makes the values at the advancing front underflow through the
    subnormal range (the reference runs on x86 without flush-to-zero). The fp32 kernels' division-
    free arithmetic flags those inputs and evaluates the affected rows a second time, inside
    a per-lane branch: this must neither change bits nor deadlock on a warp shuffle. 40 x 300 points on
    the unit box: spacings 1/39, 1/299 (checked products, then divisions for ties); 33 x 257: powers
    of two (the Laplacian tail in double)."""
    S, K = 9, 2
    o = Oracle(X, Y).set_cavity()
    o.time_step, o.num_iterations, o.num_poisson_iterations = 1e-9, S, K
    s = Solver(X, Y).set_cavity()
    s.time_step, s.num_iterations, s.num_poisson_iterations = 1e-9, S, K
    s.migrateSolve()
    o.solve()
    tiny = np.finfo(np.float32).tiny
    assert np.any((np.abs(o.u_next) < tiny) & (o.u_next != 0)), "case no longer produces subnormals"
    H.compare_with(s, o, tag=f"subnormal-front/{X}x{Y}")
    s.close()


@pytest.mark.parametrize("X,Y,divmode", [(33, 257, None), (33, 257, 1), (33, 257, 0), (40, 300, 0), (40, 300, 1)])
def test_division_modes_agree_with_the_reference(Solver, env, X, Y, divmode):
    """fp32 arithmetic modes of the predictor/corrector (Metric::mode): 2 = power-of-two spacing
    (33 x 257 points on the unit box), 1 = division-free with flagged exact fallback, 0 = IEEE
    divisions. FLUIDB200_DIVMODE can only force a more conservative mode."""
    if divmode is not None:
        env["FLUIDB200_DIVMODE"] = str(divmode)
    o = Oracle(X, Y).set_cavity()
    o.time_step, o.num_iterations, o.num_poisson_iterations = 2e-5, 3, 5
    s = Solver(X, Y).set_cavity()
    s.time_step, s.num_iterations, s.num_poisson_iterations = 2e-5, 3, 5
    s.migrateSolve()
    o.solve()
    H.compare_with(s, o, tag=f"divmode{divmode}/{X}x{Y}")
    s.close()


@pytest.mark.parametrize("dtype", ["float32", "float64"])
@pytest.mark.parametrize("kmax", [2, 5])
def test_unit_spacing(Solver, kmax, dtype):
    """domain = (X-1, Y-1) => hx = hy = 1: the benchmark's spacing (power-of-two arithmetic mode in fp32
    and fp64, and the Jacobi kernel skips the exact multiplications of rhs by hx and hy)."""
    X, Y = 48, 300
    o = Oracle(X, Y, float(X - 1), float(Y - 1), np.dtype(dtype)).set_cavity()
    o.time_step, o.num_iterations, o.num_poisson_iterations = 0.001, 3, 7
    s = Solver(X, Y, float(X - 1), float(Y - 1), np.dtype(dtype)).set_cavity()
    s.time_step, s.num_iterations, s.num_poisson_iterations = 0.001, 3, 7
    s.set_sweeps_per_launch(kmax)
    s.migrateSolve()
    o.solve()
    H.compare_with(s, o, tag=f"unit/k{kmax}/{dtype}")
    s.close()


def test_raw_dat_dumps_follow_the_reference_layout(Solver, golden, tmp_path):
    """`Cuda{U,V,P}Values.float.dat` as written by fluid_cpp/cuda_main.cu:69-80: raw little-endian
    fp32 in getter order, re-indexed by the reference's notebook as data[j*N + i] -> [i, j]."""
    rec = dict(golden["cavity41_S3_K7"])
    s = H.golden_case_setup(rec, Solver)
    s.migrateSolve()
    paths = s.write_dat("Cuda", str(tmp_path))
    want = H.load_fields(golden["cavity41_S3_K7"])
    for path, n in zip(paths, ("u", "v", "p")):
        raw = np.fromfile(path, dtype="<f4")
        assert raw.size == 41 * 41
        H.assert_bit_equal(raw.reshape(41, 41), want[n], n)
        H.assert_bit_equal(type(s).read_dat(path, 41, 41), want[n], n)
    s.close()


def _random_cases(n, seed, max_dim):
    rng = np.random.default_rng(seed)
    out = []
    for _ in range(n):
        X = int(rng.integers(3, max_dim)); Y = int(rng.integers(3, max_dim))
        out.append((X, Y, int(rng.integers(1, 14)), int(rng.integers(1, 9)), ["float32", "float64"][int(rng.integers(0, 2))]))
    return out


@pytest.mark.parametrize("X,Y,K,kmax,dtype", _random_cases(10, 2024, 260))
def test_random_shapes(Solver, X, Y, K, kmax, dtype):
    """Odd, tiny and non-multiple-of-tile shapes (3 <= X, Y < 260), random depth and blocking."""
    o = Oracle(X, Y, dtype=np.dtype(dtype)).set_cavity()
    o.time_step, o.num_iterations, o.num_poisson_iterations = 1e-5, 2, K
    s = Solver(X, Y, dtype=np.dtype(dtype)).set_cavity()
    s.time_step, s.num_iterations, s.num_poisson_iterations = 1e-5, 2, K
    s.set_sweeps_per_launch(kmax)
    s.migrateSolve()
    o.solve()
    H.compare_with(s, o, tag=f"random/{X}x{Y}/K{K}/k{kmax}/{dtype}")
    s.close()


@pytest.mark.parametrize("knob", ["FLUIDB200_NO_SPLIT", "FLUIDB200_SPLIT_MIN_TASKS"])
def test_single_kernel_and_split_launch_agree(Solver, env, knob):
    """The Jacobi launch as ONE general kernel (NO_SPLIT=1, or a grid below the split threshold) and
    as the split pair must produce the same bits as the oracle."""
    env[knob] = "1" if knob == "FLUIDB200_NO_SPLIT" else "1000000"
    X, Y = 90, 420
    o = Oracle(X, Y).set_cavity()
    o.time_step, o.num_iterations, o.num_poisson_iterations = 1e-5, 2, 13
    s = Solver(X, Y).set_cavity()
    s.time_step, s.num_iterations, s.num_poisson_iterations = 1e-5, 2, 13
    s.migrateSolve()
    o.solve()
    H.compare_with(s, o, tag=knob)
    s.close()


@pytest.mark.parametrize("kmax", [1, 2, 3, 4, 5, 6, 7, 8])
@pytest.mark.parametrize("unit", [False, True])
def test_split_launch_every_depth(Solver, kmax, unit):
    """All three kernels of the split Jacobi launch at every blocking depth: the lean FAST kernel on
    the middle strip, the lean EDGE kernel on the strips holding column 0 / Y-1 (cavity edges are
    CONST or UNSET), the general kernel on the short chunks at the x walls."""
    X, Y = 72, 300
    dom = (float(X - 1), float(Y - 1)) if unit else (1.0, 1.0)
    o = Oracle(X, Y, *dom).set_cavity()
    o.time_step, o.num_iterations, o.num_poisson_iterations = (0.001 if unit else 1e-5), 2, 2 * kmax
    s = Solver(X, Y, *dom).set_cavity()
    s.time_step, s.num_iterations, s.num_poisson_iterations = o.time_step, 2, 2 * kmax
    s.set_sweeps_per_launch(kmax)
    s.migrateSolve()
    o.solve()
    H.compare_with(s, o, tag=f"split-depth/k{kmax}/unit{unit}")
    s.close()


@pytest.mark.parametrize("depth", [1, 2, 3, 4, 6, 8])
def test_interior_flags_inside_temporal_blocking(Solver, env, depth):
    """General boundary conditions (interior obstacle flags for u, v and p, partly unset edges) with
    the pressure flags applied at every level of a `depth`-sweep launch (SerialNavierStokes.cpp:
    154-160 runs once per sweep). The solver caps masked launches at depth 2 unless told otherwise."""
    from tests.golden.make_golden import general_bcs
    env["FLUIDB200_MASK_KMAX"] = str(depth)
    X, Y, S, K = 61, 300, 3, 2 * depth
    o = general_bcs(X, Y, np.dtype("float32"))
    o.time_step, o.num_iterations, o.num_poisson_iterations = 1e-4, S, K
    s = Solver(X, Y, 1.0, 1.0, np.dtype("float32"))
    for f, setter in (("u", s.setUBoundaryCondition), ("v", s.setVBoundaryCondition), ("p", s.setPBoundaryCondition)):
        flag, val = getattr(o, f + "_set"), getattr(o, f + "_bc")
        for x, y in np.argwhere(flag):
            setter(int(x), int(y), float(val[x, y]))
    s.time_step, s.num_iterations, s.num_poisson_iterations = 1e-4, S, K
    s.set_sweeps_per_launch(8)
    s.migrateSolve()
    o.solve()
    H.compare_with(s, o, tag=f"masked/depth{depth}")
    s.close()


@pytest.mark.parametrize("tiles", [True, False])
@pytest.mark.parametrize("depth", [2, 5, 8])
def test_cavity_with_obstacles_keeps_lean_kernels(Solver, env, depth, tiles):
    """A lid-driven cavity (constant / copy edges) with two small obstacles: the masked Jacobi launch
    is split, tasks whose footprint holds no flagged cell run the lean FAST / EDGE kernels, the others
    (one in the middle strip, one in the strip holding column 0) the general kernel."""
    env["FLUIDB200_MASK_KMAX"] = str(depth)
    if not tiles:   # without the tile summary every task of the masked launch takes the general kernel
        env["FLUIDB200_NO_MASK_TILES"] = "1"
    X, Y, S, K = 96, 300, 2, 2 * depth
    def setup(s):
        s.set_cavity(1.0)
        for x in range(40, 44):
            for y in range(150, 155):
                s.setUBoundaryCondition(x, y, 0.0); s.setVBoundaryCondition(x, y, 0.0)
                if (x + y) % 2 == 0:
                    s.setPBoundaryCondition(x, y, 0.0625)
        s.setPBoundaryCondition(70, 5, -0.5)
        s.time_step, s.num_iterations, s.num_poisson_iterations = 1e-5, S, K
        return s
    o = setup(Oracle(X, Y))
    s = setup(Solver(X, Y))
    s.set_sweeps_per_launch(8)
    s.migrateSolve()
    o.solve()
    H.compare_with(s, o, tag=f"obstacles/depth{depth}/tiles{tiles}")
    s.close()


@pytest.mark.parametrize("X,Y", [(33, 257), (65, 129)])
def test_fp64_power_of_two_spacing(Solver, X, Y):
    """2^k + 1 points on the unit box in fp64: divisions become multiplications by exact reciprocals."""
    o = Oracle(X, Y, 1.0, 1.0, np.dtype("float64")).set_cavity()
    o.time_step, o.num_iterations, o.num_poisson_iterations = 2e-5, 3, 9
    s = Solver(X, Y, 1.0, 1.0, np.dtype("float64")).set_cavity()
    s.time_step, s.num_iterations, s.num_poisson_iterations = 2e-5, 3, 9
    s.migrateSolve()
    o.solve()
    H.compare_with(s, o, tag=f"f64-pow2/{X}x{Y}")
    s.close()


@pytest.mark.parametrize("X,Y,K,chunk,dtype", [(256, 100, 5, None, "float32"), (300, 77, 3, 20, "float32"), (200, 40, 2, 16, "float64"),
                                               (333, 150, 9, 44, "float32"), (400, 64, 3, 100, "float32"), (530, 40, 4, 128, "float64")])
def test_migrate_solve_pipelined_over_row_chunks(Solver, env, X, Y, K, chunk, dtype):
    """fluidb200_migrate_solve with one timestep: upload, compute and download run chunk by chunk, every chunk
    on its rows plus K+2 recomputed halo rows (the last two cases have a last chunk long enough to be cut in three).
    Host results and the state left on the device equal the plain upload + solve + download, and the oracle."""
    if chunk:
        env["FLUIDB200_E2E_CHUNK"] = str(chunk)
    rng = np.random.default_rng(X + Y)
    npdt = np.dtype(dtype)
    st = {n: (0.01 * rng.standard_normal((X, Y))).astype(npdt) for n in ("u", "v", "p")}
    def make(cls):
        s = cls(X, Y, 1.0, 1.0, npdt).set_cavity(1.0)
        s.time_step, s.num_iterations, s.num_poisson_iterations = 1e-5, 1, K
        return s
    a = make(Solver)
    a.set_sweeps_per_launch(4)
    a.migrateHostToDevice()                       # boundary description (and zero fields)
    k0 = a.launch_count()
    got = a.migrate_solve_uvp(st["u"], st["v"], st["p"])
    assert a.launch_count() - k0 > 2 + -(-K // 4), "the chunk pipeline was not taken"
    o = make(Oracle)
    for n in ("u", "v", "p"):
        getattr(o, n)[:] = st[n]
    o.set_cavity(1.0)                             # (the setters re-seed the edges of u, v, p like the upload of BCs did not)
    # what the device really held: the uploaded u, v, p as given (edges included)
    for n in ("u", "v", "p"):
        getattr(o, n)[:] = st[n]
    o.p_next[:] = o.p
    o.solve()
    for n, g in zip(("u", "v", "p"), got):
        H.assert_bit_equal(g, getattr(o, n), f"migrate_solve/{n}")
    a.migrateDeviceToHost()
    H.compare_with(a, o, fields=("u", "v", "p", "u_next", "v_next", "rhs", "p_next"), tag="migrate_solve/device")
    # a second call continues from host arrays again; and more than one timestep takes the plain sequence
    a.num_iterations = 2
    k0 = a.launch_count()
    got2 = a.migrate_solve_uvp(*got)
    o.num_iterations = 2
    o.solve()
    for n, g in zip(("u", "v", "p"), got2):
        H.assert_bit_equal(g, getattr(o, n), f"migrate_solve/second/{n}")
    a.close()


@pytest.mark.parametrize("X,Y,scale", [(37, 150, 1e-262), (64, 131, 1e+250)])
def test_fp64_numerators_out_of_range_take_the_divisions(Solver, X, Y, scale):
    """fp64 with a spacing that is not a power of two divides by FMA corrections, valid for numerators in [2^-800, 2^800):
    fields scaled far below (or above) that range, next to boundary values of order one, send the flagged rows through the
    IEEE divisions while their neighbours stay on the fast path. (1e+250: products overflow to inf / nan in the reference
    too; the comparison treats equal non-finite bit patterns and NaNs alike.)"""
    rng = np.random.default_rng(X * Y)
    st = {n: (scale * rng.standard_normal((X, Y))) for n in ("u", "v", "p")}
    pair = []
    for cls in (Solver, Oracle):
        q = cls(X, Y, 1.0, 1.0, np.dtype("float64")).set_cavity(1.0)
        for n in ("u", "v", "p"):
            (q._hostf(n) if cls is Solver else getattr(q, n))[1:-1, 1:-1] = st[n][1:-1, 1:-1]
        q.time_step, q.num_iterations, q.num_poisson_iterations = 1e-5, 2, 3
        pair.append(q)
    s, o = pair
    if hasattr(o, "p_next"):
        o.p_next[:] = o.p
    with np.errstate(all="ignore"):
        s.migrateSolve()
        o.solve()
    H.compare_with(s, o, tag=f"f64-out-of-range/{X}x{Y}/{scale:g}")
    s.close()

