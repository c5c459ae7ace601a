"""GPU-only parity at sizes the emulator cannot reach: the golden digests recorded from the
reference for the larger known-answer grids (incl. the 300^2 case with true subnormals and the
100^2 case that discriminates the fp64 Laplacian tail), oracle equality up to 2048^2, and -- at
BASELINE.json's full 16384^2 -- size-independent properties: the temporally blocked kernel is
bit-identical for every blocking depth and chunking (Jacobi is order-free), and the solve is
deterministic."""
import numpy as np
import pytest

from oracle import Oracle, fnv1a64
from tests import helpers as H

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def Solver():
    import torch
    assert torch.cuda.is_available()
    return H.solver_factory(None)


@pytest.mark.parametrize("name", ["cavity129_S100_K50", "cavity257_S20_K100", "cavity100_S50_K50",
                                  "cavity300_S10_K100", "cavity100_f64_S20_K50", "cavity41_S1000_K50"])
@pytest.mark.parametrize("kmax", [1, 4, 8])
def test_golden_digests(Solver, golden, name, kmax):
    rec = golden[name]
    s = H.golden_case_setup(rec, Solver)
    s.set_sweeps_per_launch(kmax)
    s.migrateSolve()
    for n in H.ALL_FIELDS:
        assert fnv1a64(s._hostf(n)) == rec["digest"][n], f"{name}/k{kmax}:{n}"
    for n, pts in rec["spots"].items():
        for key, hx in pts.items():
            x, y = map(int, key.split(","))
            assert float(s._hostf(n)[x, y]).hex() == hx
    s.close()


def test_reference_test_main_convention(Solver, golden):
    """fluid_cpp/tests/test_main.cpp: 41^2, 1000 steps, K=50, compare V with == at x*ny+y."""
    rec = golden["cavity41_S1000_K50"]
    s = H.golden_case_setup(rec, Solver)
    s.migrateSolve()
    tested = np.empty(41 * 41, np.float32)
    assert s.getVValues(tested) == 0
    unit = H.load_fields(rec)["v"].reshape(-1)
    errors = sum(1 for x in range(41) for y in range(41) if not (unit[x * 41 + y] == tested[x * 41 + y]))
    assert errors == 0
    s.close()


@pytest.mark.parametrize("X,Y,S,K,kmax", [(1000, 1000, 3, 50, 4), (1237, 1011, 2, 21, 8), (2048, 2048, 1, 9, 4),
                                          (509, 3001, 2, 10, 3)])
def test_oracle_equality_large(Solver, X, Y, S, K, kmax):
    dt = float(np.float32(0.125) * np.float32(1.0 / (max(X, Y) - 1)) ** 2 / np.float32(0.1))  # r = 0.125, stable
    o = Oracle(X, Y).set_cavity()
    o.time_step, o.num_iterations, o.num_poisson_iterations = dt, S, K
    s = Solver(X, Y).set_cavity()
    s.time_step, s.num_iterations, s.num_poisson_iterations = dt, S, K
    s.set_sweeps_per_launch(kmax)
    s.migrateSolve()
    o.solve()
    H.compare_with(s, o, tag=f"{X}x{Y}")
    s.close()


@pytest.mark.parametrize("dtype", ["float64"])
def test_oracle_equality_double(Solver, dtype):
    X, Y, S, K = 700, 333, 3, 17
    o = Oracle(X, Y, dtype=dtype).set_cavity()
    o.time_step, o.num_iterations, o.num_poisson_iterations = 1e-6, S, K
    s = Solver(X, Y, dtype=np.dtype(dtype)).set_cavity()
    s.time_step, s.num_iterations, s.num_poisson_iterations = 1e-6, S, K
    s.migrateSolve(); o.solve()
    H.compare_with(s, o, tag="f64")
    s.close()


def test_full_size_blocking_depth_independence(Solver):
    """16384^2 (BASELINE.json configs[2]): p, u, v after 2 steps x 8 sweeps are bit-identical for
    blocking depths 1, 4, 7 and 8 (a 100-sweep step of bench.py runs 9 launches of depth 8 and 4 of
    depth 7), and a repeated run reproduces the same digests."""
    N = 16384
    digests = {}
    for kmax in (1, 4, 7, 8, 4):
        s = Solver(N, N, float(N - 1), float(N - 1))
        s.num_iterations, s.num_poisson_iterations = 2, 8
        s.set_sweeps_per_launch(kmax)
        s.upload_cavity(1.0)
        s.solve(); s.synchronize()
        d = tuple(fnv1a64(s.device_field(n)) for n in ("p", "u", "v"))
        lid = s.device_field("u")[:, N - 1]
        assert (lid == 1.0).all()
        s.close()
        digests.setdefault(kmax, d)
        assert digests[kmax] == d, "non-deterministic"
    assert digests[1] == digests[4] == digests[7] == digests[8]


def test_device_cavity_init_equals_host_setters(Solver):
    a = Solver(300, 200); a.num_iterations, a.num_poisson_iterations = 2, 12
    a.upload_cavity(1.0); a.solve(); a.migrateDeviceToHost()
    b = Solver(300, 200).set_cavity(1.0); b.num_iterations, b.num_poisson_iterations = 2, 12
    b.migrateSolve()
    H.compare_with(a, {n: b._hostf(n) for n in H.ALL_FIELDS}, tag="init")
    a.close(); b.close()


def test_cuda_graph_replay_matches_plain_launches(Solver, golden):
    rec = dict(golden["cavity129_S100_K50"]); rec["S"] = 7
    for K in (50, 52):   # 13 (odd: ping-pong parity flips every step) and 13+... launches
        rec["K"] = K
        a = H.golden_case_setup(rec, Solver); a.set_use_graph(True); a.migrateSolve()
        b = H.golden_case_setup(rec, Solver); b.migrateSolve()
        H.compare_with(a, {n: b._hostf(n) for n in H.ALL_FIELDS}, tag=f"graph/K{K}")
        assert a.launch_count() == b.launch_count()
        a.close(); b.close()


def _random_cases(n, seed, lo, hi):
    rng = np.random.default_rng(seed)
    return [(int(rng.integers(lo, hi)), int(rng.integers(lo, hi)), int(rng.integers(1, 24)), int(rng.integers(1, 9)))
            for _ in range(n)]


@pytest.mark.parametrize("X,Y,K,kmax", _random_cases(12, 7, 130, 1500))
def test_random_shapes_gpu(Solver, X, Y, K, kmax):
    """Non-multiple-of-tile shapes between 130 and 1500 per side, random sweep counts and depths."""
    o = Oracle(X, Y).set_cavity()
    o.time_step, o.num_iterations, o.num_poisson_iterations = 1e-6, 2, K
    s = Solver(X, Y).set_cavity()
    s.time_step, s.num_iterations, s.num_poisson_iterations = 1e-6, 2, K
    s.set_sweeps_per_launch(kmax)
    s.migrateSolve()
    o.solve()
    H.compare_with(s, o, tag=f"random/{X}x{Y}/K{K}/k{kmax}")
    s.close()


@pytest.mark.parametrize("N,divmode", [(4096, None), (2048, "1")])
def test_bench_configuration_against_the_oracle(Solver, N, divmode):
    """The code path bench.py times, checked against the oracle instead of against itself: unit spacing
    (domain (N-1) x (N-1), so hx = hy = 1: the power-of-two arithmetic mode, right-hand-side scaling
    skipped), dt = 0.001, K = 100 sweeps spread over 13 launches (9 of depth 8, 4 of depth 7), the
    wave-aligned chunk plan of a large grid, one timestep, all seven fields bit-equal. 4096^2 is the
    largest grid the serial oracle finishes in about a minute; the second case forces the general
    division-free arithmetic (FLUIDB200_DIVMODE=1) on the same path."""
    import os
    saved = os.environ.get("FLUIDB200_DIVMODE")
    if divmode is not None:
        os.environ["FLUIDB200_DIVMODE"] = divmode
    try:
        s = Solver(N, N, float(N - 1), float(N - 1)).set_cavity(1.0)
        s.time_step, s.num_iterations, s.num_poisson_iterations = 0.001, 1, 100
        s.set_sweeps_per_launch(8)
        s.migrateSolve()
        assert s.launch_count() >= 2 + 13
    finally:
        if saved is None:
            os.environ.pop("FLUIDB200_DIVMODE", None)
        else:
            os.environ["FLUIDB200_DIVMODE"] = saved
    o = Oracle(N, N, float(N - 1), float(N - 1)).set_cavity(1.0)
    assert o.hx == 1.0 and o.hy == 1.0
    o.time_step, o.num_iterations, o.num_poisson_iterations = 0.001, 1, 100
    o.solve()
    H.compare_with(s, o, tag=f"bench-path/{N}/divmode{divmode}")
    s.close()


def test_solve_on_a_stream_that_cannot_be_captured(Solver, golden):
    """Small grids replay their timesteps as CUDA graphs. The legacy default stream (NULL) cannot be
    captured: solve() must fall back to plain launches without tripping over the failed capture's
    error code, and give the same bits."""
    from fluid_simulation_b200 import _lib
    rec = dict(golden["cavity41_S10_K50"])
    s = H.golden_case_setup(rec, Solver)
    s.set_use_graph(True)
    _lib.check(s._lib, s._lib.fluidb200_set_stream(s._h, None))
    s.migrateSolve()
    H.compare_with(s, H.load_fields(golden["cavity41_S10_K50"]), tag="legacy-stream")
    s.close()


def test_migrate_solve_pipeline_at_size(Solver):
    """The chunk-pipelined fluidb200_migrate_solve on a grid where chunks, halos and the split Jacobi launches are
    all in play (4096 x 2048, K = 30, pinned host memory, in place): host results equal upload + solve + download."""
    import torch
    X, Y, K = 4096, 2048, 30
    rng = np.random.default_rng(5)
    st = [(0.01 * rng.standard_normal((X, Y))).astype(np.float32) for _ in range(3)]
    a = Solver(X, Y, float(X - 1), float(Y - 1)).set_cavity(1.0)
    a.num_iterations, a.num_poisson_iterations = 1, K
    a.upload_cavity(1.0)
    pinned = [torch.from_numpy(x.copy()).pin_memory() for x in st]
    views = [t.numpy() for t in pinned]
    k0 = a.launch_count()
    got = a.migrate_solve_uvp(*views, out=views)          # in place
    assert a.launch_count() - k0 > 3 * (2 + 4), "the chunk pipeline was not taken"
    b = Solver(X, Y, float(X - 1), float(Y - 1)).set_cavity(1.0)
    b.num_iterations, b.num_poisson_iterations = 1, K
    b.upload_cavity(1.0)
    for n, x in zip(("u", "v", "p"), st):
        b._hostf(n)[:] = x
    import ctypes as C
    from fluid_simulation_b200 import _lib
    fields = (C.c_void_p * _lib.NUM_FIELDS)()
    for i, n in enumerate(("u", "v", "p")):
        fields[i] = b._hostf(n).ctypes.data
    _lib.check(b._lib, b._lib.fluidb200_upload(b._h, fields, None, None))
    b.solve()
    for n, g in zip(("u", "v", "p"), got):
        H.assert_bit_equal(g, b.device_field(n), f"pipeline/{n}")
    for n in ("u_next", "v_next", "rhs", "p_next"):
        H.assert_bit_equal(a.device_field(n), b.device_field(n), f"pipeline-device/{n}")
    a.close(); b.close()


def test_more_cells_than_a_32_bit_index_holds(Solver):
    """65538 x 32768 = 2^31 + 65536 cells (8.6 GB per field; the reference's `int` cell index, ParallelNavierStokes.cu:22,
    overflows here): u, v, p after one step of 8 sweeps are bit-identical for blocking depths 8 and 3 -- different chunk plans,
    launches and halo rows over the same cells -- compared on the device through zero-copy views; the rows beyond element 2^31
    are written (the lid drives every row) and the lid column is intact at both ends."""
    import torch
    X, Y = 65538, 32768
    if torch.cuda.mem_get_info()[0] < 110e9:
        pytest.skip("needs ~90 GB of free device memory")
    kept = None
    for kmax in (8, 3):
        s = Solver(X, Y, float(X - 1), float(Y - 1))
        s.num_iterations, s.num_poisson_iterations = 1, 8
        s.set_sweeps_per_launch(kmax)
        s.upload_cavity(1.0)
        s.solve(); s.synchronize()
        views = {n: torch.as_tensor(s.device_array(n), device="cuda") for n in ("u", "v", "p")}
        assert views["u"].shape == (X, Y)
        assert bool((views["u"][:, Y - 1] == 1.0).all())
        tail = views["u"][X - 8: X - 1, Y - 64: Y - 1]                 # interior cells beyond element 2^31, right under the lid
        assert bool((tail != 0).any()), "rows beyond 2^31 elements were not computed"
        assert bool(torch.isfinite(views["p"][X - 4:]).all())
        if kept is None:
            kept = {n: v.clone() for n, v in views.items()}
        else:
            for n, v in views.items():
                assert torch.equal(kept[n], v), f"{n} differs between blocking depths 8 and 3"
        del views
        s.close()
    del kept
    torch.cuda.empty_cache()

