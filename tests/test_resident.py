"""The resident solver (csrc/resident_kernels.cuh): small grids with cavity-style boundaries run every
timestep of solve() in ONE cooperative launch, pressure in registers. GPU only (the host emulation has
no cooperative kernels; the rest of the suite pins FLUIDB200_RESIDENT=0 so that small grids keep
exercising the streaming kernels as well).

Expected values: the golden fields recorded from the reference, and the oracle on the same inputs;
bit-exact on all seven fields, like everywhere else."""
import os

import numpy as np
import pytest

from oracle import Oracle, fnv1a64
from tests import helpers as H

pytestmark = pytest.mark.gpu


@pytest.fixture
def env():
    saved = dict(os.environ)
    os.environ["FLUIDB200_RESIDENT"] = "1"
    yield os.environ
    os.environ.clear()
    os.environ.update(saved)


@pytest.fixture(scope="module")
def Solver():
    import torch
    assert torch.cuda.is_available()
    return H.solver_factory(None)


def _took_resident_path(s, calls=1):
    assert s.resident_launch_count() == calls, "solve() did not take the resident path"


@pytest.mark.parametrize("layout", ["auto", "patch4", "patch8"])
@pytest.mark.parametrize("name", ["cavity41_S1000_K50", "cavity41_S3_K7", "cavity41_S10_K50", "cavity67x45_S25_K30",
                                  "cavity45x131_S12_K33", "cavity41_f64_S100_K50"])
def test_golden_fields_resident(Solver, env, golden, name, layout):
    """auto: grids of at most 4096 cells take the one-CTA cell layout; patch4 / patch8 force the register-patch
    layout (4 or 8 rows per thread) that larger grids use."""
    if layout != "auto":
        env["FLUIDB200_RESIDENT_R"] = layout[-1]
    rec = golden[name]
    s = H.golden_case_setup(rec, Solver)
    s.migrateSolve()
    _took_resident_path(s)
    assert s.launch_count() == 1, "one kernel for the whole solve"
    H.compare_with(s, H.load_fields(rec), tag=f"resident/{name}")
    s.close()


@pytest.mark.parametrize("name", ["cavity129_S100_K50", "cavity257_S20_K100", "cavity100_S50_K50", "cavity300_S10_K100",
                                  "cavity100_f64_S20_K50"])
def test_golden_digests_resident(Solver, env, golden, name):
    """129 and 257 put column Y-1 alone at the start of a warp (the detached-edge rule); 300^2 holds
    true subnormals (exact-division fallback); 100^2 discriminates the fp64 Laplacian tail."""
    rec = golden[name]
    s = H.golden_case_setup(rec, Solver)
    s.migrateSolve()
    _took_resident_path(s)
    for n in H.ALL_FIELDS:
        assert fnv1a64(s._hostf(n)) == rec["digest"][n], f"resident/{name}:{n}"
    s.close()


def _random_state(X, Y, dtype, seed, scale=0.01):
    rng = np.random.default_rng(seed)
    return {n: (scale * rng.standard_normal((X, Y))).astype(dtype) for n in H.ALL_FIELDS}


def _setup_pair(Solver, X, Y, dtype, seed, p_edges="cavity", domain=None, scale=0.01):
    """A GPU solver and an oracle with the same random initial fields and cavity-style boundaries.
    p_edges: 'cavity' (lid Dirichlet 0, rest copy), 'all_const', 'none', 'mixed'."""
    dtype = np.dtype(dtype)
    wx, wy = domain if domain else (1.0, 1.0)
    st = _random_state(X, Y, dtype, seed, scale)
    pair = []
    for cls in (Solver, Oracle):
        s = cls(X, Y, wx, wy, dtype)
        for n in H.ALL_FIELDS:
            (s._hostf(n) if cls is Solver else getattr(s, n))[:] = st[n]
        for x in range(X):
            s.setUBoundaryCondition(x, 0, 0.0); s.setVBoundaryCondition(x, 0, 0.02)
            s.setUBoundaryCondition(x, Y - 1, 1.0); s.setVBoundaryCondition(x, Y - 1, 0.0)
        for y in range(1, Y - 1):
            s.setUBoundaryCondition(0, y, 0.0); s.setVBoundaryCondition(0, y, -0.01)
            s.setUBoundaryCondition(X - 1, y, 0.03); s.setVBoundaryCondition(X - 1, y, 0.0)
        if p_edges in ("cavity", "all_const", "mixed"):
            for x in range(X):
                s.setPBoundaryCondition(x, Y - 1, 0.0 if p_edges == "cavity" else 0.25)
        if p_edges == "all_const":
            for x in range(X):
                s.setPBoundaryCondition(x, 0, -0.5)
            for y in range(1, Y - 1):
                s.setPBoundaryCondition(0, y, 0.125); s.setPBoundaryCondition(X - 1, y, -0.125)
        if p_edges == "mixed":
            for y in range(1, Y - 1):
                s.setPBoundaryCondition(X - 1, y, 0.5)
        pair.append(s)
    return pair


@pytest.mark.parametrize("X,Y,dtype,S,K,p_edges", [
    (41, 41, "float32", 3, 5, "cavity"), (41, 41, "float32", 2, 1, "none"), (40, 37, "float32", 3, 4, "all_const"),
    (53, 129, "float32", 2, 7, "mixed"), (64, 257, "float32", 2, 6, "none"), (33, 130, "float32", 2, 9, "cavity"),
    (35, 65, "float64", 2, 5, "none"), (47, 200, "float64", 3, 6, "mixed"), (3, 3, "float32", 2, 2, "cavity"),
    (3, 700, "float32", 2, 3, "none"), (700, 3, "float32", 2, 3, "cavity"), (5, 4, "float64", 1, 2, "all_const"),
    (64, 64, "float32", 2, 8, "none"), (63, 65, "float64", 2, 3, "mixed"), (3, 1365, "float32", 3, 2, "all_const"),
])
@pytest.mark.parametrize("layout", ["auto", "patch4"])
def test_random_initial_state_and_pressure_edge_kinds(Solver, env, X, Y, dtype, S, K, p_edges, layout):
    """Random u, v, p, u_next, v_next (edges included: the first sweep reads the wall rows as uploaded, the
    right-hand side reads the stored edge u_next / v_next), every kind of pressure edge the resident path
    accepts, both precisions, degenerate boxes."""
    if layout != "auto":
        env["FLUIDB200_RESIDENT_R"] = layout[-1]
    s, o = _setup_pair(Solver, X, Y, dtype, seed=X * 1000 + Y, p_edges=p_edges)
    for q in (s, o):
        q.time_step, q.num_iterations, q.num_poisson_iterations = 1e-4, S, K
    s.migrateSolve()
    _took_resident_path(s)
    o.solve()
    H.compare_with(s, o, tag=f"resident-random/{X}x{Y}/{dtype}/{p_edges}")
    s.close()


@pytest.mark.parametrize("X,Y,R,wy,wx,dtype", [
    (300, 200, 4, 1, 4, "float32"), (300, 200, 4, 2, 3, "float32"), (200, 520, 4, 1, 8, "float32"), (97, 1030, 8, 2, 2, "float32"),
    (150, 257, 4, 1, 2, "float32"), (120, 300, 4, 2, 5, "float64"), (31, 129, 4, 1, 3, "float64"), (64, 64, 4, 1, 1, "float32"),
    (200, 130, 8, 2, 4, "float32"), (90, 700, 8, 1, 1, "float32"),
])
def test_forced_tilings_exchange_across_tiles(Solver, env, X, Y, R, wy, wx, dtype):
    """Small grids cut into many tiles on purpose (FLUIDB200_RESIDENT_R / _WY / _WX): rows and columns
    crossing tile borders through L2, partial last tiles, tiles of a single warp."""
    env["FLUIDB200_RESIDENT_R"], env["FLUIDB200_RESIDENT_WY"], env["FLUIDB200_RESIDENT_WX"] = str(R), str(wy), str(wx)
    s, o = _setup_pair(Solver, X, Y, dtype, seed=R * 100 + wy * 10 + wx)
    for q in (s, o):
        q.time_step, q.num_iterations, q.num_poisson_iterations = 1e-5, 3, 11
    s.migrateSolve()
    _took_resident_path(s)
    o.solve()
    H.compare_with(s, o, tag=f"resident-tiles/{X}x{Y}/R{R}/{wy}x{wx}")
    s.close()


def test_1000x1000_against_the_oracle(Solver, env):
    """BASELINE.json configs[1]'s grid with its 50 sweeps per step, stable time step, three steps."""
    X = Y = 1000
    dt = float(np.float32(0.125) * np.float32(1.0 / (X - 1)) ** 2 / np.float32(0.1))
    o = Oracle(X, Y).set_cavity()
    s = Solver(X, Y).set_cavity()
    for q in (s, o):
        q.time_step, q.num_iterations, q.num_poisson_iterations = dt, 3, 50
    s.migrateSolve()
    _took_resident_path(s)
    o.solve()
    H.compare_with(s, o, tag="resident/1000x1000")
    s.close()


@pytest.mark.parametrize("K", [1, 2, 3, 50])
def test_resident_continues_from_device_state_and_matches_streaming(Solver, env, golden, K):
    """Two solve() calls of S steps equal one of 2S, and the streaming kernels give the same bits
    (K < 3 takes the extra barrier before the corrector's writes)."""
    rec = dict(golden["cavity129_S100_K50"]); rec["K"] = K; rec["S"] = 6
    a = H.golden_case_setup(rec, Solver); a.migrateSolve()
    rec["S"] = 3
    b = H.golden_case_setup(rec, Solver); b.migrateHostToDevice(); b.solve(); b.solve(); b.migrateDeviceToHost()
    _took_resident_path(b, 2)
    c = H.golden_case_setup(rec, Solver); c.set_resident(False); c.num_iterations = 6; c.migrateSolve()
    assert c.resident_launch_count() == 0
    want = {n: a._hostf(n) for n in H.ALL_FIELDS}
    H.compare_with(b, want, tag=f"resident-continue/K{K}")
    H.compare_with(c, want, tag=f"resident-vs-streaming/K{K}")
    for q in (a, b, c):
        q.close()


def test_what_the_resident_path_declines(Solver, env, golden):
    """Interior flags, per-cell edges, a single sweep in total: solve() takes the streaming kernels."""
    s = H.golden_case_setup(golden["general53x37_S20_K30"], Solver)
    s.num_iterations = 2
    s.migrateSolve()
    assert s.resident_launch_count() == 0
    s.close()
    rec = dict(golden["cavity41_S1_K1"])
    s = H.golden_case_setup(rec, Solver)
    s.migrateSolve()
    assert s.resident_launch_count() == 0
    H.compare_with(s, H.load_fields(golden["cavity41_S1_K1"]), tag="single-sweep")
    s.close()


@pytest.mark.parametrize("X,Y,layout", [(40, 300, "auto"), (33, 257, "auto"), (33, 257, "patch8"), (40, 300, "patch4")])
def test_subnormal_front_resident(Solver, env, X, Y, layout):
    """The front of a starting flow underflows through the subnormal range: the flagged cells take their second evaluation
    per cell here (checked products for spacings 1/39, 1/299; the double Laplacian tail for the power-of-two spacings of 33 x 257)."""
    if layout != "auto":
        env["FLUIDB200_RESIDENT_R"] = layout[-1]
    S, K = 9, 2
    o = Oracle(X, Y).set_cavity()
    o.time_step, o.num_iterations, o.num_poisson_iterations = 1e-9, S, K
    s = Solver(X, Y).set_cavity()
    s.time_step, s.num_iterations, s.num_poisson_iterations = 1e-9, S, K
    s.migrateSolve()
    _took_resident_path(s)
    o.solve()
    tiny = np.finfo(np.float32).tiny
    assert np.any((np.abs(o.u_next) < tiny) & (o.u_next != 0)), "case no longer produces subnormals"
    H.compare_with(s, o, tag=f"resident/subnormal-front/{X}x{Y}/{layout}")
    s.close()


@pytest.mark.parametrize("X,Y,layout", [(41, 41, "auto"), (47, 200, "patch4"), (70, 150, "patch8")])
def test_fp64_numerators_out_of_range_take_the_divisions(Solver, env, X, Y, layout):
    """fp64 with a spacing that is not a power of two divides by FMA corrections (Ar<double>::quot), valid for numerators in
    [2^-800, 2^800): a state scaled to 1e-262 (next to boundary values of order one) puts most differences below that, so
    the flagged cells go through the IEEE divisions while their neighbours do not."""
    if layout != "auto":
        env["FLUIDB200_RESIDENT_R"] = layout[-1]
    s, o = _setup_pair(Solver, X, Y, "float64", seed=X + Y, p_edges="mixed", scale=1e-262)
    for q in (s, o):
        q.time_step, q.num_iterations, q.num_poisson_iterations = 1e-4, 2, 4
    s.migrateSolve()
    _took_resident_path(s)
    o.solve()
    H.compare_with(s, o, tag=f"resident-f64-tiny/{X}x{Y}/{layout}")
    s.close()

