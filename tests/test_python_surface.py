"""The Python face of the drop-in: ``solve_navier_stokes`` with the signature of the
reference's only Python entry point (python/navier_stokes.py:24-26), the class mirror, zero-copy device views.

Cross-check of SURVEY.md section 4 item 3: the reference's NumPy solver run for S + 1 iterations equals the C++
``<double>`` solver (here: the oracle, pinned to it) run for S steps to a few 1e-15 away from the four corners -- the
C++ setters seed the moving lid at t = 0 while the script starts from rest, and the corners follow different rules.
The NumPy result is a committed fixture (tests/golden/numpy41_S11_K50.npz, made by tests/golden/make_numpy_golden.py
from /root/reference); where the reference is present it is also regenerated live and compared."""
import os

import numpy as np
import pytest

from oracle import Oracle
from tests import helpers as H

S, K, N = 10, 50, 41
TOL = 4e-15   # a few float64 ulps of fields that reach ~4 (SURVEY.md measured 1.2e-15 on its run; this one: 1.6e-15)


def _numpy_fixture():
    return np.load(os.path.join(H.GOLDEN_DIR, "numpy41_S11_K50.npz"))


def _interior_mask():
    m = np.ones((N, N), bool)
    for c in ((0, 0), (0, N - 1), (N - 1, 0), (N - 1, N - 1)):
        m[c] = False
    return m


def _oracle_double():
    o = Oracle(N, N, 1.0, 1.0, np.float64).set_cavity(1.0)
    o.num_iterations, o.num_poisson_iterations = S, K
    return o.solve()


def test_numpy_reference_equals_cpp_double_with_the_one_step_offset():
    ref, o, m = _numpy_fixture(), _oracle_double(), _interior_mask()
    for n in ("u", "v", "p"):
        want = ref[n]                      # [y, x]
        got = getattr(o, n).T              # (X, Y) -> [y, x]
        assert np.abs(got - want)[m].max() <= TOL, n


def test_numpy_fixture_is_what_the_reference_script_produces():
    if not os.path.isdir("/root/reference/python"):
        pytest.skip("the reference tree is not on this machine")
    from tests.golden.make_numpy_golden import reference_solver
    _, _, u, v, p = reference_solver()(num_elements=N, num_iterations=S + 1, num_poisson_iterations=K)
    ref = _numpy_fixture()
    for n, a in (("u", u), ("v", v), ("p", p)):
        assert np.array_equal(a, ref[n]), n


BACKENDS = [pytest.param("emu"), pytest.param("cuda", marks=pytest.mark.gpu)]


@pytest.fixture(params=BACKENDS)
def libpath(request):
    if request.param == "emu":
        return H.build_emu()
    import torch
    assert torch.cuda.is_available()
    return None


def test_solve_navier_stokes_signature_and_values(libpath):
    """Same call as the reference's script; C++ semantics: bit-equal to the <double> oracle, hence within a few 1e-15
    of the NumPy script run one iteration longer (corners excluded); (X, Y) mesh and [y, x] orientation as there."""
    import inspect

    from fluid_simulation_b200 import solve_navier_stokes
    names = list(inspect.signature(solve_navier_stokes).parameters)[:9]
    assert names == ["num_elements", "domain_size", "num_iterations", "time_step", "kinematic_viscosity", "density",
                     "top_velocity", "num_poisson_iterations", "stability_safety_factor"]
    X, Y, u, v, p = solve_navier_stokes(N, 1.0, S, 0.001, 0.1, 1.0, 1.0, K, 0.5, _libpath=libpath)
    o, ref, m = _oracle_double(), _numpy_fixture(), _interior_mask()
    assert X.shape == (N, N) and np.array_equal(X[0], np.linspace(0.0, 1.0, N)) and np.array_equal(Y[:, 0], X[0])
    for n, a in (("u", u), ("v", v), ("p", p)):
        assert a.dtype == np.float64 and a.flags.c_contiguous
        H.assert_bit_equal(a, np.ascontiguousarray(getattr(o, n).T), f"solve_navier_stokes:{n}")
        assert np.abs(a - ref[n])[m].max() <= TOL
    assert np.all(u[-1, :] == 1.0)          # the lid is the last row, like in the script (navier_stokes.py:86)
    # float32 on request
    _, _, u32, _, _ = solve_navier_stokes(N, num_iterations=3, dtype=np.float32, _libpath=libpath)
    o32 = Oracle(N, N).set_cavity(1.0)
    o32.num_iterations = 3
    o32.solve()
    H.assert_bit_equal(u32, np.ascontiguousarray(o32.u.T), "solve_navier_stokes:f32")


def test_tiled_class_mirror(libpath):
    from fluid_simulation_b200 import TiledNavierStokes
    s = TiledNavierStokes(N, N, 1.0, 1.0, dtype=np.float32, _libpath=libpath).set_cavity(1.0)
    s.num_iterations, s.num_poisson_iterations = 4, 13
    s.migrateSolve()
    o = Oracle(N, N).set_cavity(1.0)
    o.num_iterations, o.num_poisson_iterations = 4, 13
    o.solve()
    H.compare_with(s, o, tag="TiledNavierStokes")
    out = np.empty(N * N, np.float32)
    assert s.getPValues(out) == 0 and np.array_equal(out.view(np.uint32), o.p.reshape(-1).view(np.uint32))
    s.close()


@pytest.mark.gpu
def test_zero_copy_device_views():
    import torch

    from fluid_simulation_b200 import ParallelNavierStokes
    s = ParallelNavierStokes(300, 77, 1.0, 1.0).set_cavity(1.0)
    s.time_step, s.num_iterations, s.num_poisson_iterations = 1e-5, 2, 7
    s.migrateHostToDevice()
    s.solve(); s.synchronize()
    for n in ("u", "v", "p", "rhs"):
        view = s.device_array(n)
        assert view.__cuda_array_interface__["strides"][0] == view.pitch * 4 and view.pitch % 32 == 0
        t = torch.as_tensor(view, device="cuda")
        assert t.shape == (300, 77) and t.data_ptr() == view.__cuda_array_interface__["data"][0]   # no copy
        assert np.array_equal(t.cpu().numpy().view(np.uint32), s.device_field(n).view(np.uint32))
    # writes through the view reach the solver
    t = torch.as_tensor(s.device_array("u"), device="cuda")
    t[5, 5] = 123.0
    torch.cuda.synchronize()
    assert s.device_field("u")[5, 5] == 123.0
    s.close()
