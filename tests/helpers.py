"""Shared helpers for the parity tests (test infrastructure; may import oracle/)."""
import os

import numpy as np

GOLDEN_DIR = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")


def bits(a):
    """Bit pattern view for exact (==, sign-of-zero and NaN-payload aware) comparison."""
    a = np.ascontiguousarray(a)
    return a.view(np.uint32 if a.dtype == np.float32 else np.uint64)


def assert_bit_equal(got, want, name=""):
    got = np.asarray(got)
    want = np.asarray(want)
    assert got.shape == want.shape, f"{name}: shape {got.shape} != {want.shape}"
    bad = bits(got) != bits(want)
    if bad.any():
        idx = np.argwhere(bad)
        x, y = idx[0]
        raise AssertionError(
            f"{name}: {bad.sum()} of {bad.size} cells differ; first at [x={x},y={y}] "
            f"got {float(got[x, y]).hex()} want {float(want[x, y]).hex()}; "
            f"max abs diff {np.nanmax(np.abs(got.astype(np.float64) - want.astype(np.float64)))}")


def golden_case_setup(rec, cls):
    """Build a solver-like object `cls(X, Y, wx, wy, dtype)` configured as golden case `rec`."""
    from tests.golden.make_golden import general_bcs
    dtype = np.dtype(rec["dtype"])
    if rec["general"]:
        src = general_bcs(rec["X"], rec["Y"], dtype)
        s = cls(rec["X"], rec["Y"], 1.0, 1.0, dtype)
        for f, setter in (("u", s.setUBoundaryCondition), ("v", s.setVBoundaryCondition),
                          ("p", s.setPBoundaryCondition)):
            flag, val = getattr(src, f + "_set"), getattr(src, f + "_bc")
            for x, y in np.argwhere(flag):
                setter(int(x), int(y), float(val[x, y]))
    else:
        s = cls(rec["X"], rec["Y"], 1.0, 1.0, dtype)
        s.set_cavity(1.0)
    s.time_step = float.fromhex(rec["dt_hex"])
    s.num_iterations = rec["S"]
    s.num_poisson_iterations = rec["K"]
    return s


def load_fields(rec):
    return np.load(os.path.join(GOLDEN_DIR, rec["fields_file"]))


# ---------------------------------------------------------------------------------------------
# running the PRODUCT's C ABI (real CUDA library, or tests/emu's host build of the same code)
# ---------------------------------------------------------------------------------------------
EMU_DIR = os.path.join(os.path.dirname(os.path.abspath(__file__)), "emu")
EMU_LIB = os.path.join(EMU_DIR, "libfluidb200_emu.so")


def build_emu():
    import subprocess
    subprocess.check_call(["make", "-s", "-C", EMU_DIR])
    return EMU_LIB


def solver_factory(libpath):
    from fluid_simulation_b200.solver import ParallelNavierStokes

    class _S(ParallelNavierStokes):
        def __init__(self, X, Y, wx=1.0, wy=1.0, dtype=np.float32):
            super().__init__(X, Y, wx, wy, dtype, _libpath=libpath)
    return _S


ALL_FIELDS = ("u", "v", "p", "u_next", "v_next", "p_next", "rhs")


def edges_only_bcs(X, Y, dtype):
    """Per-cell (ARRAY-mode) boundary values/flags on all four edges, partly unset, unset corners,
    and NO interior flags: exercises the boundary logic inside the temporally blocked pipeline."""
    from oracle import Oracle
    o = Oracle(X, Y, 1.0, 1.0, dtype)
    for x in range(X):
        if x % 3 != 0 and 0 < x < X - 1:
            o.setUBoundaryCondition(x, Y - 1, 1.0 + 0.01 * x)
            o.setVBoundaryCondition(x, Y - 1, 0.0)
        if x % 4 != 1 and x != X - 1:
            o.setPBoundaryCondition(x, Y - 1, 0.01 * (x % 5))
        if x % 5 != 2 and x != 0:
            o.setUBoundaryCondition(x, 0, 0.0)
            o.setVBoundaryCondition(x, 0, 0.02)
        if x % 7 == 3:
            o.setPBoundaryCondition(x, 0, 0.25)
    for y in range(1, Y - 1):
        if y % 3 != 1:
            o.setUBoundaryCondition(0, y, 0.0)
            o.setVBoundaryCondition(0, y, -0.01)
        if y % 4 != 2:
            o.setUBoundaryCondition(X - 1, y, 0.03)
            o.setVBoundaryCondition(X - 1, y, 0.0)
        if y % 6 == 0:
            o.setPBoundaryCondition(0, y, -0.125)
        if y % 9 == 4:
            o.setPBoundaryCondition(X - 1, y, 0.5)
    return o


def copy_bcs(src, dst):
    """Replay the setter calls recorded in `src` (an Oracle) on `dst` (any solver-like object)."""
    for f, setter in (("u", dst.setUBoundaryCondition), ("v", dst.setVBoundaryCondition),
                      ("p", dst.setPBoundaryCondition)):
        flag, val = getattr(src, f + "_set"), getattr(src, f + "_bc")
        for x, y in np.argwhere(flag):
            setter(int(x), int(y), float(val[x, y]))
    for n in ("time_step", "kinematic_viscosity", "density", "num_iterations", "num_poisson_iterations"):
        setattr(dst, n, getattr(src, n))
    return dst


def compare_with(solver, want, fields=ALL_FIELDS, tag=""):
    """`want` is an Oracle or a dict/npz of golden fields."""
    for n in fields:
        w = want[n] if not hasattr(want, n) else getattr(want, n)
        assert_bit_equal(solver._hostf(n), w, f"{tag}:{n}")
