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
