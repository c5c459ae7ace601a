#!/usr/bin/env python
"""Golden vectors from the reference's NumPy solver (python/navier_stokes.py:24-99), generated where
/root/reference exists:   python tests/golden/make_numpy_golden.py

The reference script imports matplotlib at module level (absent here and not needed by the solver), so an
empty stand-in module is put in sys.modules first; nothing of the reference is copied. Output:
tests/golden/numpy41_S11_K50.npz with u, v, p as the function returns them ([y, x] float64 arrays).
SURVEY.md section 4 item 3: this run of S + 1 = 11 iterations equals the C++ <double> solver with S = 10 steps to
about 1e-15 away from the four corner cells (the C++ setters seed the moving lid at t = 0, the script starts from rest)."""
import os
import sys
import types

import numpy as np

REF = "/root/reference/python"
HERE = os.path.dirname(os.path.abspath(__file__))


def reference_solver():
    for name in ("matplotlib", "matplotlib.pyplot"):
        sys.modules.setdefault(name, types.ModuleType(name))
    sys.path.insert(0, REF)
    try:
        import navier_stokes
        return navier_stokes.solve_navier_stokes
    finally:
        sys.path.remove(REF)


def main():
    solve = reference_solver()
    X, Y, u, v, p = solve(num_elements=41, num_iterations=11, num_poisson_iterations=50)
    np.savez_compressed(os.path.join(HERE, "numpy41_S11_K50.npz"), u=u, v=v, p=p)
    print("wrote numpy41_S11_K50.npz", u.shape, u.dtype, float(np.abs(u).max()))


if __name__ == "__main__":
    main()
