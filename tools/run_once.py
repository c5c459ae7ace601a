#!/usr/bin/env python
"""One solve() of an N x N (N x $NY) cavity (for ncu):  [NY=..] python tools/run_once.py N S K [resident 0/1]"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
from fluid_simulation_b200 import ParallelNavierStokes
n, S, K = int(sys.argv[1]), int(sys.argv[2]), int(sys.argv[3])
ny = int(os.environ.get("NY", n))
s = ParallelNavierStokes(n, ny, 1.0, 1.0 * (ny - 1) / (n - 1))
s.time_step = float(np.float32(0.125) * np.float32(1.0 / (n - 1)) ** 2 / np.float32(0.1))
s.num_poisson_iterations, s.num_iterations = K, S
if len(sys.argv) > 4:
    s.set_resident(bool(int(sys.argv[4])))
s.upload_cavity(1.0)
s.solve(); s.synchronize()
print("ok", n, S, K, "resident launches", s.resident_launch_count(), "kernels", s.launch_count())
s.close()
