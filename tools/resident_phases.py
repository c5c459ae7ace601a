#!/usr/bin/env python
"""Cycles per phase of the resident solver, launch after launch (the state carries over, so later launches are the steady regime).
    FLUIDB200_RESIDENT_PROF=1 python tools/resident_phases.py [N] [launches] [steps per launch]"""
import ctypes as C, json, os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
from fluid_simulation_b200 import ParallelNavierStokes
os.environ.setdefault("FLUIDB200_RESIDENT_PROF", "1")
n = int(sys.argv[1]) if len(sys.argv) > 1 else 1000
L = int(sys.argv[2]) if len(sys.argv) > 2 else 6
S = int(sys.argv[3]) if len(sys.argv) > 3 else 200
names = ("halo+barrier", "tentative", "rhs", "sweeps", "corrector", "writeback")
ny = int(os.environ.get("NY", n))
s = ParallelNavierStokes(n, ny, 1.0, 1.0 * (ny - 1) / (n - 1))
s.time_step = float(np.float32(0.125) * np.float32(1.0 / (n - 1)) ** 2 / np.float32(0.1))
s.set_resident(True); s.upload_cavity(1.0)
for K in [50] * L + [10, 90, 50]:
    s.num_poisson_iterations, s.num_iterations = K, S
    s.synchronize(); t = time.perf_counter(); s.solve(); s.synchronize()
    us = (time.perf_counter() - t) * 1e6 / S
    buf = (C.c_uint64 * 24)()
    s._lib.fluidb200_resident_profile(s._h, buf)
    print(json.dumps({"grid": [n, ny], "K": K, "steps": S, "us_per_step": round(us, 1), "cycles_per_step": {k: round((buf[i] + (sum(buf[8:13]) if k == "sweeps" else 0)) / S) for i, k in enumerate(names)},
                      "cycles_per_sweep": {k: round(buf[8 + i] / S / K) for i, k in enumerate(("wait", "edge rows", "inner rows", "fix+cols", "barrier"))},
                      "polls_per_sweep": round(buf[13] / S / K, 2)}), flush=True)
s.close()
