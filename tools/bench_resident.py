#!/usr/bin/env python
"""Resident solver vs streaming kernels on small grids: milliseconds for S timesteps x K sweeps.
    python tools/bench_resident.py [N ...]      (FLUIDB200_RESIDENT_R/_WY/_WX force a tiling)"""
import json, os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
from fluid_simulation_b200 import ParallelNavierStokes
sizes = [int(a) for a in sys.argv[1:]] or [41, 129, 300, 1000]
S, K = int(os.environ.get("S", 1000)), int(os.environ.get("K", 50))
for n in sizes:
    out = {"grid": n if int(os.environ.get("NY", n)) == n else [n, int(os.environ.get("NY", n))], "steps": S, "sweeps": K}
    for resident in (True, False):
        ny = int(os.environ.get("NY", n))
        s = ParallelNavierStokes(n, ny, 1.0, 1.0 * (ny - 1) / (n - 1))
        s.time_step = float(np.float32(0.125) * np.float32(1.0 / (n - 1)) ** 2 / np.float32(0.1))
        s.num_poisson_iterations = K
        s.set_resident(resident); s.upload_cavity(1.0)
        s.num_iterations = 10; s.solve(); s.synchronize()
        s.num_iterations = S; t = time.perf_counter(); s.solve(); s.synchronize()
        ms = (time.perf_counter() - t) * 1e3
        key = "resident" if resident else "streaming"
        out[key + "_ms"] = round(ms, 2)
        out[key + "_us_per_step"] = round(ms * 1e3 / S, 2)
        if resident:
            out["took_resident_path"] = s.resident_launch_count() == 2
            if os.environ.get("FLUIDB200_RESIDENT_PROF"):
                import ctypes as C
                buf = (C.c_uint64 * 24)()
                s._lib.fluidb200_resident_profile(s._h, buf)
                out["cycles_per_step"] = {k: round(buf[i] / S) for i, k in enumerate(("halo+barrier", "tentative", "rhs", "sweeps", "corrector", "writeback"))}
        s.close()
    print(json.dumps(out), flush=True)
