#!/usr/bin/env python
"""Microseconds per Jacobi launch for each blocking depth k (16384^2 fp32 cavity unless N / DTYPE=f64 are set).

    python tools/jacobi_depth.py [k ...]        # default 4 5 6 7 8

One line of JSON: {"k": {"us_per_launch": .., "gcell_sweeps_per_s": ..}, ...}. Timed with CUDA events on the
solver's stream over 3 repetitions of 8 launches, after a warm-up.
"""
import json
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch
from fluid_simulation_b200 import ParallelNavierStokes, _lib

N = int(os.environ.get("N", 16384))
NX, NY = int(os.environ.get("NX", N)), int(os.environ.get("NY", N))   # (a slab-shaped grid: NX rows of NY columns)
DT = np.float64 if os.environ.get("DTYPE", "f32") == "f64" else np.float32
ks = [int(a) for a in sys.argv[1:]] or [4, 5, 6, 7, 8]
lib = _lib.load()
s = ParallelNavierStokes(NX, NY, float(NX - 1), float(NY - 1), dtype=DT)
s.num_poisson_iterations, s.num_iterations = 100, 1
st = torch.cuda.current_stream()
_lib.check(lib, lib.fluidb200_set_stream(s._h, st.cuda_stream))
s.upload_cavity(1.0)
s._push_params()
_lib.check(lib, lib.fluidb200_step_predictor(s._h))
out = {}
for k in ks:
    s.set_sweeps_per_launch(k)
    _lib.check(lib, lib.fluidb200_step_poisson(s._h, 4 * k))
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    reps, launches = 3, 8
    e0.record(st)
    for _ in range(reps):
        _lib.check(lib, lib.fluidb200_step_poisson(s._h, launches * k))
    e1.record(st)
    torch.cuda.synchronize()
    us = e0.elapsed_time(e1) * 1e3 / (reps * launches)
    out[k] = {"us_per_launch": round(us, 1), "gcell_sweeps_per_s": round(NX * NY * k / us / 1e3, 1)}
s.close()
print(json.dumps(out))
