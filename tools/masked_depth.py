#!/usr/bin/env python
"""Microseconds per sweep of the general-BC (interior pressure flags) Jacobi path for each blocking depth.

    N=8192 python tools/masked_depth.py [k ...]        # default 1 2 3 4 6 8
"""
import ctypes as C
import json
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch
from fluid_simulation_b200 import ParallelNavierStokes, _lib

N = int(os.environ.get("N", 8192))
ks = [int(a) for a in sys.argv[1:]] or [1, 2, 3, 4, 6, 8]
lib = _lib.load()
s = ParallelNavierStokes(N, N, float(N - 1), float(N - 1), dtype=np.float32)
s.num_poisson_iterations, s.num_iterations = 100, 1
st = torch.cuda.current_stream()
_lib.check(lib, lib.fluidb200_set_stream(s._h, st.cuda_stream))
s.upload_cavity(1.0)
pset = np.zeros((N, N), np.uint8)
pval = np.zeros((N, N), np.float32)
pset[:, N - 1] = 1                                   # cavity: p = 0 on the lid
pset[N // 3:N // 3 + 64, N // 2:N // 2 + 64] = 1     # an obstacle with fixed pressure
pval[N // 3:N // 3 + 64, N // 2:N // 2 + 64] = 0.0625
vals = (C.c_void_p * 3)(None, None, pval.ctypes.data)
sets = (C.c_void_p * 3)(None, None, pset.ctypes.data)
_lib.check(lib, lib.fluidb200_upload(s._h, None, C.cast(vals, C.POINTER(C.c_void_p)), C.cast(sets, C.POINTER(C.c_void_p))))
s._push_params()
_lib.check(lib, lib.fluidb200_step_predictor(s._h))
s.set_sweeps_per_launch(8)
out = {}
for k in ks:
    os.environ["FLUIDB200_MASK_KMAX"] = str(k)   # (also switches the sparse-mask depth rule off)
    _lib.check(lib, lib.fluidb200_step_poisson(s._h, 2 * k))
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(st)
    _lib.check(lib, lib.fluidb200_step_poisson(s._h, 8 * k))
    e1.record(st)
    torch.cuda.synchronize()
    us = e0.elapsed_time(e1) * 1e3 / 8
    out[k] = {"us_per_launch": round(us, 1), "us_per_sweep": round(us / k, 1), "gcell_sweeps_per_s": round(N * N * k / us / 1e3, 1)}
s.close()
print(json.dumps(out))
