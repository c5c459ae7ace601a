#!/usr/bin/env python
"""fp64 build of the path at 8192^2, for the record: ms per phase."""
import os, sys, json
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
from fluid_simulation_b200 import ParallelNavierStokes, _lib
N = int(os.environ.get("N", 8192)); K = 100; kmax = int(os.environ.get("FLUIDB200_KMAX", 4))
lib = _lib.load()
out = {}
for dtype in (np.float64, np.float32):
    s = ParallelNavierStokes(N, N, float(N - 1), float(N - 1), dtype=dtype)
    s.num_poisson_iterations, s.num_iterations = K, 1
    s.set_sweeps_per_launch(kmax)
    st = torch.cuda.current_stream()
    _lib.check(lib, lib.fluidb200_set_stream(s._h, st.cuda_stream))
    s.upload_cavity(1.0); s._push_params()
    def step(ev=None):
        if ev: ev[0].record(st)
        lib.fluidb200_step_predictor(s._h)
        if ev: ev[1].record(st)
        lib.fluidb200_step_poisson(s._h, K)
        if ev: ev[2].record(st)
        lib.fluidb200_step_corrector(s._h)
        if ev: ev[3].record(st)
    for _ in range(3): step()
    torch.cuda.synchronize()
    evs = [[torch.cuda.Event(enable_timing=True) for _ in range(4)] for _ in range(5)]
    for e in evs: step(e)
    torch.cuda.synchronize()
    pred = sum(e[0].elapsed_time(e[1]) for e in evs) / 5; jac = sum(e[1].elapsed_time(e[2]) for e in evs) / 5
    corr = sum(e[2].elapsed_time(e[3]) for e in evs) / 5
    out[np.dtype(dtype).name] = {"grid": N, "kmax": kmax, "predictor_ms": pred, "jacobi_ms_100_sweeps": jac, "corrector_ms": corr,
                                 "gcell_sweeps_per_s": N * N * K / (jac * 1e-3) / 1e9,
                                 "jacobi_algorithmic_GBps": 3 * np.dtype(dtype).itemsize * N * N * K / (jac * 1e-3) / 1e9}
    s.close()
print(json.dumps(out))
