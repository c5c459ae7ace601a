#!/usr/bin/env python
"""Milliseconds per phase of one timestep (predictor, K sweeps, corrector) at N x N, CUDA events, 5 steps after 3.
    [FLUIDB200_LIB=other.so] [FLUIDB200_DIVMODE=1] [DTYPE=f64] python tools/phase_times.py [N] [K]"""
import json, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
from fluid_simulation_b200 import ParallelNavierStokes, _lib
N = int(sys.argv[1]) if len(sys.argv) > 1 else 16384
K = int(sys.argv[2]) if len(sys.argv) > 2 else 100
lib = _lib.load()
DT = np.float64 if os.environ.get("DTYPE", "f32") == "f64" else np.float32
s = ParallelNavierStokes(N, N, float(N - 1), float(N - 1), dtype=DT)
s.num_poisson_iterations, s.num_iterations = K, 1
st = torch.cuda.current_stream()
_lib.check(lib, lib.fluidb200_set_stream(s._h, st.cuda_stream))
s.upload_cavity(1.0); s._push_params()
def step(ev=None):
    if ev: ev[0].record(st)
    _lib.check(lib, lib.fluidb200_step_predictor(s._h))
    if ev: ev[1].record(st)
    _lib.check(lib, lib.fluidb200_step_poisson(s._h, K))
    if ev: ev[2].record(st)
    _lib.check(lib, lib.fluidb200_step_corrector(s._h))
    if ev: ev[3].record(st)
for _ in range(3): step()
torch.cuda.synchronize()
evs = [[torch.cuda.Event(enable_timing=True) for _ in range(4)] for _ in range(5)]
for e in evs: step(e)
torch.cuda.synchronize()
m = lambda i: round(sum(e[i].elapsed_time(e[i + 1]) for e in evs) / len(evs), 4)
print(json.dumps({"dtype": os.environ.get("DTYPE", "f32"), "lib": os.environ.get("FLUIDB200_LIB", "default"), "divmode": os.environ.get("FLUIDB200_DIVMODE", "auto"), "N": N, "K": K,
                  "predictor_ms": m(0), "poisson_ms": m(1), "corrector_ms": m(2), "step_ms": round(m(0) + m(1) + m(2), 4)}))
s.close()
