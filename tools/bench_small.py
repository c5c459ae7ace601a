#!/usr/bin/env python
"""Small grids (BASELINE.json configs[0] and [1]): milliseconds for 1000 timesteps x 50 sweeps, plain launches vs
CUDA-graph replay, for several blocking depths.    N=41 python tools/bench_small.py [k ...]"""
import os, sys, time, json
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
from fluid_simulation_b200 import ParallelNavierStokes
n = int(os.environ.get("N", 1000)); out = {}
ks = [int(a) for a in sys.argv[1:]] or [2, 3, 5, 8]
for k in ks:
    for graph in (False, True):
        s = ParallelNavierStokes(n, n, 1.0, 1.0)
        s.time_step = float(np.float32(0.125) * np.float32(1.0 / (n - 1)) ** 2 / np.float32(0.1))
        s.num_poisson_iterations = 50
        s.set_sweeps_per_launch(k); s.set_use_graph(graph); s.upload_cavity(1.0)
        s.num_iterations = 10; s.solve(); s.synchronize()
        s.num_iterations = 1000; t = time.perf_counter(); s.solve(); s.synchronize()
        out[f"k{k}_{'graph' if graph else 'plain'}"] = round((time.perf_counter() - t) * 1e3, 1); s.close()
print(json.dumps(out))
