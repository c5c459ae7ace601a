#!/usr/bin/env python
"""One process, one slab per GPU (fluidb200_multi_*): parity with the CPU oracle and timing.
    python tools/check_multi_gpu.py [ndev]        (under gpurun --gpus N)
Prints one JSON line; exits non-zero on a mismatch."""
import json, os, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import numpy as np
import torch
from fluid_simulation_b200 import MultiGpuNavierStokes
from oracle import Oracle

ndev = int(sys.argv[1]) if len(sys.argv) > 1 else torch.cuda.device_count()
res = {"devices": ndev, "parity": [], "timing": []}
for X, Y, S, K, kmax in [(1200, 1000, 2, 23, 8), (777, 2100, 2, 16, 5), (2048, 1024, 1, 16, 8)]:
    os.environ["FLUIDB200_KMAX"] = str(kmax)
    dt = float(np.float32(0.125) * np.float32(1.0 / (max(X, Y) - 1)) ** 2 / np.float32(0.1))
    m = MultiGpuNavierStokes(X, Y, 1.0, 1.0, np.float32, devices=tuple(range(ndev)))
    m.time_step, m.num_iterations, m.num_poisson_iterations = dt, S, K
    m.upload_cavity(1.0); m.solve()
    o = Oracle(X, Y).set_cavity(1.0)
    o.time_step, o.num_iterations, o.num_poisson_iterations = dt, S, K
    o.solve()
    ok = all(np.array_equal(m.device_field(n).view(np.uint32), getattr(o, n).view(np.uint32)) for n in ("u", "v", "p"))
    res["parity"].append({"grid": f"{X}x{Y}", "steps": S, "sweeps": K, "depth": kmax, "bit_equal": bool(ok)})
    m.close()
os.environ["FLUIDB200_KMAX"] = "8"
for X, Y, what in [(16384 * ndev, 16384, "weak: 16384 rows per GPU"), (32768, 32768, "strong: 32768^2")]:
    m = MultiGpuNavierStokes(X, Y, float(X - 1), float(Y - 1), np.float32, devices=tuple(range(ndev)))
    m.time_step, m.num_poisson_iterations = 0.001, 100
    m.upload_cavity(1.0)
    m.num_iterations = 2; m.solve(); m.synchronize()
    m.num_iterations = 5
    t0 = time.perf_counter(); m.solve(); t_issue = time.perf_counter() - t0; m.synchronize()
    ms = (time.perf_counter() - t0) * 1e3 / 5
    res["timing"].append({"what": what, "grid": f"{X}x{Y}", "ms_per_step": round(ms, 3), "gcell_updates_per_s": round(X * Y / ms / 1e6, 2),
                          "host_issue_ms_per_step": round(t_issue * 1e3 / 5, 3), "timing": "host clock around solve() + synchronize(), 5 steps"})
    m.close()
print(json.dumps(res))
sys.exit(0 if all(p["bit_equal"] for p in res["parity"]) else 1)
