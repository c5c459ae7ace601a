#!/usr/bin/env python
"""Real multi-GPU parity check of the slab driver (one process per GPU over NCCL):

    python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 --master-port 29533 tools/check_slabs_gpu.py

Every rank solves its x-slab of a cavity; rank 0 gathers u, v, p and compares them bit for bit with the CPU
oracle (tests/ is where the oracle may be used; this tool is a test that needs more than one GPU, which the
gpu-marked pytest suite cannot assume). Prints one JSON line and exits non-zero on a mismatch."""
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import numpy as np
import torch
import torch.distributed as dist

from fluid_simulation_b200.distributed import SlabSolver

rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
local = int(os.environ.get("LOCAL_RANK", rank))
torch.cuda.set_device(local)
dist.init_process_group("nccl", device_id=torch.device("cuda", local))
results = []
for X, Y, S, K, kmax, dom in [(1200, 1000, 2, 23, 8, 1.0), (777, 2100, 2, 16, 5, 1.0), (2048, 1024, 1, 16, 8, None)]:
    wx, wy = (float(X - 1), float(Y - 1)) if dom is None else (dom, dom)
    dt = float(np.float32(0.125) * np.float32(min(wx / (X - 1), wy / (Y - 1))) ** 2 / np.float32(0.1))
    s = SlabSolver(X, Y, wx, wy, np.float32, rank=rank, world=world, device=f"cuda:{local}", kmax=kmax)
    s.time_step, s.num_poisson_iterations, s.num_iterations = dt, K, S
    s.init_cavity(1.0)
    s.solve()
    got = {n: s.gather(n) for n in ("u", "v", "p")}
    if rank == 0:
        from oracle import Oracle
        o = Oracle(X, Y, wx, wy).set_cavity(1.0)
        o.time_step, o.num_iterations, o.num_poisson_iterations = dt, S, K
        o.solve()
        ok = all(np.array_equal(got[n].view(np.uint32), np.asarray(getattr(o, n)).reshape(X, Y).view(np.uint32))
                 for n in ("u", "v", "p"))
        results.append({"grid": f"{X}x{Y}", "steps": S, "sweeps": K, "depth": kmax, "ranks": world, "bit_equal": bool(ok)})
    dist.barrier()
dist.destroy_process_group()
if rank == 0:
    print(json.dumps(results))
    sys.exit(0 if all(r["bit_equal"] for r in results) else 1)
