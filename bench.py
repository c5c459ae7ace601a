#!/usr/bin/env python
"""bench.py -- headline benchmark of the cavity projection timestep on B200.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl b200|reference]
    (N > 1: python -m torch.distributed.run --nproc-per-node N ... bench.py --gpus N ...)

A "step" is ONE projection timestep (fused predictor, 100 Jacobi sweeps, fused corrector) of
the 16384 x 16384 fp32 lid-driven cavity per GPU (BASELINE.json configs[2]; weak scaling:
16384*N x 16384 over N x-slabs). Prints ONE JSON line (rank 0).

  value     Gcell-updates/s, device-resident state, CUDA events on the launching stream
  roofline  the Jacobi launch (k blocked sweeps): achieved = compulsory DRAM bytes of a launch (12 B x cells) / mean launch
            time measured with events inside the same timed steps; frac = that / MEASURED_PEAKS.json; .algorithmic = the
            12 B/cell/sweep figure (k times larger); .traffic = ncu DRAM bytes per launch (profiles/, dropped when the
            kernel sources changed since the capture); .other_kernels, .whole_step
  e2e       same metric through the C ABI with HOST buffers: fluidb200_migrate_solve on pinned arrays, every step
            (upload, compute, download pipelined over row chunks); .separate_calls = upload / solve / download unpipelined
  cpu_baseline  the reference's threaded C++ solver (oracle/_ref, its fixed 8 threads) and, under .serial, its serial
            solver (1 core), each on a bounded sample, on this box's host cores
  headline_general_spacing   the same step with a spacing that is not a power of two (general division-free arithmetic)
  config0_41x41, config1_1000x1000   BASELINE.json configs[0], [1] (1000 steps x 50 sweeps): resident solver, streaming
            kernels, the reference's CPU solvers and the reference's CUDA solver compiled for sm_100a, same box
  N > 1: value = weak scaling (16384 rows per GPU); config.strong = 32768^2 over the same ranks + on one GPU
--impl reference times the reference's own CPU implementation on this box's host cores.
"""
from __future__ import annotations

import argparse
import ctypes as C
import json
import os
import statistics
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

import numpy as np  # noqa: E402

GRID = int(os.environ.get("FLUIDB200_BENCH_GRID", 16384))
SWEEPS = int(os.environ.get("FLUIDB200_BENCH_SWEEPS", 100))
KMAX = int(os.environ.get("FLUIDB200_KMAX", 8))
METRIC = "Gcell-updates/s per timestep (cavity 16384x16384 fp32, 100 Poisson sweeps/step)"
UNIT = "Gcell-updates/s"
# stable, well-conditioned parameters for a 16384^2 grid: hx = hy = 1
DT, NU, RHO, LID = 0.001, 0.1, 1.0, 1.0


def peaks():
    try:
        with open(os.path.join(ROOT, "MEASURED_PEAKS.json")) as f:
            return float(json.load(f)["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
    except Exception:
        return 6650.0, "fallback (B200_PROFILING.md 6.65 TB/s)"


def kernel_source_hash():
    """sha1 over the kernel sources: stamps an ncu traffic table so that it is dropped when the kernels change."""
    import hashlib
    h = hashlib.sha1()
    d = os.path.join(ROOT, "fluid_simulation_b200", "csrc")
    for f in ("warp_kernels.cuh", "warp_args.h", "kernels.cu", "api.cpp"):
        with open(os.path.join(d, f), "rb") as fh:
            h.update(fh.read())
    return h.hexdigest()


def ncu_traffic(kernels):
    """Mean dram__bytes_read.sum + dram__bytes_write.sum per launch over `kernels`, from the newest table under
    profiles/ whose source stamp matches the kernel sources of this tree; (None, why) otherwise."""
    import glob
    for path in sorted(glob.glob(os.path.join(ROOT, "profiles", "ncu_traffic_r*.json")), reverse=True):
        try:
            with open(path) as f:
                table = json.load(f)
            if table.get("source_sha1") != kernel_source_hash():
                continue
            return sum(table[k] for k in kernels) / len(kernels), os.path.basename(path)
        except Exception:
            continue
    return None, "no ncu capture on file for these kernel sources"


class ClockSampler:
    """nvidia-smi clocks / throttle reasons sampled DURING the timed region."""
    Q = ("clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index=0):
        self.index, self.rows, self.proc = index, [], None

    def start(self):
        try:
            self.proc = subprocess.Popen(
                ["nvidia-smi", f"--id={self.index}", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits",
                 "-lms", "100"], stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.thread = threading.Thread(target=self._read, daemon=True)
            self.thread.start()
        except Exception:
            self.proc = None
        return self

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append([c.strip() for c in line.split(",")])

    def stop(self):
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.15)
        self.proc.terminate()
        try:
            self.proc.wait(timeout=5)
        except Exception:
            self.proc.kill()
        sm = [float(r[0]) for r in self.rows if len(r) >= 6 and r[0].replace(".", "").isdigit()]
        mx = [float(r[1]) for r in self.rows if len(r) >= 6 and r[1].replace(".", "").isdigit()]
        names = ("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap")
        reasons = sorted({n for r in self.rows if len(r) >= 6 for n, v in zip(names, r[2:6]) if v == "Active"})
        return {"sm_mhz": statistics.median(sm) if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "reasons": reasons, "samples": len(sm)}


# ---------------------------------------------------------------------------------------------
# reference arm / cpu baseline: the reference's own CPU solvers (oracle/_ref), bounded sample
# ---------------------------------------------------------------------------------------------
def cpu_reference_run(n, sweeps, steps, kind="threaded"):
    """Time `steps` timesteps of the reference CPU solver on an n x n cavity. Returns
    (Gcell-updates/s, descriptor). Uses oracle/_ref when present, else the oracle port."""
    import oracle
    dt = float(np.float32(DT))
    if oracle.have_ref():
        _, ms = oracle.ref_solve(kind, n, n, wx=float(n - 1), wy=float(n - 1), dt=dt, nu=NU, rho=RHO,
                                 K=sweeps, S=steps, want=())
        cores = 8 if kind == "threaded" else 1   # THREADED_GRID_SIZE, ThreadedNavierStokes.h:13
        what = "reference"
    else:
        o = oracle.Oracle(n, n, float(n - 1), float(n - 1)).set_cavity(LID)
        o.time_step, o.num_poisson_iterations, o.num_iterations = dt, sweeps, steps
        t = time.perf_counter(); o.solve(); ms = (time.perf_counter() - t) * 1e3
        cores, what, kind = 1, "port", "serial"
    val = n * n * steps / (ms * 1e-3) / 1e9
    return val, ms, {"kind": what, "cores": cores,
                     "sample": f"{n}x{n} cavity (hx=hy=1, dt={DT}), {sweeps} sweeps/step, {steps} step(s), "
                               f"{kind} C++ solver, solve() wall time {ms / 1e3:.1f} s; host has {os.cpu_count()} cpus"}


def run_reference_arm(args):
    rank = int(os.environ.get("RANK", 0))
    if rank != 0:
        return
    n = int(os.environ.get("FLUIDB200_REF_GRID", 1024))
    for _ in range(args.warmup):
        cpu_reference_run(n, SWEEPS, 1)
    t0 = time.perf_counter()
    total_ms = 0.0
    for _ in range(args.steps):
        _, ms, desc = cpu_reference_run(n, SWEEPS, 1)
        total_ms += ms
    val = n * n * args.steps / (total_ms * 1e-3) / 1e9
    desc["value"] = val
    desc["unit"] = UNIT
    try:   # the serial solver (SerialNavierStokes<float>::solve(), one core) on the same sample, once
        sv, _, sdesc = cpu_reference_run(n, SWEEPS, 1, kind="serial")
        sdesc.update({"value": sv, "unit": UNIT})
        desc["serial"] = sdesc
    except Exception as ex:
        desc["serial"] = {"value": None, "sample": repr(ex)}
    wall = time.perf_counter() - t0
    print(json.dumps({
        "impl": "reference", "metric": METRIC, "value": val, "unit": UNIT, "n_gpus": args.gpus,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": total_ms / args.steps,
        "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": {"workload": f"cavity {GRID}x{GRID} fp32, {SWEEPS} Poisson sweeps/step -- timed on a {n}x{n} sample of it "
                               f"(same spacing, dt and sweeps; per-cell throughput; the {GRID}^2 grid itself needs "
                               f"{GRID * GRID * 92 / 1e9:.0f} GB of the reference's 92-byte cells and ~450 s per step)",
                   "sample": desc["sample"], "wall_s": round(wall, 2)},
        "cpu_baseline": desc,
        "e2e": {"value": val, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }))


# ---------------------------------------------------------------------------------------------
# our arm
# ---------------------------------------------------------------------------------------------
def run_single(args):
    import torch

    from fluid_simulation_b200 import ParallelNavierStokes, _lib

    torch.cuda.set_device(0)
    lib = _lib.load()
    _lib.check(lib, lib.fluidb200_set_device(0))
    X = Y = GRID
    s = ParallelNavierStokes(X, Y, float(X - 1), float(Y - 1), dtype=np.float32)
    s.time_step, s.kinematic_viscosity, s.density = DT, NU, RHO
    s.num_poisson_iterations, s.num_iterations = SWEEPS, 1
    s.set_sweeps_per_launch(KMAX)
    stream = torch.cuda.current_stream()
    _lib.check(lib, lib.fluidb200_set_stream(s._h, stream.cuda_stream))  # so torch events bracket our launches
    s.upload_cavity(LID)
    s._push_params()
    h = s._h

    def step(ev=None):
        if ev is not None:
            ev[2].record(stream)
        _lib.check(lib, lib.fluidb200_step_predictor(h))
        if ev is not None:
            ev[0].record(stream)
        _lib.check(lib, lib.fluidb200_step_poisson(h, SWEEPS))
        if ev is not None:
            ev[1].record(stream)
        _lib.check(lib, lib.fluidb200_step_corrector(h))
        if ev is not None:
            ev[3].record(stream)

    for _ in range(max(args.warmup, 3)):
        step()
    torch.cuda.synchronize()

    clocks = ClockSampler(0).start()
    time.sleep(0.3)
    phase_ev = [tuple(torch.cuda.Event(enable_timing=True) for _ in range(4)) for _ in range(args.steps)]
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    launches0 = s.launch_count()
    torch.cuda.synchronize()
    e0.record(stream)
    for i in range(args.steps):
        step(phase_ev[i])
    e1.record(stream)
    torch.cuda.synchronize()
    clk = clocks.stop()
    launches = s.launch_count() - launches0
    total_ms = e0.elapsed_time(e1)
    jac_ms = sum(e[0].elapsed_time(e[1]) for e in phase_ev)
    pred_ms = sum(e[2].elapsed_time(e[0]) for e in phase_ev) / args.steps
    corr_ms = sum(e[1].elapsed_time(e[3]) for e in phase_ev) / args.steps
    cells = X * Y
    value = cells * args.steps / (total_ms * 1e-3) / 1e9

    # The solver spreads the sweeps of a step evenly over ceil(SWEEPS / KMAX) launches (100 sweeps at depth 8:
    # 9 launches of 8 and 4 of 7), and a launch costs about the same at either depth, so the roofline is
    # quoted for the MEAN launch: mean duration, mean depth.
    n_launch = (SWEEPS + KMAX - 1) // KMAX
    mean_depth = SWEEPS / n_launch
    jac_launch_ms = jac_ms / args.steps / n_launch
    alg_bytes = 12.0 * cells * mean_depth
    achieved = alg_bytes / (jac_launch_ms * 1e-3) / 1e9
    peak, peak_src = peaks()
    kernel = f"jacobi_fast_kernel<float,4,{KMAX}>"   # + the edge and wall kernels of the same split launch
    # ncu dram bytes per launch for exactly these kernels and this grid, when captures are on file: the mean over
    # the launches of one step (same rule as step_poisson: launch i takes ceil(left / launches left) sweeps)
    left, per_launch = SWEEPS, []
    for i in range(n_launch):
        k = -(-left // (n_launch - i))
        left -= k
        per_launch.append(f"jacobi_fast_kernel<float,4,{k}>@{X}x{Y}")
    traffic, traffic_src = ncu_traffic(per_launch)
    # A launch of k blocked sweeps must move each cell's p (read + write) and rhs (read) once: 12 B/cell per LAUNCH is the
    # compulsory DRAM traffic, and `frac` is that over the launch time over the measured copy peak -- a physical fraction
    # (<= 1). The algorithmic figure (12 B/cell per SWEEP, section 8(d) of SURVEY.md) is k times larger and is kept beside it.
    compulsory = 12.0 * cells
    dram_ach = compulsory / (jac_launch_ms * 1e-3) / 1e9
    roofline = {"bound": "hbm", "kernel": kernel, "achieved": dram_ach, "peak": peak, "unit": "GB/s", "frac": dram_ach / peak,
                "traffic": traffic, "traffic_source": traffic_src,
                "traffic_GBps": traffic / (jac_launch_ms * 1e-3) / 1e9 if traffic else None,
                "traffic_frac": traffic / (jac_launch_ms * 1e-3) / 1e9 / peak if traffic else None,
                "compulsory_bytes_per_launch": compulsory,
                "algorithmic": {"bytes_per_launch": alg_bytes, "achieved": achieved, "frac": achieved / peak,
                                "note": "12 B/cell/sweep x sweeps per launch / launch time: what a sweep-at-a-time kernel "
                                        "would have to move; exceeds the peak by the blocking depth, by design"},
                "peak_source": peak_src, "sweeps_per_launch": KMAX, "launches_per_step": n_launch, "mean_depth": mean_depth,
                "launch_ms": jac_launch_ms,
                "gcell_sweeps_per_s": cells * SWEEPS * args.steps / (jac_ms * 1e-3) / 1e9,
                "share_of_step": jac_ms / total_ms,
                "other_kernels": {"predictor_ms": pred_ms, "predictor_GBps": 20.0 * cells / (pred_ms * 1e-3) / 1e9,
                                  "predictor_frac": 20.0 * cells / (pred_ms * 1e-3) / 1e9 / peak,
                                  "corrector_ms": corr_ms, "corrector_GBps": 20.0 * cells / (corr_ms * 1e-3) / 1e9,
                                  "corrector_frac": 20.0 * cells / (corr_ms * 1e-3) / 1e9 / peak},
                "whole_step": {"bytes": (40.0 + 12.0 * n_launch) * cells, "GBps": (40.0 + 12.0 * n_launch) * cells / (total_ms / args.steps * 1e-3) / 1e9,
                               "frac": (40.0 + 12.0 * n_launch) * cells / (total_ms / args.steps * 1e-3) / 1e9 / peak,
                               "note": "compulsory DRAM bytes of a step (20 + 12 per Jacobi launch + 20 B/cell) over the step time"}}

    # ---- the same step with a spacing that is NOT a power of two (the reference's own h = 1/(n-1) is not): the general
    # division-free arithmetic with its flagged exact fallback, and the rhs scaling inside the sweeps
    general = None
    if not os.environ.get("FLUIDB200_SKIP_GENERAL"):
        saved = os.environ.get("FLUIDB200_DIVMODE")
        os.environ["FLUIDB200_DIVMODE"] = "1"
        try:
            step(); torch.cuda.synchronize()
            g0, g1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            gsteps = max(2, min(args.steps, 4))
            g0.record(stream)
            for _ in range(gsteps):
                step()
            g1.record(stream)
            torch.cuda.synchronize()
            gms = g0.elapsed_time(g1) / gsteps
            general = {"ms_per_step": gms, "value": cells / (gms * 1e-3) / 1e9, "unit": UNIT, "steps": gsteps,
                       "what": "FLUIDB200_DIVMODE=1: division-free arithmetic for any spacing (flagged exact fallback) in "
                               "predictor and corrector; the Jacobi sweeps are the same kernels"}
        finally:
            if saved is None:
                os.environ.pop("FLUIDB200_DIVMODE", None)
            else:
                os.environ["FLUIDB200_DIVMODE"] = saved

    # ---- e2e: host buffers in, host buffers out, every step ---------------------------------
    e2e_steps = max(1, min(args.steps, int(os.environ.get("FLUIDB200_E2E_STEPS", 3))))
    host = [torch.zeros((X, Y), dtype=torch.float32).pin_memory() for _ in range(3)]
    host[0][:, Y - 1] = LID
    fields = (C.c_void_p * _lib.NUM_FIELDS)()
    for i in range(3):
        fields[i] = host[i].data_ptr()

    io = (C.c_void_p * 3)(*[host[i].data_ptr() for i in range(3)])

    def e2e_step():       # the reference-facing call: migrateSolve on host arrays (one timestep, in place)
        _lib.check(lib, lib.fluidb200_migrate_solve(h, io, io))

    def e2e_step_serial():   # the same work as three separate calls: upload, solve, download x 3 (no overlap)
        _lib.check(lib, lib.fluidb200_upload(h, fields, None, None))
        _lib.check(lib, lib.fluidb200_solve(h))
        for i in range(3):
            _lib.check(lib, lib.fluidb200_download(h, i, host[i].data_ptr()))

    e2e_step()  # warm-up (allocates the pipeline's buffers)
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    for _ in range(e2e_steps):
        e2e_step()
    torch.cuda.synchronize()
    e2e_s = time.perf_counter() - t0
    e2e = {"value": cells * e2e_steps / e2e_s / 1e9, "unit": UNIT, "h2d_bytes_per_step": 3 * cells * 4,
           "d2h_bytes_per_step": 3 * cells * 4, "steps": e2e_steps, "ms_per_step": e2e_s / e2e_steps * 1e3,
           "link_GBps_each_way": 3 * cells * 4 / (e2e_s / e2e_steps) / 1e9,
           "api": "fluidb200_migrate_solve(u,v,p pinned host -> u,v,p pinned host), one timestep per call: upload, compute and "
                  "download pipelined over row chunks with K+2 recomputed halo rows (csrc/api.cpp)"}
    assert bool(torch.isfinite(host[0]).all()), "non-finite field after the e2e steps"
    e2e_step_serial()
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    for _ in range(e2e_steps):
        e2e_step_serial()
    torch.cuda.synchronize()
    ser_s = time.perf_counter() - t0
    e2e["separate_calls"] = {"value": cells * e2e_steps / ser_s / 1e9, "unit": UNIT, "ms_per_step": ser_s / e2e_steps * 1e3,
                             "api": "fluidb200_upload + fluidb200_solve + 3 x fluidb200_download, nothing overlapped (round 1's e2e)"}

    # The same calls from TWO host threads, each with its own solver handle, stream and pinned buffers (two
    # independent simulations in flight): the upload of one overlaps the download of the other on the full-duplex
    # link and both hide the compute. Reported beside the single-instance figure, which stays e2e.value.
    if not os.environ.get("FLUIDB200_SKIP_E2E_PAIR"):
        import threading
        s2 = ParallelNavierStokes(X, Y, float(X - 1), float(Y - 1), dtype=np.float32)
        s2.time_step, s2.kinematic_viscosity, s2.density = DT, NU, RHO
        s2.num_poisson_iterations, s2.num_iterations = SWEEPS, 1
        s2.set_sweeps_per_launch(KMAX)
        s2.upload_cavity(LID)
        s2._push_params()
        host2 = [torch.zeros((X, Y), dtype=torch.float32).pin_memory() for _ in range(3)]
        host2[0][:, Y - 1] = LID
        fields2 = (C.c_void_p * _lib.NUM_FIELDS)()
        for i in range(3):
            fields2[i] = host2[i].data_ptr()
        own = torch.cuda.Stream()   # the first handle was timed on torch's default stream; give it a stream of its own
        _lib.check(lib, lib.fluidb200_set_stream(h, own.cuda_stream))

        def worker(handle, flds, hst, n):
            ptrs = (C.c_void_p * 3)(*[hst[i].data_ptr() for i in range(3)])
            for _ in range(n):
                _lib.check(lib, lib.fluidb200_migrate_solve(handle, ptrs, ptrs))

        worker(s2._h, fields2, host2, 1)  # warm-up of the second handle
        torch.cuda.synchronize()
        per = max(2, e2e_steps)
        th = [threading.Thread(target=worker, args=(h, fields, host, per)),
              threading.Thread(target=worker, args=(s2._h, fields2, host2, per))]
        t0 = time.perf_counter()
        for t in th:
            t.start()
        for t in th:
            t.join()
        torch.cuda.synchronize()
        pair_s = time.perf_counter() - t0
        e2e["two_instances"] = {"value": cells * 2 * per / pair_s / 1e9, "unit": UNIT, "steps": 2 * per,
                                "ms_per_step": pair_s / (2 * per) * 1e3,
                                "api": "the same call from two host threads, one solver handle each"}
        assert bool(torch.isfinite(host2[0]).all())
        s2.close()
    s.close()

    # ---- BASELINE.json configs[0] and [1]: 41 x 41 and 1000 x 1000, 1000 timesteps x 50 sweeps (the shapes of the reference's
    # own test and of its published numbers). The default path (resident solver: one cooperative launch per solve()), and
    # the streaming kernels with and without CUDA-graph replay beside it. 1000^2 uses a stable dt (the reference's dt = 0.001
    # diverges to NaN at that resolution); 41^2 the reference's defaults.
    def small_grid(n, dt, reference_cuda_steps):
        rec = {"grid": f"{n}x{n}", "timesteps": 1000, "sweeps_per_step": 50, "dt": dt,
               "timing": "host clock around solve() + synchronize(), as fluid_cpp/cuda_main.cu:96-99"}
        sm = ParallelNavierStokes(n, n, 1.0, 1.0, dtype=np.float32)   # (on its own stream: the legacy stream cannot be captured)
        sm.time_step, sm.num_poisson_iterations = dt, 50
        for name, resident, graph in (("resident", True, False), ("streaming_graph", False, True), ("streaming_plain", False, False)):
            sm.set_resident(resident)
            sm.set_use_graph(graph)
            sm.upload_cavity(LID)
            sm.num_iterations = 10
            sm.solve(); sm.synchronize()      # warm-up (and graph capture)
            sm.upload_cavity(LID)
            sm.num_iterations = 1000
            k0 = sm.launch_count()
            t0 = time.perf_counter()
            sm.solve()
            sm.synchronize()
            ms = (time.perf_counter() - t0) * 1e3
            rec[name] = {"ms_total": ms, "us_per_step": ms, "gcell_updates_per_s": n * n * 1000 / (ms * 1e-3) / 1e9,
                         "kernel_launches": sm.launch_count() - k0}
            if name == "resident":
                rec["took_resident_path"] = sm.resident_launch_count() == 2
                rec["_v"] = sm.device_field("v")
        sm.close()
        rec["ms_total"] = rec["resident"]["ms_total"]
        rec["gcell_updates_per_s"] = rec["resident"]["gcell_updates_per_s"]
        # the reference's own code on this box: CPU (serial = 1 core; threaded = its fixed 8 threads) and its CUDA solver
        try:
            import oracle
            cpu_steps = 1000 if n <= 64 else 2
            for kind in ("serial", "threaded"):
                out, ms = oracle.ref_solve(kind, n, n, dt=dt, K=50, S=cpu_steps, want=("v",) if cpu_steps == 1000 else ())
                rec["reference_cpu_" + kind] = {"ms_total": ms * 1000 / cpu_steps, "steps_run": cpu_steps,
                                                "cores": 1 if kind == "serial" else 8, "extrapolated": cpu_steps != 1000}
                if kind == "serial" and cpu_steps == 1000:
                    rec["v_bit_equal_to_reference_serial"] = bool(np.array_equal(out["v"].view(np.uint32), rec["_v"].view(np.uint32)))
            if oracle.have_ref_cuda():
                ms, _ = oracle.ref_cuda_cavity(n, n, dt=dt, K=50, S=reference_cuda_steps, warmup_steps=2)
                rec["reference_cuda_sm100a"] = {
                    "ms_total": ms * 1000 / reference_cuda_steps, "steps_run": reference_cuda_steps,
                    "extrapolated": reference_cuda_steps != 1000,
                    "what": "the reference's ParallelNavierStokes<float> (fluid_cpp/cuda/ParallelNavierStokes.cu) compiled unmodified "
                            "with nvcc -gencode arch=compute_100a,code=sm_100a; solve() + cudaDeviceSynchronize() after migrateHostToDevice()"}
        except Exception as ex:
            rec["reference_error"] = repr(ex)
        rec.pop("_v", None)
        return rec

    small = small0 = None
    if not os.environ.get("FLUIDB200_SKIP_SMALL"):
        small0 = small_grid(41, 0.001, 1000)
        small = small_grid(1000, float(np.float32(0.125) * np.float32(1.0 / 999) ** 2 / np.float32(0.1)), 5)

    # ---- cpu baseline: bounded sample of the same workload on the host cores ---------------
    try:
        ngrid = int(os.environ.get("FLUIDB200_CPU_GRID", 1536))
        cv, _, cdesc = cpu_reference_run(ngrid, SWEEPS, 2)
        cdesc.update({"value": cv, "unit": UNIT})
        sv, _, sdesc = cpu_reference_run(ngrid // 2, SWEEPS, 1, kind="serial")
        sdesc.update({"value": sv, "unit": UNIT})
        cdesc["serial"] = sdesc
    except Exception as ex:  # the checker is optional for the measurement itself
        cdesc = {"value": None, "unit": UNIT, "cores": 0, "kind": "unavailable", "sample": repr(ex)}

    return {
        "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": 1, "steps": args.steps, "warmup": max(args.warmup, 3),
        "ms_per_step": total_ms / args.steps, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
        "dtype": "f32", "data": "synthetic",
        "config": {"workload": f"cavity {X}x{Y} fp32, {SWEEPS} Poisson sweeps/step, 1 GPU (BASELINE.json configs[2])",
                   "dt": DT, "nu": NU, "rho": RHO, "hx": 1.0, "hy": 1.0, "sweeps_per_launch": KMAX,
                   "division_mode": os.environ.get("FLUIDB200_DIVMODE", "2 (power-of-two spacing: exact scalings; "
                                                   "FLUIDB200_DIVMODE=1 forces the general division-free mode)"),
                   "l2": "fields are 1.07 GB each, >> 126 MB L2, no flush needed",
                   "algorithmic_GBps_step": (40 + 12 * SWEEPS) * cells * args.steps / (total_ms * 1e-3) / 1e9},
        "roofline": roofline, "cpu_baseline": cdesc, "e2e": e2e, "gpu_launches": int(launches), "clocks": clk,
        "headline_general_spacing": general, "config0_41x41": small0, "config1_1000x1000": small,
    }


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    args = ap.parse_args()
    if args.impl == "reference":
        run_reference_arm(args)
        return
    if args.gpus > 1 or int(os.environ.get("WORLD_SIZE", 1)) > 1:
        from fluid_simulation_b200.distributed import bench_distributed
        out = bench_distributed(args, GRID, SWEEPS, KMAX, dict(dt=DT, nu=NU, rho=RHO, lid=LID), METRIC, UNIT)
    else:
        out = run_single(args)
    if out is not None:
        print(json.dumps(out))


if __name__ == "__main__":
    main()
