"""Multi-GPU driver: the grid is cut into x-slabs, one process (rank) per GPU.

The reference has no multi-device path at all (no NCCL/MPI/cudaSetDevice anywhere, SURVEY.md
section 2.2); this is new work on top of the same kernels. Because the pressure solve is true Jacobi
(``SerialNavierStokes.cpp:140-198``) and every kernel here reproduces the serial arithmetic bit for
bit, the sharded result is bit-identical to the single-GPU one -- which is the correctness test.

Layout. Row ``x`` is ``pitch`` contiguous elements, so a slab ``[x0, x1)`` is one contiguous range
and a halo "line" is one contiguous run of ``pitch`` elements. Every rank stores its owned rows
plus ``H = kmax + 1`` halo rows per interior side; both y-edges (floor, lid) exist on every rank,
the x = 0 / x = X-1 walls only on the first / last rank, so all boundary rules stay rank-local.

Per timestep, with ``m = ceil(K / kmax)`` Jacobi launches of depth ``k <= kmax``:

  1. exchange u, v halos (depth kmax+1)                                     [1 exchange]
  2. predictor: u*, v* on owned +- kmax rows, rhs on owned +- (kmax-1) rows  (recomputed, not sent)
  3. m x { k sweeps on the two k-row edge strips of the owned range -> start sending them;
           k sweeps on the interior of the owned range, overlapping the transfer;
           wait for the neighbours' strips }                                 [m exchanges]
  4. corrector on the owned rows (p halo of depth 1 is already there)

Only point-to-point transfers between neighbours are needed (``torch.distributed`` isend/irecv:
NCCL over NVLink on GPUs, gloo on CPU for the tests); there is no reduction anywhere because the
iteration count is fixed. On CPU the "device" library is tests/emu's host build of the same C ABI
(test infrastructure), selected explicitly by the caller; the product path has no CPU fallback.
"""
from __future__ import annotations

import ctypes as C
import os

import numpy as np

from . import _lib

_FIELDS = ("u", "v", "p", "u_next", "v_next", "p_next", "rhs")


def slab_bounds(X, world, rank):
    """Owned rows [x0, x1) of `rank`: contiguous, sizes differ by at most one row."""
    return (rank * X) // world, ((rank + 1) * X) // world


class _DeviceBlock:
    """A cudaMalloc'ed range (fluidb200_device_alloc) exposed to torch through __cuda_array_interface__."""

    def __init__(self, ptr, shape, typestr):
        self.__cuda_array_interface__ = {"shape": tuple(shape), "typestr": typestr, "data": (int(ptr), False),
                                         "version": 2, "strides": None}


class SlabSolver:
    """One rank's slab of a lid-driven cavity (boundary set-up of fluid_cpp/main.cpp:43-55)."""

    def __init__(self, X, Y, domain_x, domain_y, dtype=np.float32, *, rank=0, world=1, group=None,
                 device="cuda", kmax=8, libpath=None, overlap=True):
        import torch
        self.torch = torch
        self.dist = torch.distributed if world > 1 else None
        self.group = group
        self.X, self.Y, self.rank, self.world = int(X), int(Y), int(rank), int(world)
        self.npdt = np.dtype(dtype)
        self.tdt = torch.float32 if self.npdt == np.float32 else torch.float64
        self.device = torch.device(device)
        self.on_gpu = self.device.type == "cuda"
        self.lib = _lib.load(libpath)
        if self.on_gpu:
            _lib.check(self.lib, self.lib.fluidb200_set_device(self.device.index or 0))
        self.kmax = int(kmax)
        self.overlap = bool(overlap) and self.on_gpu
        self.x0, self.x1 = slab_bounds(self.X, self.world, self.rank)
        self.H = self.kmax + 1
        if world > 1 and (self.x1 - self.x0) < 2 * self.H:
            raise ValueError(f"slab of {self.x1 - self.x0} rows is too thin for halo depth {self.H}")
        self.lo, self.hi = max(0, self.x0 - self.H), min(self.X, self.x1 + self.H)
        self.rows = self.hi - self.lo
        self.pitch = (self.Y + 31) // 32 * 32
        dt = int(self.npdt == np.float64)
        hx = self.lib.fluidb200_element_length(dt, float(domain_x), self.X)
        hy = self.lib.fluidb200_element_length(dt, float(domain_y), self.Y)
        self.slab = _lib.Slab(dt, self.X, self.Y, self.pitch, self.lo, self.rows, hx, hy)
        # Halo transport on GPUs: "push" (default) = every rank copies the rows its neighbour needs straight into
        # the neighbour's slab over NVLink (CUDA IPC mapping) and raises a flag there, all stream-ordered;
        # "nccl" = isend / irecv pairs through torch.distributed (also what the CPU tests use, over gloo).
        self.halo = "nccl"
        if self.on_gpu and world > 1 and os.environ.get("FLUIDB200_HALO", "push") == "push":
            self.halo = "push"
        self._block = None
        if self.halo == "push":
            self._setup_push_links()
        else:
            self.f = {n: torch.zeros((self.rows, self.pitch), dtype=self.tdt, device=self.device) for n in _FIELDS}
        self.phys = {n: i for i, n in enumerate(_FIELDS)}   # which physical array a name refers to (p / p_next swap)
        self.seq = 0
        self._fresh = True
        self.edges = {}
        self.time_step, self.kinematic_viscosity, self.density = 0.001, 0.1, 1.0
        self.num_poisson_iterations, self.num_iterations = 50, 1
        self.lid_velocity = 1.0
        self.launches = 0
        self.halo_bytes = 0
        if self.on_gpu:
            self.main = torch.cuda.current_stream(self.device)
            self.comm = torch.cuda.Stream(self.device)
            self.edge = torch.cuda.Stream(self.device, priority=-1)   # edge strips jump the queue,
            self.edge2 = torch.cuda.Stream(self.device, priority=-1)  # one stream per side so the two run side by side
            if os.environ.get("FLUIDB200_ONE_EDGE_STREAM"):          # (measurement knob)
                self.edge2 = self.edge
        else:
            self.main = self.comm = self.edge = self.edge2 = None
        self._set_edges(1.0)

    # -- one-sided halo links ----------------------------------------------------------------------
    def _setup_push_links(self):
        torch, lib = self.torch, self.lib
        item = self.npdt.itemsize
        self._field_bytes = self.rows * self.pitch * item
        total = len(_FIELDS) * self._field_bytes + 256           # + two flag words, 128 bytes apart
        base = C.c_void_p()
        _lib.check(lib, lib.fluidb200_device_alloc(C.byref(base), total))
        self._block = base.value
        typestr = "<f4" if item == 4 else "<f8"
        self._holders = [_DeviceBlock(self._block + i * self._field_bytes, (self.rows, self.pitch), typestr) for i in range(len(_FIELDS))]
        self.f = {n: torch.as_tensor(h, device=self.device) for n, h in zip(_FIELDS, self._holders)}
        self._flag = {"lo": self._block + len(_FIELDS) * self._field_bytes, "hi": self._block + len(_FIELDS) * self._field_bytes + 128}
        handle = C.create_string_buffer(64)
        _lib.check(lib, lib.fluidb200_ipc_export(C.c_void_p(self._block), handle))
        mine = (bytes(handle.raw), self.lo, self.rows)
        every = [None] * self.world
        self.dist.all_gather_object(every, mine, group=self.group)
        self._peer = {}
        for side, r in (("lo", self.rank - 1), ("hi", self.rank + 1)):
            if 0 <= r < self.world:
                h, plo, prows = every[r]
                ptr = C.c_void_p()
                _lib.check(lib, lib.fluidb200_ipc_open(h, C.byref(ptr)))
                fb = prows * self.pitch * item
                # I am the `hi` neighbour of the rank below me, and the `lo` neighbour of the rank above
                flag = ptr.value + len(_FIELDS) * fb + (128 if side == "lo" else 0)
                self._peer[side] = {"base": ptr.value, "lo": plo, "field_bytes": fb, "flag": flag}

    def _peer_rows_ptr(self, side, name, g0):
        """Address, in the neighbour's slab, of global row g0 of the physical array `name` currently refers to."""
        p = self._peer[side]
        return p["base"] + self.phys[name] * p["field_bytes"] + (g0 - p["lo"]) * self.pitch * self.npdt.itemsize

    def _push(self, names, depth, seq, streams=None):
        """Queue the copies of my `depth` outermost owned rows of each field into both neighbours' halo rows, each
        side followed by its flag. streams: {"lo": stream, "hi": stream} (default: the main stream)."""
        lib, item = self.lib, self.npdt.itemsize
        nbytes = depth * self.pitch * item
        for side, g0 in (("lo", self.x0), ("hi", self.x1 - depth)):
            if side not in self._peer:
                continue
            st = (streams or {}).get(side) or self.main
            stp = C.c_void_p(st.cuda_stream)
            for i, n in enumerate(names):
                src = C.c_void_p(self._rows(n, g0, g0 + depth).data_ptr())
                dst = C.c_void_p(self._peer_rows_ptr(side, n, g0))
                if i < len(names) - 1:
                    _lib.check(lib, lib.fluidb200_halo_copy(dst, src, nbytes, stp))
                else:
                    _lib.check(lib, lib.fluidb200_halo_push(dst, src, nbytes, C.c_void_p(self._peer[side]["flag"]), seq, stp))
                self.halo_bytes += nbytes
            self.launches += 1   # (the flag kernel)

    def _wait_pushed(self, seq, stream=None):
        st = C.c_void_p((stream or self.main).cuda_stream)
        for side in ("lo", "hi"):
            if side in self._peer:
                _lib.check(self.lib, self.lib.fluidb200_halo_wait(C.c_void_p(self._flag[side]), seq, st))

    # ------------------------------------------------------------------------------------------
    def _set_edges(self, lid):
        for name in ("u", "v", "p"):
            e = _lib.Edges()
            _lib.check(self.lib, self.lib.fluidb200_cavity_edges(name.encode(), float(lid), C.byref(e)))
            self.edges[name] = e

    def _stream(self):
        return C.c_void_p(self.main.cuda_stream) if self.on_gpu else None

    def _ptr(self, name):
        return C.c_void_p(self.f[name].data_ptr())

    def init_cavity(self, lid_velocity=1.0):
        """Zero state + cavity boundary values on this rank's rows (device-side, no host arrays)."""
        self.lid_velocity = float(lid_velocity)
        self._set_edges(lid_velocity)
        _lib.check(self.lib, self.lib.fluidb200_init_cavity(
            C.byref(self.slab), *[self._ptr(n) for n in _FIELDS], float(lid_velocity), self._stream()))
        self.launches += 1
        self._fresh = True

    # -- halo exchange ---------------------------------------------------------------------------
    def _rows(self, name, g0, g1):
        """Contiguous view of global rows [g0, g1) of a field."""
        return self.f[name][g0 - self.lo: g1 - self.lo]

    def _post_exchange(self, names, depth):
        """Start sending the `depth` outermost owned rows of each field to both neighbours and
        receiving theirs into the halo rows. Returns the list of work handles."""
        if self.world == 1:
            return []
        dist = self.dist
        ops = []
        for n in names:
            if self.rank > 0:
                ops.append(dist.P2POp(dist.isend, self._rows(n, self.x0, self.x0 + depth), self.rank - 1, self.group))
                ops.append(dist.P2POp(dist.irecv, self._rows(n, self.x0 - depth, self.x0), self.rank - 1, self.group))
            if self.rank < self.world - 1:
                ops.append(dist.P2POp(dist.isend, self._rows(n, self.x1 - depth, self.x1), self.rank + 1, self.group))
                ops.append(dist.P2POp(dist.irecv, self._rows(n, self.x1, self.x1 + depth), self.rank + 1, self.group))
        per_row = self.pitch * self.npdt.itemsize
        self.halo_bytes += sum(1 for o in ops if o.op is dist.isend) * depth * per_row
        return dist.batch_isend_irecv(ops) if ops else []

    @staticmethod
    def _wait(works):
        for w in works:
            w.wait()

    def exchange(self, names, depth):
        if self.halo == "push":
            self.seq += 1
            self._push(names, depth, self.seq)
            self._wait_pushed(self.seq)
            return
        self._wait(self._post_exchange(names, depth))

    # -- phases ----------------------------------------------------------------------------------
    def _params(self):
        T = self.npdt.type
        return float(T(self.time_step)), float(T(self.kinematic_viscosity)), float(T(self.density))

    def predictor(self):
        dt, nu, rho = self._params()
        k = self.kmax
        star_lo, star_hi = max(0, self.x0 - k), min(self.X, self.x1 + k)
        rhs_lo, rhs_hi = max(0, self.x0 - k + 1), min(self.X, self.x1 + k - 1)
        if self.world == 1:
            star_lo, star_hi, rhs_lo, rhs_hi = 0, self.X, 0, self.X
        _lib.check(self.lib, self.lib.fluidb200_predictor(
            C.byref(self.slab), self._ptr("u"), self._ptr("v"), self._ptr("u_next"), self._ptr("v_next"),
            self._ptr("rhs"), dt, nu, rho, star_lo, star_hi, rhs_lo, rhs_hi, self._stream()))
        self.launches += 1

    def _sweeps(self, k, lo, hi, stream=None):
        """k sweeps: reads f['p'], writes global rows [lo, hi) of f['p_next']."""
        st = self._stream() if stream is None else C.c_void_p(stream.cuda_stream)
        _lib.check(self.lib, self.lib.fluidb200_poisson_sweeps(
            C.byref(self.slab), self._ptr("p"), self._ptr("p_next"), self._ptr("rhs"), C.byref(self.edges["p"]),
            None, k, lo, hi, st))
        self.launches += 1

    def _swap_p(self):
        self.f["p"], self.f["p_next"] = self.f["p_next"], self.f["p"]   # ping-pong instead of p = p_next
        self.phys["p"], self.phys["p_next"] = self.phys["p_next"], self.phys["p"]

    def poisson(self, sweeps):
        """`sweeps` Jacobi sweeps, at most kmax per launch, one halo exchange after every launch.
        The exchange always moves kmax rows: enough for whatever depth the next launch uses, and for
        the corrector's p[x +- 1]."""
        torch = self.torch
        left = int(sweeps)
        d = self.kmax
        x0, x1 = self.x0, self.x1
        nlaunch = -(-left // self.kmax)   # launches of nearly equal depth, like step_poisson in csrc/api.cpp
        for i in range(nlaunch):
            k = -(-left // (nlaunch - i))
            left -= k
            if self.world == 1:
                self._sweeps(k, 0, self.X)
                self._swap_p()
                continue
            if not self.overlap:
                self._sweeps(k, x0, x1)
                self.exchange(["p_next"], d)
                self._swap_p()
                continue
            # Three independent pieces of one sweep launch: all read f['p'], write disjoint rows of f['p_next'].
            # 1. the d-row strips the neighbours are waiting for (none on a wall side), on a HIGH-PRIORITY
            #    stream so their few CTAs are scheduled ahead of the interior's
            lo_strip = (x0, x0 + d) if self.rank > 0 else None
            hi_strip = (x1 - d, x1) if self.rank < self.world - 1 else None
            start = torch.cuda.Event()
            start.record(self.main)
            ready = []
            for strip, stream in ((lo_strip, self.edge), (hi_strip, self.edge2)):
                if strip:
                    stream.wait_event(start)
                    self._sweeps(k, *strip, stream=stream)
                    ev = torch.cuda.Event()
                    ev.record(stream)
                    ready.append(ev)
            if self.halo == "push":
                # each strip goes to its neighbour from the stream that computed it, the flag right behind it; the
                # interior sweeps meanwhile; the next launch starts when both strips are done here and both
                # neighbours' strips have landed
                self.seq += 1
                self._push(["p_next"], d, self.seq, {"lo": self.edge, "hi": self.edge2})
                self._sweeps(k, lo_strip[1] if lo_strip else x0, hi_strip[0] if hi_strip else x1)
                for ev in ready:
                    self.main.wait_event(ev)
                self._wait_pushed(self.seq)
                self._swap_p()
                continue
            # 2. the interior of the owned range, concurrently, on the main stream
            self._sweeps(k, lo_strip[1] if lo_strip else x0, hi_strip[0] if hi_strip else x1)
            # 3. ship the strips on the communication stream while the interior runs
            for ev in ready:
                self.comm.wait_event(ev)
            with torch.cuda.stream(self.comm):
                works = self._post_exchange(["p_next"], d)
                self._wait(works)
                done = torch.cuda.Event()
                done.record(self.comm)
            self.main.wait_event(done)   # (comm waited for `ready`, so the strips are complete too)
            self._swap_p()

    def corrector(self):
        dt, _, rho = self._params()
        lo, hi = (0, self.X) if self.world == 1 else (self.x0, self.x1)
        _lib.check(self.lib, self.lib.fluidb200_corrector(
            C.byref(self.slab), self._ptr("p"), self._ptr("u_next"), self._ptr("v_next"), self._ptr("u"),
            self._ptr("v"), C.byref(self.edges["u"]), C.byref(self.edges["v"]), None, None, dt, rho, lo, hi,
            self._stream()))
        self.launches += 1

    def step(self):
        """One projection timestep (phase order of SerialNavierStokes.cpp:49-67)."""
        mark = self._phase_marks()
        if self.world > 1 and self.halo == "push" and self._fresh:
            # a neighbour's pushes may only arrive once this rank's set-up (init kernels, uploads) is done
            self.synchronize()
            self.dist.barrier(group=self.group)
            self._fresh = False
        if self.world > 1:
            if self.halo == "push" and int(self.num_poisson_iterations) == 0:
                # (with no sweep launches between them nothing orders one step's pushes behind the neighbour's reads)
                self.synchronize()
                self.dist.barrier(group=self.group)
            self.exchange(["u", "v"], self.H)
        mark("exchange_uv")
        self.predictor()
        mark("predictor")
        self.poisson(self.num_poisson_iterations)
        mark("poisson")
        self.corrector()
        mark("corrector")

    def _phase_marks(self):
        """FLUIDB200_SLAB_TIMING=1: CUDA events on the main stream around the phases of a step, collected
        for phase_ms(). Otherwise a no-op. (The first phase also absorbs any wait for a slower neighbour.)"""
        if not (self.on_gpu and os.environ.get("FLUIDB200_SLAB_TIMING")):
            return lambda name: None
        torch = self.torch
        if not hasattr(self, "_phase_events"):
            self._phase_events = []
        prev = [torch.cuda.Event(enable_timing=True)]
        prev[0].record(self.main)

        def mark(name):
            ev = torch.cuda.Event(enable_timing=True)
            ev.record(self.main)
            self._phase_events.append((name, prev[0], ev))
            prev[0] = ev
        return mark

    def phase_ms(self):
        """Milliseconds per phase summed over the steps since the last call (synchronizes)."""
        self.synchronize()
        out = {}
        for name, a, b in getattr(self, "_phase_events", []):
            out[name] = out.get(name, 0.0) + a.elapsed_time(b)
        self._phase_events = []
        return out

    def solve(self):
        for _ in range(int(self.num_iterations)):
            self.step()

    def synchronize(self):
        if self.on_gpu:
            self.torch.cuda.synchronize(self.device)

    # -- results ---------------------------------------------------------------------------------
    def owned(self, name):
        """This rank's owned rows of a field as a dense (x1-x0, Y) NumPy array."""
        self.synchronize()
        return self._rows(name, self.x0, self.x1)[:, : self.Y].cpu().numpy()

    def gather(self, name, dst=0):
        """The whole field on rank `dst` (None elsewhere); for tests and small grids."""
        mine = self.owned(name)
        if self.world == 1:
            return mine
        out = [None] * self.world if self.rank == dst else None
        self.dist.gather_object(mine, out, dst=dst, group=self.group)
        return np.concatenate(out, axis=0) if self.rank == dst else None

    def upload_owned(self, name, host_tensor):
        """Host (pinned) tensor of this rank's owned rows, shape (x1-x0, Y) -> device slab."""
        self._fresh = True
        self._rows(name, self.x0, self.x1)[:, : self.Y].copy_(host_tensor, non_blocking=True)

    def download_owned(self, name, host_tensor):
        host_tensor.copy_(self._rows(name, self.x0, self.x1)[:, : self.Y], non_blocking=True)


# ---------------------------------------------------------------------------------------------
def bench_distributed(args, grid, sweeps, kmax, prm, metric, unit):
    """bench.py's N > 1 arm. Default: weak scaling, `grid` rows per GPU (X = grid * N, Y = grid).
    FLUIDB200_SCALING=strong fixes the total at FLUIDB200_STRONG_GRID^2 (default 32768^2,
    BASELINE.json configs[3]) and splits its rows over the N ranks."""
    import os
    import time

    import torch
    import torch.distributed as dist

    rank = int(os.environ.get("RANK", 0))
    world = int(os.environ.get("WORLD_SIZE", 1))
    local = int(os.environ.get("LOCAL_RANK", rank))
    torch.cuda.set_device(local)
    # NCCL's own stream at high priority: the send/recv kernels of a halo exchange are a handful of CTAs that
    # would otherwise queue behind the interior sweep's thousands
    opts = None
    if not os.environ.get("FLUIDB200_NCCL_LOWPRIO"):                 # (measurement knob)
        try:
            opts = dist.ProcessGroupNCCL.Options()
            opts.is_high_priority_stream = True
        except Exception:
            opts = None
    dist.init_process_group("nccl", device_id=torch.device("cuda", local), pg_options=opts)
    strong = os.environ.get("FLUIDB200_SCALING", "weak") == "strong"
    if strong:
        X = Y = int(os.environ.get("FLUIDB200_STRONG_GRID", 32768))
        X = int(os.environ.get("FLUIDB200_STRONG_ROWS", X))   # (a shallower box, to study small slabs on few GPUs)
    else:
        X, Y = grid * world, grid
    s = SlabSolver(X, Y, float(X - 1), float(Y - 1), np.float32, rank=rank, world=world,
                   device=f"cuda:{local}", kmax=kmax)
    s.time_step, s.kinematic_viscosity, s.density = prm["dt"], prm["nu"], prm["rho"]
    s.num_poisson_iterations, s.num_iterations = sweeps, 1
    s.init_cavity(prm["lid"])
    warm = max(args.warmup, 3)
    for _ in range(warm):
        s.step()
    torch.cuda.synchronize()
    dist.barrier()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    l0, hb0 = s.launches, s.halo_bytes
    e0.record(s.main)
    t_issue = time.perf_counter()
    for _ in range(args.steps):
        s.step()
    t_issue = time.perf_counter() - t_issue   # host time to ISSUE the steps (nothing waits for the GPU in there)
    e1.record(s.main)
    torch.cuda.synchronize()
    dist.barrier()
    ms = torch.tensor([e0.elapsed_time(e1)], device=s.device)
    dist.all_reduce(ms, op=dist.ReduceOp.MAX)
    total_ms = float(ms.item())
    launches = torch.tensor([s.launches - l0], device=s.device)
    dist.all_reduce(launches)
    launches_total = int(launches.item())
    cells = X * Y

    # e2e: every step each rank uploads its owned u, v, p from pinned host memory and reads them back
    own = s.x1 - s.x0
    host = [torch.zeros((own, Y), dtype=torch.float32).pin_memory() for _ in range(3)]
    host[0][:, Y - 1] = prm["lid"]
    e2e_steps = max(1, min(args.steps, int(os.environ.get("FLUIDB200_E2E_STEPS", 3))))

    def e2e_step():
        for n, h in zip(("u", "v", "p"), host):
            s.upload_owned(n, h)
        s.step()
        for n, h in zip(("u", "v", "p"), host):
            s.download_owned(n, h)
        torch.cuda.synchronize()

    e2e_step()
    dist.barrier()
    t0 = time.perf_counter()
    for _ in range(e2e_steps):
        e2e_step()
    dist.barrier()
    t = torch.tensor([time.perf_counter() - t0], device=s.device)
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    e2e_s = float(t.item())
    finite = torch.tensor([int(bool(torch.isfinite(host[0]).all()))], device=s.device)
    dist.all_reduce(finite, op=dist.ReduceOp.MIN)
    assert int(finite.item()) == 1, "non-finite field after the e2e steps"
    halo_per_step = (s.halo_bytes - hb0) / max(1, args.steps + e2e_steps + 1)
    weak_rows = s.x1 - s.x0
    halo_mode = s.halo
    weak_desc = f"cavity {X}x{Y} fp32 ({weak_rows} rows per GPU, x-slabs), {sweeps} Poisson sweeps/step"

    # ---- strong scaling beside it (BASELINE.json configs[3]): a FIXED 32768^2 grid cut over the same ranks, and the same
    # grid on one GPU (rank 0 alone, the others wait), so that one invocation yields the speed-up
    strong_rec = None
    if not strong and not os.environ.get("FLUIDB200_SKIP_STRONG"):
        del s, host
        torch.cuda.empty_cache()
        SX = SY = int(os.environ.get("FLUIDB200_STRONG_GRID", 32768))
        ssteps = max(2, min(args.steps, 5))

        def timed(solver, n):
            for _ in range(2):
                solver.step()
            torch.cuda.synchronize()
            a0, a1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            a0.record(solver.main)
            for _ in range(n):
                solver.step()
            a1.record(solver.main)
            torch.cuda.synchronize()
            return a0.elapsed_time(a1) / n

        ss = SlabSolver(SX, SY, float(SX - 1), float(SY - 1), np.float32, rank=rank, world=world, device=f"cuda:{local}", kmax=kmax)
        ss.time_step, ss.kinematic_viscosity, ss.density = prm["dt"], prm["nu"], prm["rho"]
        ss.num_poisson_iterations = sweeps
        ss.init_cavity(prm["lid"])
        dist.barrier()
        t = torch.tensor([timed(ss, ssteps)], device=ss.device)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        strong_ms = float(t.item())
        srows = ss.x1 - ss.x0
        del ss
        torch.cuda.empty_cache()
        one_ms = None
        if rank == 0:
            try:
                s1 = SlabSolver(SX, SY, float(SX - 1), float(SY - 1), np.float32, rank=0, world=1, device=f"cuda:{local}", kmax=kmax)
                s1.time_step, s1.kinematic_viscosity, s1.density = prm["dt"], prm["nu"], prm["rho"]
                s1.num_poisson_iterations = sweeps
                s1.init_cavity(prm["lid"])
                one_ms = timed(s1, max(2, ssteps // 2))
                del s1
                torch.cuda.empty_cache()
            except Exception as ex:   # (e.g. another tenant's memory on the device)
                one_ms = None
                print(f"strong scaling: single-GPU {SX}^2 run failed: {ex!r}", file=__import__("sys").stderr)
        dist.barrier()
        strong_rec = {"grid": f"{SX}x{SY}", "rows_per_gpu": srows, "steps": ssteps, "ms_per_step": strong_ms,
                      "gcell_updates_per_s": SX * SY / (strong_ms * 1e-3) / 1e9,
                      "one_gpu_ms_per_step": one_ms, "speedup_vs_one_gpu": (one_ms / strong_ms) if one_ms else None,
                      "efficiency": (one_ms / strong_ms / world) if one_ms else None,
                      "timing": "CUDA events on each rank's launching stream, max over ranks; one-GPU figure: rank 0 alone, same build"}
    dist.destroy_process_group()
    if rank != 0:
        return None
    return {
        "metric": metric, "value": cells * args.steps / (total_ms * 1e-3) / 1e9, "unit": unit, "n_gpus": world,
        "steps": args.steps, "warmup": warm, "ms_per_step": total_ms / args.steps, "higher_is_better": True,
        "scaling": "strong" if strong else "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": {"workload": weak_desc, "strong": strong_rec,
                   "dt": prm["dt"], "nu": prm["nu"], "rho": prm["rho"], "hx": 1.0, "hy": 1.0, "sweeps_per_launch": kmax,
                   "host_issue_ms_per_step": t_issue / args.steps * 1e3,
                   "parallelism": f"x-slabs x{world}, halo depth {kmax + 1}, " + (
                       "edge strips pushed into the neighbour's slab over NVLink (IPC-mapped peer memory + flag, stream-ordered)"
                       if halo_mode == "push" else "NCCL send/recv") + " overlapped with the interior sweeps; NCCL for bootstrap",
                   "halo_bytes_per_rank_per_step": halo_per_step,
                   "l2": f"fields are {weak_rows * Y * 4 / 1e9:.2f} GB per GPU, >> 126 MB L2, no flush needed"},
        "roofline": None, "cpu_baseline": None,
        "e2e": {"value": cells * e2e_steps / e2e_s / 1e9, "unit": unit, "h2d_bytes_per_step": 3 * cells * 4,
                "d2h_bytes_per_step": 3 * cells * 4, "steps": e2e_steps, "ms_per_step": e2e_s / e2e_steps * 1e3,
                "host_link_GBps_aggregate": 2 * 3 * cells * 4 / (e2e_s / e2e_steps) / 1e9,
                "api": "per rank: copy owned u,v,p from pinned host + SlabSolver.step() + copy back (all ranks share the "
                       "host's memory and PCIe complex: host_link_GBps_aggregate is what they moved through it together)"},
        "gpu_launches": int(launches_total), "clocks": None,
    }
