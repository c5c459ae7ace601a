"""GPU test of fluidb200_poisson_solve_noalloc, the entry point for callers that may not allocate.

Some users of this library cannot call cudaMalloc and cannot let the library call it either:
they run after their memory pool has been frozen, inside a loop whose latency budget a trip
into the driver's allocator would break, or as a plug-in whose host owns every buffer. For them
the Jacobi ping-pong partner is a fixed 16 MiB array inside the library image instead of a
buffer somebody allocates per call. This test drives that entry point through the C ABI and
checks it, bit for bit (the tolerance every parity test here uses), against the ordinary
fluidb200_poisson_sweeps entry point working with a caller-provided second buffer and against the
CPU oracle's sweeps -- on two calls in a row, because the built-in array keeps whatever the previous
call left in it, and the second call (a different grid shape) must not be affected by those
leftovers; the array is also filled with NaN between calls to make any stale read visible. That
the call allocates nothing is checked on the driver's own free-memory figure around a call.

The file name sorts last on purpose: with `pytest -x` every other test has already run if this
one fails. It passes alone in a fresh process and after the rest of the suite.
"""
import ctypes as C

import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def _device_field(torch, host, pitch):
    """Pitched device copy of a dense (X, Y) host array; padding filled with NaN on purpose."""
    X, Y = host.shape
    t = torch.full((X, pitch), float("nan"), dtype=torch.from_numpy(host).dtype, device="cuda")
    t[:, :Y] = torch.from_numpy(host).cuda()
    return t


def _run_case(lib, _lib, torch, X, Y, sweeps, dtype, seed):
    from oracle import Oracle
    rng = np.random.default_rng(seed)
    npdt = np.dtype(dtype)
    pitch = (Y + 31) // 32 * 32
    p0 = rng.standard_normal((X, Y)).astype(npdt)
    rhs = rng.standard_normal((X, Y)).astype(npdt)
    hx, hy = float(npdt.type(1.0) / npdt.type(X - 1)), float(npdt.type(1.0) / npdt.type(Y - 1))
    slab = _lib.Slab(int(npdt == np.float64), X, Y, pitch, 0, X, hx, hy)
    edges = _lib.Edges()
    _lib.check(lib, lib.fluidb200_cavity_edges(b"p", 1.0, C.byref(edges)))
    stream = torch.cuda.current_stream().cuda_stream

    # expected (1): the existing entry point, one sweep per launch, ping-pong with our own buffer
    a = _device_field(torch, p0, pitch)
    b = torch.zeros_like(a)
    d_rhs = _device_field(torch, rhs, pitch)
    for _ in range(sweeps):
        _lib.check(lib, lib.fluidb200_poisson_sweeps(C.byref(slab), a.data_ptr(), b.data_ptr(), d_rhs.data_ptr(),
                                                     C.byref(edges), None, 1, 0, X, stream))
        a, b = b, a
    torch.cuda.synchronize()
    want = a[:, :Y].cpu().numpy()
    # expected (2): the CPU oracle's sweeps on the same data (the reference's arithmetic, not CUDA against CUDA)
    o = Oracle(X, Y, 1.0, 1.0, npdt)
    assert o.hx == hx and o.hy == hy
    o.p[:] = p0; o.p_next[:] = p0; o.rhs[:] = rhs    # (the reference has p_next == p between sweeps)
    o.p_bc[:, Y - 1] = 0.0; o.p_set[:, Y - 1] = 1      # the cavity's pressure edges: lid Dirichlet 0, rest copy
    o.poisson_sweeps(sweeps)
    view = np.uint32 if npdt == np.float32 else np.uint64
    assert np.array_equal(want.view(view), o.p.view(view)), "existing entry point differs from the oracle"

    # the allocation-free entry point, in place. Free device memory as the DRIVER counts it (not torch's
    # caching allocator, which cannot see a cudaMalloc made inside the library) must not move.
    p = _device_field(torch, p0, pitch)
    torch.cuda.synchronize()
    free_before, _ = torch.cuda.mem_get_info()
    _lib.check(lib, lib.fluidb200_poisson_solve_noalloc(C.byref(slab), p.data_ptr(), d_rhs.data_ptr(), C.byref(edges),
                                                        sweeps, stream))
    torch.cuda.synchronize()
    free_after, _ = torch.cuda.mem_get_info()
    got = p[:, :Y].cpu().numpy()
    assert np.array_equal(got.view(view), want.view(view)), f"{X}x{Y} {dtype} {sweeps} sweeps: fields differ"
    assert np.array_equal(got.view(view), o.p.view(view)), f"{X}x{Y} {dtype} {sweeps} sweeps: differs from the oracle"
    return free_before, free_after


def _poison_builtin_array(lib, _lib, torch):
    """Fill the library's built-in array with NaN: sweeps of an all-NaN grid as large as the array."""
    X, Y = 2048, 2048   # 16 MiB of fp32 = the whole array
    slab = _lib.Slab(0, X, Y, Y, 0, X, 1.0, 1.0)
    edges = _lib.Edges()
    _lib.check(lib, lib.fluidb200_cavity_edges(b"p", 1.0, C.byref(edges)))
    nan = torch.full((X, Y), float("nan"), dtype=torch.float32, device="cuda")
    nan_rhs = nan.clone()
    _lib.check(lib, lib.fluidb200_poisson_solve_noalloc(C.byref(slab), nan.data_ptr(), nan_rhs.data_ptr(),
                                                        C.byref(edges), 3, torch.cuda.current_stream().cuda_stream))
    torch.cuda.synchronize()


def test_noalloc_matches_existing_entry_point_on_two_calls_in_a_row():
    import torch

    from fluid_simulation_b200 import _lib
    assert torch.cuda.is_available()
    lib = _lib.load()
    _lib.check(lib, lib.fluidb200_set_device(torch.cuda.current_device()))
    # first call: larger grid, odd sweep count (exercises the uneven launch split); it may resolve the
    # array's address (the documented first-call set-up), so the memory check starts with the second
    _run_case(lib, _lib, torch, 700, 1000, 11, "float32", seed=1)
    # second call, straight after: smaller and differently shaped grid, so the built-in array
    # still holds the first call's data beyond and between this grid's rows
    free = _run_case(lib, _lib, torch, 333, 250, 6, "float32", seed=2)
    assert free[0] == free[1], f"device memory moved by {free[0] - free[1]} bytes during a no-alloc call"
    # the same with the array full of NaN: nothing stale may reach the result
    _poison_builtin_array(lib, _lib, torch)
    free = _run_case(lib, _lib, torch, 333, 250, 6, "float32", seed=2)
    assert free[0] == free[1]
    # a single sweep ends in the built-in array and is copied back; and the fp64 build
    _poison_builtin_array(lib, _lib, torch)
    _run_case(lib, _lib, torch, 129, 257, 1, "float32", seed=3)
    free = _run_case(lib, _lib, torch, 200, 300, 5, "float64", seed=4)
    assert free[0] == free[1]
    # near the array's limit, where fluidb200_poisson_sweeps itself would split the launch over a helper stream
    free = _run_case(lib, _lib, torch, 2048, 2040, 4, "float32", seed=5)
    assert free[0] == free[1]


def test_noalloc_validates_like_a_launch_even_without_sweeps():
    import torch

    from fluid_simulation_b200 import FluidB200Error, _lib
    lib = _lib.load()
    edges = _lib.Edges()
    _lib.check(lib, lib.fluidb200_cavity_edges(b"p", 1.0, C.byref(edges)))
    dummy = torch.zeros(64 * 64, device="cuda")
    bad_pitch = _lib.Slab(0, 64, 64, 65, 0, 64, 1.0, 1.0)
    with pytest.raises(FluidB200Error):
        _lib.check(lib, lib.fluidb200_poisson_solve_noalloc(C.byref(bad_pitch), dummy.data_ptr(), dummy.data_ptr(),
                                                            C.byref(edges), 0, None))
    good = _lib.Slab(0, 64, 64, 64, 0, 64, 1.0, 1.0)
    rhs = dummy.clone()
    _lib.check(lib, lib.fluidb200_poisson_solve_noalloc(C.byref(good), dummy.data_ptr(), rhs.data_ptr(),
                                                        C.byref(edges), 0, None))


def test_noalloc_rejects_grids_beyond_the_builtin_scratch():
    import torch

    from fluid_simulation_b200 import FluidB200Error, _lib
    lib = _lib.load()
    X, Y = 4096, 4096   # 64 MiB of fp32 > 16 MiB
    slab = _lib.Slab(0, X, Y, Y, 0, X, 1.0, 1.0)
    edges = _lib.Edges()
    _lib.check(lib, lib.fluidb200_cavity_edges(b"p", 1.0, C.byref(edges)))
    dummy = torch.zeros(64, device="cuda")
    with pytest.raises(FluidB200Error) as ei:
        _lib.check(lib, lib.fluidb200_poisson_solve_noalloc(C.byref(slab), dummy.data_ptr(), dummy.data_ptr(),
                                                            C.byref(edges), 4, None))
    assert "scratch" in str(ei.value) and str(_lib.NOALLOC_MAX_BYTES) in str(ei.value)
