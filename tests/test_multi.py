"""The multi-GPU handle of the C ABI (fluidb200_multi_*, csrc/multi.cpp): one process, the grid cut into x-slabs.

CPU: one slab through the host emulation (the handle's bookkeeping, scatter / gather, phase order).
GPU (one device is all the gpu-marked suite may assume): one slab through the real library, scatter / gather over
several slabs. Runs with one slab per GPU -- halo pushes, flags, stream waits -- are made by tools/check_multi_gpu.py
under `gpurun --gpus N`, against the same oracle; their results are kept in profiles/. (Slabs that wait for each
other must not share a device: their streams can end up behind one another in the same hardware queue.)"""
import numpy as np
import pytest

from oracle import Oracle
from tests import helpers as H


def _oracle(X, Y, dtype, dt, S, K):
    o = Oracle(X, Y, 1.0, 1.0, np.dtype(dtype)).set_cavity(1.0)
    o.time_step, o.num_iterations, o.num_poisson_iterations = dt, S, K
    return o.solve()


def test_single_slab_through_the_host_emulation(monkeypatch):
    from fluid_simulation_b200 import MultiGpuNavierStokes
    monkeypatch.setenv("FLUIDB200_KMAX", "3")
    X, Y, S, K = 60, 45, 3, 7
    m = MultiGpuNavierStokes(X, Y, 1.0, 1.0, np.float32, devices=(0,), _libpath=H.build_emu())
    m.time_step, m.num_iterations, m.num_poisson_iterations = 2e-4, S, K
    m.upload_cavity(1.0)
    m.solve()
    o = _oracle(X, Y, "float32", 2e-4, S, K)
    for n in ("u", "v", "p", "u_next", "v_next", "rhs", "p_next"):
        H.assert_bit_equal(m.device_field(n), getattr(o, n), f"multi/1:{n}")
    m.close()


@pytest.mark.gpu
@pytest.mark.parametrize("X,Y,S,K,kmax,dtype", [(400, 300, 2, 9, 4, "float32"), (301, 257, 3, 5, 5, "float64"), (2048, 1536, 1, 17, 8, "float32")])
def test_one_slab_on_the_gpu_equals_the_oracle(monkeypatch, X, Y, S, K, kmax, dtype):
    from fluid_simulation_b200 import MultiGpuNavierStokes
    monkeypatch.setenv("FLUIDB200_KMAX", str(kmax))
    m = MultiGpuNavierStokes(X, Y, 1.0, 1.0, np.dtype(dtype), devices=(0,))
    m.time_step, m.num_iterations, m.num_poisson_iterations = 1e-5, S, K
    m.upload_cavity(1.0)
    m.solve()
    m.solve()     # continues from the device state
    o = _oracle(X, Y, dtype, 1e-5, 2 * S, K)
    for n in ("u", "v", "p", "u_next", "v_next", "rhs", "p_next"):
        H.assert_bit_equal(m.device_field(n), getattr(o, n), f"multi/gpu:{n}")
    assert m.launch_count() > 0
    m.close()


@pytest.mark.gpu
def test_multi_upload_scatter_and_download_gather():
    from fluid_simulation_b200 import MultiGpuNavierStokes
    X, Y = 300, 64
    rng = np.random.default_rng(3)
    st = {n: rng.standard_normal((X, Y)).astype(np.float32) for n in ("u", "v", "p")}
    m = MultiGpuNavierStokes(X, Y, 1.0, 1.0, np.float32, devices=(0, 0, 0))   # (no solve: nothing waits for anything)
    m.upload(**st)
    for n in st:
        assert np.array_equal(m.device_field(n), st[n])
    m.close()
