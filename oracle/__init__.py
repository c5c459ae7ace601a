"""CPU oracle for the cavity projection timestep -- TEST INFRASTRUCTURE ONLY.

Only ``tests/``, ``__graft_entry__.smoke()`` and ``bench.py``'s ``cpu_baseline`` /
``--impl reference`` legs may import this package, and only as the checker (or as the
reported CPU arm). Nothing in ``fluid_simulation_b200`` imports it.

Two back ends, both driven through ctypes:

* :class:`Oracle` -- ``oracle/liboracle.so``: our plain-C restatement of
  ``SerialNavierStokes<T>::solve()`` (``/root/reference/fluid_cpp/cpp/SerialNavierStokes.cpp:49-67``).
* :func:`ref_solve` -- ``oracle/_ref/libref_fluid.so``: the reference's own Serial /
  Threaded solvers compiled unmodified from ``/root/reference`` (build container only;
  the file travels to the GPU box prebuilt).

Field layout everywhere: flat ``i = x*Y + y`` (``fluid_cpp/tests/test_main.cpp:76``), i.e. a
NumPy array of shape ``(X, Y)``.
"""
from __future__ import annotations

import ctypes as C
import os
import subprocess

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
LIB_ORACLE = os.path.join(HERE, "liboracle.so")
LIB_REF = os.path.join(HERE, "_ref", "libref_fluid.so")
LIB_REF_CUDA = os.path.join(HERE, "_ref", "libref_cuda.so")
REFERENCE_ROOT = "/root/reference/fluid_cpp"


def build(ref: bool = True) -> None:
    """Compile liboracle.so and, when the reference sources exist, oracle/_ref."""
    subprocess.check_call(["make", "-s", "-C", HERE, "oracle"])
    if ref and os.path.isdir(os.path.join(REFERENCE_ROOT, "cpp")):
        subprocess.check_call(["make", "-s", "-C", HERE, "ref"])
        if os.path.exists(os.environ.get("NVCC", "/usr/local/cuda/bin/nvcc")):
            subprocess.check_call(["make", "-s", "-C", HERE, "ref_cuda"], stderr=subprocess.DEVNULL)


def have_ref() -> bool:
    return os.path.exists(LIB_REF)


def _ctype(dtype):
    return C.c_float if np.dtype(dtype) == np.float32 else C.c_double


def _make_state_struct(ct):
    class State(C.Structure):
        _fields_ = (
            [("X", C.c_int), ("Y", C.c_int), ("hx", ct), ("hy", ct), ("dt", ct), ("nu", ct),
             ("rho", ct), ("K", C.c_int), ("S", C.c_int)]
            + [(n, C.c_void_p) for n in ("u", "v", "p", "u_next", "v_next", "p_next", "rhs",
                                         "u_bc", "v_bc", "p_bc", "u_set", "v_set", "p_set")]
        )
    return State


_StateF32 = _make_state_struct(C.c_float)
_StateF64 = _make_state_struct(C.c_double)
_lib = None


def lib():
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_ORACLE):
            build(ref=False)
        _lib = C.CDLL(LIB_ORACLE)
        _lib.oracle_fnv1a64.restype = C.c_uint64
        _lib.oracle_fnv1a64.argtypes = [C.c_void_p, C.c_size_t]
        _lib.oracle_element_length_f32.restype = C.c_float
        _lib.oracle_element_length_f32.argtypes = [C.c_float, C.c_int]
        _lib.oracle_element_length_f64.restype = C.c_double
        _lib.oracle_element_length_f64.argtypes = [C.c_double, C.c_int]
    return _lib


def fnv1a64(arr: np.ndarray) -> str:
    """FNV-1a-64 of the raw little-endian bytes, as 16 hex digits."""
    a = np.ascontiguousarray(arr)
    return "%016x" % lib().oracle_fnv1a64(a.ctypes.data, a.nbytes)


FIELDS = ("u", "v", "p", "u_next", "v_next", "p_next", "rhs")


class Oracle:
    """SoA state + the phases of the serial reference, one C call per phase.

    Mirrors the reference's public surface: ``time_step``, ``kinematic_viscosity``,
    ``density``, ``num_poisson_iterations``, ``num_iterations``
    (``NavierStokesSolver.h:22-26``) and ``set{U,V,P}BoundaryCondition``
    (``NavierStokesSolver.cpp:63-85`` -- the setter also seeds the live field).
    """

    def __init__(self, X, Y, wx=1.0, wy=1.0, dtype=np.float32):
        self.dtype = np.dtype(dtype)
        self.X, self.Y = int(X), int(Y)
        self._sfx = "f32" if self.dtype == np.float32 else "f64"
        el = getattr(lib(), "oracle_element_length_" + self._sfx)
        self.hx = el(wx, self.X)
        self.hy = el(wy, self.Y)
        self.time_step = 0.001
        self.kinematic_viscosity = 0.1
        self.density = 1.0
        self.num_poisson_iterations = 50
        self.num_iterations = 1000
        shape = (self.X, self.Y)
        for n in FIELDS + ("u_bc", "v_bc", "p_bc"):
            setattr(self, n, np.zeros(shape, self.dtype))
        for n in ("u_set", "v_set", "p_set"):
            setattr(self, n, np.zeros(shape, np.uint8))

    # -- reference-shaped setters ------------------------------------------------
    def setUBoundaryCondition(self, x, y, bc):
        self.u_bc[x, y] = bc; self.u[x, y] = bc; self.u_set[x, y] = 1
        return 0

    def setVBoundaryCondition(self, x, y, bc):
        self.v_bc[x, y] = bc; self.v[x, y] = bc; self.v_set[x, y] = 1
        return 0

    def setPBoundaryCondition(self, x, y, bc):
        self.p_bc[x, y] = bc; self.p[x, y] = bc; self.p_set[x, y] = 1
        return 0

    def set_cavity(self, lid_velocity=1.0):
        """Boundary set-up of fluid_cpp/main.cpp:43-55 (vectorised)."""
        X, Y = self.X, self.Y
        for f, top in (("u", lid_velocity), ("v", 0.0)):
            bc, live, flag = getattr(self, f + "_bc"), getattr(self, f), getattr(self, f + "_set")
            for sl, val in (((slice(None), 0), 0.0), ((slice(None), Y - 1), top),
                            ((0, slice(1, Y - 1)), 0.0), ((X - 1, slice(1, Y - 1)), 0.0)):
                bc[sl] = val; live[sl] = val; flag[sl] = 1
        self.p_bc[:, Y - 1] = 0.0; self.p[:, Y - 1] = 0.0; self.p_set[:, Y - 1] = 1
        return self

    # -- phases --------------------------------------------------------------------
    def _state(self):
        St = _StateF32 if self.dtype == np.float32 else _StateF64
        s = St()
        s.X, s.Y = self.X, self.Y
        s.hx, s.hy = self.hx, self.hy
        s.dt = float(self.dtype.type(self.time_step))
        s.nu = float(self.dtype.type(self.kinematic_viscosity))
        s.rho = float(self.dtype.type(self.density))
        s.K, s.S = int(self.num_poisson_iterations), int(self.num_iterations)
        for n in FIELDS + ("u_bc", "v_bc", "p_bc", "u_set", "v_set", "p_set"):
            a = getattr(self, n)
            assert a.flags.c_contiguous
            setattr(s, n, a.ctypes.data)
        return s

    def _call(self, name, *args):
        s = self._state()
        getattr(lib(), f"{name}_{self._sfx}")(C.byref(s), *args)
        return self

    def solve(self):
        return self._call("oracle_solve")

    def predictor(self):
        return self._call("oracle_predictor")

    def poisson_sweeps(self, sweeps):
        return self._call("oracle_poisson_sweeps", C.c_int(int(sweeps)))

    def corrector(self):
        return self._call("oracle_corrector")


# ------------------------------------------------------------------------------------------
_ref = None


def _ref_lib():
    global _ref
    if _ref is None:
        if not os.path.exists(LIB_REF):
            raise FileNotFoundError(
                f"{LIB_REF} missing: run `make -C oracle ref` where /root/reference exists")
        _ref = C.CDLL(LIB_REF)
        _ref.ref_solve.restype = C.c_int
        _ref.ref_solve.argtypes = ([C.c_int] * 4 + [C.c_double] * 5 + [C.c_int] * 2
                                   + [C.c_void_p] * 13 + [C.POINTER(C.c_double)])
    return _ref


REF_KINDS = {"serial": 0, "safe": 1, "threaded": 2}


def ref_solve(kind, X, Y, *, wx=1.0, wy=1.0, dt=0.001, nu=0.1, rho=1.0, K=50, S=1000,
              dtype=np.float32, bcs=None, lid_velocity=1.0, want=("u", "v", "p")):
    """Run the reference's own CPU solver. ``bcs`` is an :class:`Oracle`-like object whose
    ``*_set`` / ``*_bc`` arrays say which setters to call (default: the cavity).
    Returns ``(fields: dict, elapsed_ms)``; elapsed covers ``solve()`` only, like
    ``fluid_cpp/main.cpp:94-98``."""
    dtype = np.dtype(dtype)
    if bcs is None:
        bcs = Oracle(X, Y, wx, wy, dtype).set_cavity(lid_velocity)
    out = {n: np.zeros((X, Y), dtype) for n in want}
    order = ("u", "v", "p", "u_next", "v_next", "p_next", "rhs")
    ptrs = [out[n].ctypes.data if n in out else None for n in order]
    ms = C.c_double(0.0)
    rc = _ref_lib().ref_solve(
        REF_KINDS[kind], int(dtype == np.float64), X, Y, wx, wy, dt, nu, rho, K, S,
        bcs.u_set.ctypes.data, bcs.u_bc.ctypes.data, bcs.v_set.ctypes.data, bcs.v_bc.ctypes.data,
        bcs.p_set.ctypes.data, bcs.p_bc.ctypes.data, *ptrs, C.byref(ms))
    if rc != 0:
        raise RuntimeError(f"ref_solve returned {rc}")
    return out, ms.value


# ------------------------------------------------------------------------------------------
def have_ref_cuda() -> bool:
    return os.path.exists(LIB_REF_CUDA)


_ref_cuda = None


def ref_cuda_cavity(X, Y, *, wx=1.0, wy=1.0, dt=0.001, nu=0.1, rho=1.0, K=50, S=1000, warmup_steps=2, want_v=False):
    """The reference's own CUDA solver, ``ParallelNavierStokes<float>`` compiled unmodified for sm_100a
    (``oracle/_ref/libref_cuda.so``), on a lid-driven cavity: returns ``(elapsed_ms, v or None)``; elapsed
    covers ``solve(); cudaDeviceSynchronize()`` like ``fluid_cpp/cuda_main.cu:96-99``, after
    ``migrateHostToDevice()``. A same-box comparison for bench.py, never an oracle."""
    global _ref_cuda
    if _ref_cuda is None:
        _ref_cuda = C.CDLL(LIB_REF_CUDA)
        _ref_cuda.ref_cuda_cavity.restype = C.c_int
        _ref_cuda.ref_cuda_cavity.argtypes = [C.c_int, C.c_int] + [C.c_double] * 5 + [C.c_int] * 3 + [C.c_void_p, C.POINTER(C.c_double)]
    v = np.zeros((X, Y), np.float32) if want_v else None
    ms = C.c_double(0.0)
    rc = _ref_cuda.ref_cuda_cavity(X, Y, wx, wy, dt, nu, rho, K, S, warmup_steps, v.ctypes.data if want_v else None, C.byref(ms))
    if rc != 0:
        raise RuntimeError(f"ref_cuda_cavity: CUDA error {rc}")
    return ms.value, v
