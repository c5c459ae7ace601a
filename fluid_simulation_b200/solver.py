"""Python mirror of the reference's solver classes over the C ABI.

The reference ships no Python binding (its readme lists one as a goal); this module offers the
same class surface as ``fluid_cpp`` so the parity tests read like
``fluid_cpp/tests/test_main.cpp``:

* public fields ``time_step, kinematic_viscosity, density, num_poisson_iterations,
  num_iterations, stability_safety_factor`` (``fluid_cpp/cpp/NavierStokesSolver.h:22-27``),
* ``set{U,V,P}BoundaryCondition(x, y, BC)`` -- host state only, also seeds the live field
  (``NavierStokesSolver.cpp:63-85``),
* ``migrateHostToDevice / solve / migrateDeviceToHost / migrateSolve``
  (``fluid_cpp/cuda/ParallelNavierStokes.cuh:43-46``),
* ``get{U,V,P}Values, getUNextValues, getRightHand, getPNext`` (``NavierStokesSolver.cpp:87-145``).

Getters return / fill flat arrays with index ``x*Y + y``. The reference writes
``output[x*box_dimension_x + y]`` (``NavierStokesSolver.cpp:91``), which is the same thing for the
square grids it is only ever used with and overruns the buffer otherwise; ``x*Y + y`` is the
layout its device array (``ParallelNavierStokes.cu:22``) and its test (``test_main.cpp:76``) use.
"""
from __future__ import annotations

import ctypes as C

import numpy as np

from . import _lib

_FIELD_NAMES = ("u", "v", "p", "u_next", "v_next", "p_next", "rhs")


class ParallelNavierStokes:
    """Drop-in for ``ParallelNavierStokes<T>`` (``fluid_cpp/cuda/ParallelNavierStokes.cuh:16-47``)."""

    def __init__(self, box_dim_x, box_dim_y, domain_size_x=1.0, domain_size_y=1.0, dtype=np.float32,
                 _libpath=None):
        self.dtype = np.dtype(dtype)
        if self.dtype not in (np.dtype(np.float32), np.dtype(np.float64)):
            raise TypeError("dtype must be float32 or float64 (the reference instantiates exactly these)")
        self._lib = _lib.load(_libpath)
        self.box_dimension_x, self.box_dimension_y = int(box_dim_x), int(box_dim_y)
        T = self.dtype.type
        self.domain_size_x, self.domain_size_y = T(domain_size_x), T(domain_size_y)
        self.time_step = 0.001
        self.kinematic_viscosity = 0.1
        self.density = 1.0
        self.num_poisson_iterations = 50
        self.num_iterations = 1000
        self.stability_safety_factor = 0.5  # unused by every C++ solver of the reference
        self._h = C.c_void_p()
        _lib.check(self._lib, self._lib.fluidb200_create(
            C.byref(self._h), int(self.dtype == np.float64), self.box_dimension_x, self.box_dimension_y,
            float(self.domain_size_x), float(self.domain_size_y)))
        # host state: allocated lazily so 16384^2 grids driven by upload_cavity() never pay for it
        self._host = {}
        self._bc_val = {}
        self._bc_set = {}
        self._bc_dirty = False
        self._fields_dirty = set()

    # -- lifetime ---------------------------------------------------------------------------
    def close(self):
        if getattr(self, "_h", None) is not None and self._h.value:
            self._lib.fluidb200_destroy(self._h)
            self._h = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    # -- host state ---------------------------------------------------------------------------
    @property
    def shape(self):
        return (self.box_dimension_x, self.box_dimension_y)

    def _hostf(self, name):
        if name not in self._host:
            self._host[name] = np.zeros(self.shape, self.dtype)
        return self._host[name]

    def _bc(self, f):
        if f not in self._bc_val:
            self._bc_val[f] = np.zeros(self.shape, self.dtype)
            self._bc_set[f] = np.zeros(self.shape, np.uint8)
        return self._bc_val[f], self._bc_set[f]

    def _set_bc(self, f, x, y, bc):
        val, flag = self._bc(f)
        val[x, y] = bc
        flag[x, y] = 1
        self._hostf(f)[x, y] = bc  # the setter also overwrites the live field
        self._bc_dirty = True
        self._fields_dirty.add(f)
        return 0

    def setUBoundaryCondition(self, x, y, BC):
        return self._set_bc("u", x, y, BC)

    def setVBoundaryCondition(self, x, y, BC):
        return self._set_bc("v", x, y, BC)

    def setPBoundaryCondition(self, x, y, BC):
        return self._set_bc("p", x, y, BC)

    def setDomainSize(self, domain_size_x, domain_size_y):
        T = self.dtype.type
        self.domain_size_x, self.domain_size_y = T(domain_size_x), T(domain_size_y)
        _lib.check(self._lib, self._lib.fluidb200_set_domain(self._h, float(self.domain_size_x),
                                                             float(self.domain_size_y)))
        return 0

    def set_cavity(self, lid_velocity=1.0):
        """Vectorised form of the boundary loops of fluid_cpp/main.cpp:43-55 (host state)."""
        X, Y = self.shape
        for f, top in (("u", lid_velocity), ("v", 0.0)):
            val, flag = self._bc(f)
            live = self._hostf(f)
            for sl, v in (((slice(None), 0), 0.0), ((slice(None), Y - 1), top),
                          ((0, slice(1, Y - 1)), 0.0), ((X - 1, slice(1, Y - 1)), 0.0)):
                val[sl] = v; live[sl] = v; flag[sl] = 1
            self._fields_dirty.add(f)
        val, flag = self._bc("p")
        val[:, Y - 1] = 0.0; flag[:, Y - 1] = 1
        self._hostf("p")[:, Y - 1] = 0.0
        self._fields_dirty.add("p")
        self._bc_dirty = True
        return self

    # -- device ---------------------------------------------------------------------------
    def _push_params(self):
        T = self.dtype.type
        _lib.check(self._lib, self._lib.fluidb200_set_params(
            self._h, float(T(self.time_step)), float(T(self.kinematic_viscosity)), float(T(self.density)),
            int(self.num_poisson_iterations), int(self.num_iterations)))

    def migrateHostToDevice(self):
        """ParallelNavierStokes.cu:244-249: the whole host grid (all seven fields the kernels
        touch, plus boundary values and flags) goes to the device."""
        fields = (C.c_void_p * _lib.NUM_FIELDS)()
        keep = []
        for i, n in enumerate(_FIELD_NAMES):
            a = np.ascontiguousarray(self._hostf(n))
            keep.append(a)
            fields[i] = a.ctypes.data
        bcv = (C.c_void_p * 3)()
        bcs = (C.c_void_p * 3)()
        for i, f in enumerate(("u", "v", "p")):
            val, flag = self._bc(f)
            bcv[i] = val.ctypes.data
            bcs[i] = flag.ctypes.data
        _lib.check(self._lib, self._lib.fluidb200_upload(self._h, fields, bcv, bcs))
        self._bc_dirty = False
        self._fields_dirty.clear()

    def upload_cavity(self, lid_velocity=1.0):
        """Device-side cavity set-up (no host arrays): for grids too large for a host copy."""
        _lib.check(self._lib, self._lib.fluidb200_upload_cavity(self._h, float(lid_velocity)))

    def solve(self):
        """ParallelNavierStokes.cu:269-288: asynchronous; advances the DEVICE state."""
        self._push_params()
        _lib.check(self._lib, self._lib.fluidb200_solve(self._h))

    def synchronize(self):
        _lib.check(self._lib, self._lib.fluidb200_synchronize(self._h))

    def migrateDeviceToHost(self):
        """ParallelNavierStokes.cu:252-257 (blocking)."""
        for i, n in enumerate(_FIELD_NAMES):
            a = self._hostf(n)
            _lib.check(self._lib, self._lib.fluidb200_download(self._h, i, a.ctypes.data))

    def migrateSolve(self):
        self.migrateHostToDevice()
        self.solve()
        self.migrateDeviceToHost()

    def migrate_solve_uvp(self, u, v, p, out=None):
        """``migrateSolve`` on caller-owned host arrays of shape (X, Y): upload u, v, p, run ``solve()``, return
        the new (u, v, p) -- written into ``out`` (three arrays) when given. Boundary description and the other
        device fields stay as they are. One timestep of a tall grid is pipelined over row chunks (upload,
        compute and download of different chunks overlap); pass pinned memory (e.g. ``torch`` pinned tensors
        viewed as NumPy) to get that overlap."""
        self._push_params()
        ins = [np.ascontiguousarray(a, dtype=self.dtype) for a in (u, v, p)]
        outs = list(out) if out is not None else [np.empty(self.shape, self.dtype) for _ in range(3)]
        pi = (C.c_void_p * 3)(*[a.ctypes.data for a in ins])
        po = (C.c_void_p * 3)(*[a.ctypes.data for a in outs])
        _lib.check(self._lib, self._lib.fluidb200_migrate_solve(self._h, pi, po))
        return tuple(outs)

    # phase-level access (the reference's protected wrappers, ParallelNavierStokes.cu:297-334)
    def unifiedApproximateTimeStepAndRightHand(self):
        self._push_params()
        _lib.check(self._lib, self._lib.fluidb200_step_predictor(self._h))

    def poissonSweeps(self, sweeps):
        self._push_params()
        _lib.check(self._lib, self._lib.fluidb200_step_poisson(self._h, int(sweeps)))

    def unifiedVelocityCorrection(self):
        self._push_params()
        _lib.check(self._lib, self._lib.fluidb200_step_corrector(self._h))

    def set_sweeps_per_launch(self, k):
        _lib.check(self._lib, self._lib.fluidb200_set_sweeps_per_launch(self._h, int(k)))

    def set_use_graph(self, enable=True):
        _lib.check(self._lib, self._lib.fluidb200_set_use_graph(self._h, int(bool(enable))))

    def set_resident(self, enable=True):
        """Small grids with cavity-style boundaries: run solve() as one cooperative launch (default on)."""
        _lib.check(self._lib, self._lib.fluidb200_set_resident(self._h, -1 if enable is None else int(bool(enable))))

    def resident_launch_count(self):
        return int(self._lib.fluidb200_resident_launch_count(self._h))

    def launch_count(self):
        return int(self._lib.fluidb200_launch_count(self._h))

    def device_array(self, name):
        """Zero-copy view of one DEVICE field: an object with ``__cuda_array_interface__`` (shape (X, Y), row
        stride = the device pitch), accepted by ``torch.as_tensor(..., device="cuda")``, CuPy and Numba without
        a copy. The view follows the solver's memory: it is valid until the next call that swaps or frees the
        field's buffer (``solve()`` swaps p and p_next; ``close()`` frees everything), so take it again after those."""
        slab = _lib.Slab()
        _lib.check(self._lib, self._lib.fluidb200_get_slab(self._h, C.byref(slab)))
        ptr = self._lib.fluidb200_field_ptr(self._h, _FIELD_NAMES.index(name))
        if not ptr:
            raise ValueError(f"no such field: {name}")
        item = self.dtype.itemsize

        class _View:
            pass
        v = _View()
        v.pitch = int(slab.pitch)
        v.__cuda_array_interface__ = {"shape": self.shape, "typestr": "<f4" if item == 4 else "<f8", "data": (int(ptr), False),
                                      "version": 2, "strides": (int(slab.pitch) * item, item)}
        return v

    def device_field(self, name):
        """Download one field straight from the device (shape (X, Y))."""
        out = np.empty(self.shape, self.dtype)
        _lib.check(self._lib, self._lib.fluidb200_download(self._h, _FIELD_NAMES.index(name), out.ctypes.data))
        return out

    # -- getters (host state, like the reference) ----------------------------------------------
    def _get(self, name, output):
        a = self._hostf(name)
        if output is None:
            return a.reshape(-1).copy()
        np.asarray(output).reshape(-1)[: a.size] = a.reshape(-1)
        return 0

    def getUValues(self, output=None):
        return self._get("u", output)

    def getVValues(self, output=None):
        return self._get("v", output)

    def getPValues(self, output=None):
        return self._get("p", output)

    def getUNextValues(self, output=None):
        return self._get("u_next", output)

    def getRightHand(self, output=None):
        return self._get("rhs", output)

    def getPNext(self, output=None):
        return self._get("p_next", output)

    # -- raw dumps, as the reference's drivers write them -------------------------------------
    def write_dat(self, prefix="Cuda", directory="."):
        """Write ``{prefix}UValues.float.dat`` / ``VValues`` / ``PValues``: raw little-endian values in
        getter order (``i = x*Y + y``), the files ``fluid_cpp/cuda_main.cu:69-80`` (prefix ``Cuda``) and
        ``fluid_cpp/main.cpp:68-78`` (prefix ``Cpp``) produce and ``figures/plot_c_time_data.ipynb``
        reads. Host state is written, so call ``migrateDeviceToHost()`` first, like the getters.
        The file name keeps the reference's ``.float.dat`` suffix for fp32 and uses ``.double.dat``
        for fp64. Returns the three paths."""
        import os
        suffix = "float" if self.dtype == np.float32 else "double"
        paths = []
        for letter, name in (("U", "u"), ("V", "v"), ("P", "p")):
            path = os.path.join(directory, f"{prefix}{letter}Values.{suffix}.dat")
            self._hostf(name).astype(self.dtype.newbyteorder("<"), copy=False).tofile(path)
            paths.append(path)
        return paths

    @staticmethod
    def read_dat(path, num_x_bins, num_y_bins, dtype=np.float32):
        """Load one such file as an ``(X, Y)`` array (``.T`` gives the ``[y, x]`` orientation of the
        reference's NumPy solver and of its notebook, cell 6)."""
        return np.fromfile(path, dtype=np.dtype(dtype).newbyteorder("<")).reshape(num_x_bins, num_y_bins)


class MultiGpuNavierStokes:
    """The x-slab multi-GPU solver of the C ABI (``fluidb200_multi_*``): one process, one slab per entry of
    ``devices`` (an entry may repeat). Cavity boundaries; same public fields and call sequence as
    ``ParallelNavierStokes``; results are bit-identical to it."""

    def __init__(self, box_dim_x, box_dim_y, domain_size_x=1.0, domain_size_y=1.0, dtype=np.float32, devices=(0,), _libpath=None):
        self.dtype = np.dtype(dtype)
        self._lib = _lib.load(_libpath)
        self.box_dimension_x, self.box_dimension_y = int(box_dim_x), int(box_dim_y)
        self.time_step, self.kinematic_viscosity, self.density = 0.001, 0.1, 1.0
        self.num_poisson_iterations, self.num_iterations = 50, 1000
        T = self.dtype.type
        devs = (C.c_int * len(devices))(*[int(d) for d in devices])
        self._h = C.c_void_p()
        _lib.check(self._lib, self._lib.fluidb200_multi_create(
            C.byref(self._h), int(self.dtype == np.float64), self.box_dimension_x, self.box_dimension_y,
            float(T(domain_size_x)), float(T(domain_size_y)), len(devices), devs))

    @property
    def shape(self):
        return (self.box_dimension_x, self.box_dimension_y)

    def upload_cavity(self, lid_velocity=1.0):
        _lib.check(self._lib, self._lib.fluidb200_multi_upload_cavity(self._h, float(lid_velocity)))

    def upload(self, **fields):
        """Dense (X, Y) host arrays by field name (u, v, p, u_next, v_next, p_next, rhs)."""
        ptrs = (C.c_void_p * _lib.NUM_FIELDS)()
        keep = []
        for i, n in enumerate(_FIELD_NAMES):
            if n in fields:
                a = np.ascontiguousarray(fields[n], dtype=self.dtype)
                keep.append(a)
                ptrs[i] = a.ctypes.data
        _lib.check(self._lib, self._lib.fluidb200_multi_upload(self._h, ptrs))

    def solve(self):
        T = self.dtype.type
        _lib.check(self._lib, self._lib.fluidb200_multi_set_params(
            self._h, float(T(self.time_step)), float(T(self.kinematic_viscosity)), float(T(self.density)),
            int(self.num_poisson_iterations), int(self.num_iterations)))
        _lib.check(self._lib, self._lib.fluidb200_multi_solve(self._h))

    def synchronize(self):
        _lib.check(self._lib, self._lib.fluidb200_multi_synchronize(self._h))

    def device_field(self, name):
        out = np.empty(self.shape, self.dtype)
        _lib.check(self._lib, self._lib.fluidb200_multi_download(self._h, _FIELD_NAMES.index(name), out.ctypes.data))
        return out

    def launch_count(self):
        return int(self._lib.fluidb200_multi_launch_count(self._h))

    def close(self):
        if getattr(self, "_h", None) is not None and self._h.value:
            self._lib.fluidb200_multi_destroy(self._h)
            self._h = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


class TiledNavierStokes(ParallelNavierStokes):
    """Drop-in for ``TiledNavierStokes<T>`` (``fluid_cpp/cuda/TiledNavierStokes.cuh:10-23``).

    The reference's tiled kernels read the wrong left neighbour (``TiledNavierStokes.cu:35,109``)
    and so do not reproduce the serial solver; the name and signatures are kept, the behaviour
    is the correct (serial-equal) one. Here "tiled" selects the deepest temporal blocking."""

    def __init__(self, *a, **k):
        super().__init__(*a, **k)
        self.set_sweeps_per_launch(6)


def solve_navier_stokes(num_elements=41, domain_size=1.0, num_iterations=1000, time_step=0.001,
                        kinematic_viscosity=0.1, density=1., top_velocity=1., num_poisson_iterations=50,
                        stability_safety_factor=0.5, dtype=np.float64, _libpath=None):
    """Signature of the reference's only Python entry point (``python/navier_stokes.py:24-26``),
    returning ``(X, Y, u, v, p)`` with ``[y, x]`` arrays like it does (``:99``), computed by the
    CUDA solver with the C++ solver's semantics. Differences from the NumPy script that follow
    from using the C++ semantics: the lid is already moving at t=0, so
    ``num_iterations = S`` here corresponds to ``S + 1`` there, and the four corner cells differ.
    """
    n = int(num_elements)
    s = ParallelNavierStokes(n, n, domain_size, domain_size, dtype=dtype, _libpath=_libpath)
    s.time_step, s.kinematic_viscosity, s.density = time_step, kinematic_viscosity, density
    s.num_iterations, s.num_poisson_iterations = int(num_iterations), int(num_poisson_iterations)
    s.stability_safety_factor = stability_safety_factor
    s.set_cavity(top_velocity)
    s.migrateSolve()
    x = np.linspace(0.0, domain_size, n)
    X, Y = np.meshgrid(x, x)
    out = tuple(np.ascontiguousarray(s._hostf(f).T) for f in ("u", "v", "p"))
    s.close()
    return (X, Y) + out
