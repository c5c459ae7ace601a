"""ctypes binding of the C ABI in include/fluidb200.h.

The product library is ``fluid_simulation_b200/libfluidb200.so`` (built in-tree by
``__graft_entry__.build()`` from ``csrc/``). There is NO fallback: if the library is missing
or cannot be loaded this module raises, it never substitutes a CPU implementation.
"""
from __future__ import annotations

import ctypes as C
import os

HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(HERE, "libfluidb200.so")

F32, F64 = 0, 1
FIELD_U, FIELD_V, FIELD_P, FIELD_U_NEXT, FIELD_V_NEXT, FIELD_P_NEXT, FIELD_RHS = range(7)
NUM_FIELDS = 7
EDGE_UNSET, EDGE_CONST, EDGE_ARRAY = 0, 1, 2
EDGE_BOTTOM, EDGE_TOP, EDGE_LEFT, EDGE_RIGHT = 0, 1, 2, 3
MAX_SWEEPS_PER_LAUNCH = 8
NOALLOC_MAX_BYTES = 16 << 20


class Slab(C.Structure):
    """struct fluidb200_slab"""
    _fields_ = [("dtype", C.c_int), ("X", C.c_int), ("Y", C.c_int), ("pitch", C.c_int64),
                ("row0", C.c_int), ("rows", C.c_int), ("hx", C.c_double), ("hy", C.c_double)]


class Edges(C.Structure):
    """struct fluidb200_edges"""
    _fields_ = [("mode", C.c_int * 4), ("cval", C.c_double * 4), ("val", C.c_void_p * 4),
                ("set", C.c_void_p * 4)]


class Mask(C.Structure):
    """struct fluidb200_mask"""
    _fields_ = [("set", C.c_void_p), ("val", C.c_void_p), ("tiles", C.c_void_p)]


class FluidB200Error(RuntimeError):
    def __init__(self, code, msg):
        super().__init__(f"fluidb200 error {code}: {msg}")
        self.code = code


_SIGS = {
    "fluidb200_predictor": (C.c_int, [C.POINTER(Slab)] + [C.c_void_p] * 5 + [C.c_double] * 3 + [C.c_int] * 4 + [C.c_void_p]),
    "fluidb200_poisson_sweeps": (C.c_int, [C.POINTER(Slab), C.c_void_p, C.c_void_p, C.c_void_p, C.POINTER(Edges),
                                           C.POINTER(Mask), C.c_int, C.c_int, C.c_int, C.c_void_p]),
    "fluidb200_corrector": (C.c_int, [C.POINTER(Slab)] + [C.c_void_p] * 5 + [C.POINTER(Edges)] * 2 + [C.POINTER(Mask)] * 2
                            + [C.c_double] * 2 + [C.c_int] * 2 + [C.c_void_p]),
    "fluidb200_poisson_solve_noalloc": (C.c_int, [C.POINTER(Slab), C.c_void_p, C.c_void_p, C.POINTER(Edges), C.c_int, C.c_void_p]),
    "fluidb200_init_cavity": (C.c_int, [C.POINTER(Slab)] + [C.c_void_p] * 7 + [C.c_double, C.c_void_p]),
    "fluidb200_cavity_edges": (C.c_int, [C.c_char, C.c_double, C.POINTER(Edges)]),
    "fluidb200_create": (C.c_int, [C.POINTER(C.c_void_p), C.c_int, C.c_int, C.c_int, C.c_double, C.c_double]),
    "fluidb200_destroy": (C.c_int, [C.c_void_p]),
    "fluidb200_set_params": (C.c_int, [C.c_void_p, C.c_double, C.c_double, C.c_double, C.c_int, C.c_int]),
    "fluidb200_set_domain": (C.c_int, [C.c_void_p, C.c_double, C.c_double]),
    "fluidb200_upload": (C.c_int, [C.c_void_p, C.POINTER(C.c_void_p), C.POINTER(C.c_void_p), C.POINTER(C.c_void_p)]),
    "fluidb200_upload_cavity": (C.c_int, [C.c_void_p, C.c_double]),
    "fluidb200_solve": (C.c_int, [C.c_void_p]),
    "fluidb200_step_predictor": (C.c_int, [C.c_void_p]),
    "fluidb200_step_poisson": (C.c_int, [C.c_void_p, C.c_int]),
    "fluidb200_step_corrector": (C.c_int, [C.c_void_p]),
    "fluidb200_synchronize": (C.c_int, [C.c_void_p]),
    "fluidb200_download": (C.c_int, [C.c_void_p, C.c_int, C.c_void_p]),
    "fluidb200_migrate_solve": (C.c_int, [C.c_void_p, C.POINTER(C.c_void_p), C.POINTER(C.c_void_p)]),
    "fluidb200_get_slab": (C.c_int, [C.c_void_p, C.POINTER(Slab)]),
    "fluidb200_field_ptr": (C.c_void_p, [C.c_void_p, C.c_int]),
    "fluidb200_set_stream": (C.c_int, [C.c_void_p, C.c_void_p]),
    "fluidb200_set_sweeps_per_launch": (C.c_int, [C.c_void_p, C.c_int]),
    "fluidb200_set_use_graph": (C.c_int, [C.c_void_p, C.c_int]),
    "fluidb200_launch_count": (C.c_int64, [C.c_void_p]),
    "fluidb200_set_resident": (C.c_int, [C.c_void_p, C.c_int]),
    "fluidb200_resident_launch_count": (C.c_int64, [C.c_void_p]),
    "fluidb200_resident_profile": (C.c_int, [C.c_void_p, C.POINTER(C.c_uint64)]),
    "fluidb200_device_alloc": (C.c_int, [C.POINTER(C.c_void_p), C.c_size_t]),
    "fluidb200_device_free": (C.c_int, [C.c_void_p]),
    "fluidb200_ipc_export": (C.c_int, [C.c_void_p, C.c_char_p]),
    "fluidb200_ipc_open": (C.c_int, [C.c_char_p, C.POINTER(C.c_void_p)]),
    "fluidb200_ipc_close": (C.c_int, [C.c_void_p]),
    "fluidb200_peer_enable": (C.c_int, [C.c_int]),
    "fluidb200_halo_push": (C.c_int, [C.c_void_p, C.c_void_p, C.c_size_t, C.c_void_p, C.c_uint32, C.c_void_p]),
    "fluidb200_halo_copy": (C.c_int, [C.c_void_p, C.c_void_p, C.c_size_t, C.c_void_p]),
    "fluidb200_halo_wait": (C.c_int, [C.c_void_p, C.c_uint32, C.c_void_p]),
    "fluidb200_multi_create": (C.c_int, [C.POINTER(C.c_void_p), C.c_int, C.c_int, C.c_int, C.c_double, C.c_double, C.c_int, C.POINTER(C.c_int)]),
    "fluidb200_multi_destroy": (C.c_int, [C.c_void_p]),
    "fluidb200_multi_set_params": (C.c_int, [C.c_void_p, C.c_double, C.c_double, C.c_double, C.c_int, C.c_int]),
    "fluidb200_multi_upload_cavity": (C.c_int, [C.c_void_p, C.c_double]),
    "fluidb200_multi_upload": (C.c_int, [C.c_void_p, C.POINTER(C.c_void_p)]),
    "fluidb200_multi_solve": (C.c_int, [C.c_void_p]),
    "fluidb200_multi_synchronize": (C.c_int, [C.c_void_p]),
    "fluidb200_multi_download": (C.c_int, [C.c_void_p, C.c_int, C.c_void_p]),
    "fluidb200_multi_launch_count": (C.c_int64, [C.c_void_p]),
    "fluidb200_multi_num_slabs": (C.c_int, [C.c_void_p]),
    "fluidb200_set_device": (C.c_int, [C.c_int]),
    "fluidb200_last_error": (C.c_char_p, []),
    "fluidb200_report_error": (C.c_int, [C.c_int, C.c_char_p]),
    "fluidb200_version": (C.c_char_p, []),
    "fluidb200_element_length": (C.c_double, [C.c_int, C.c_double, C.c_int]),
}
EXPORTS = tuple(_SIGS)

_lib = None


def load(path: str | None = None):
    """Load (once) and return the C library. ``path`` overrides the in-tree product library;
    only the test-suite uses that, to point the same binding at tests/emu's host build."""
    global _lib
    if _lib is not None and path is None:
        return _lib
    p = path or os.environ.get("FLUIDB200_LIB") or LIB_PATH   # (FLUIDB200_LIB: another build of the same library, for measurements)
    if not os.path.exists(p):
        raise ImportError(
            f"{p} not found: the CUDA library must be built first "
            "(python -c 'import __graft_entry__ as g; g.build()'). There is no CPU fallback.")
    lib = C.CDLL(p)
    for name, (res, args) in _SIGS.items():
        fn = getattr(lib, name)  # AttributeError here = the library does not export the ABI
        fn.restype = res
        fn.argtypes = args
    if path is None:
        _lib = lib
    return lib


def check(lib, rc):
    if rc != 0:
        raise FluidB200Error(rc, lib.fluidb200_last_error().decode())
