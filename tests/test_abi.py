"""CPU-only: the C-ABI library loads and exports every symbol include/fluidb200.h declares.
No compute call is made (there is no GPU here)."""
import ctypes
import os
import re

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
HEADER = os.path.join(ROOT, "include", "fluidb200.h")


def declared_symbols():
    src = open(HEADER).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(fluidb200_[a-z0-9_]+)\s*\(", src)))


def test_header_declares_the_expected_surface():
    syms = declared_symbols()
    for must in ("fluidb200_predictor", "fluidb200_poisson_sweeps", "fluidb200_corrector",
                 "fluidb200_create", "fluidb200_solve", "fluidb200_upload", "fluidb200_download"):
        assert must in syms


def test_library_exports_every_declared_symbol():
    import __graft_entry__ as g
    lib_path = g.build_cuda()
    lib = ctypes.CDLL(lib_path)
    missing = [s for s in declared_symbols() if not hasattr(lib, s)]
    assert not missing, f"declared in include/fluidb200.h but not exported: {missing}"


def test_binding_covers_the_header():
    from fluid_simulation_b200 import _lib
    assert sorted(_lib.EXPORTS) == declared_symbols()


def test_no_gpu_calls_fail_loudly_not_silently():
    """Without a device, creating a solver must raise (no CPU fallback)."""
    import numpy as np
    import torch
    if torch.cuda.is_available():
        pytest.skip("a GPU is present")
    from fluid_simulation_b200 import FluidB200Error, ParallelNavierStokes
    with pytest.raises(FluidB200Error):
        ParallelNavierStokes(41, 41, 1.0, 1.0, dtype=np.float32)


def test_product_package_never_imports_oracle():
    pkg = os.path.join(ROOT, "fluid_simulation_b200")
    for dirpath, _, files in os.walk(pkg):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".cpp", ".h")):
                text = open(os.path.join(dirpath, f)).read()
                assert "import oracle" not in text and "from oracle" not in text, f
                assert "liboracle" not in text and "libfluidb200_emu" not in text, f
