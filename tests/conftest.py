import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
# The library sends a Jacobi launch through its split path (lean fast kernel + general kernel) only when
# the grid is large enough to profit; the tests want that path exercised on every grid that has
# interior strips and chunks at all.
os.environ.setdefault("FLUIDB200_SPLIT_MIN_TASKS", "1")
# Small cavity grids would all take the resident solver (one cooperative launch per solve(), tests/test_resident.py);
# the rest of the suite is about the streaming kernels, so it keeps them on every grid size.
os.environ.setdefault("FLUIDB200_RESIDENT", "0")
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box)")


@pytest.fixture(scope="session")
def golden():
    import json
    with open(os.path.join(ROOT, "tests", "golden", "golden.json")) as f:
        return json.load(f)
