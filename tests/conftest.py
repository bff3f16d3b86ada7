import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

GOLDEN = os.path.join(ROOT, "tests", "golden")
GOLDEN_NAMES = ["case6", "case5", "galewsky", "dcmip31", "dcmip21", "dcmip31_n4", "dcmip31_n5", "dcmip31_n6", "dcmip21_n5"]
# further reference configurations pinned on the CPU only (oracle vs the reference's output); not yet part of the GPU list
EXTRA_ORACLE_PINS = ["case2"]


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box)")


@pytest.fixture(scope="session")
def golden():
    """name -> (problem, extras) loaded lazily from tests/golden/*.npz."""
    from wxfactory_b200.problem import load_problem

    cache = {}

    def get(name):
        if name not in cache:
            cache[name] = load_problem(os.path.join(GOLDEN, name + ".npz"))
        return cache[name]

    return get
