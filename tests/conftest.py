import os
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "oracle"))      # the oracle is test infrastructure
GOLDEN = os.path.join(ROOT, "tests", "golden")


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box)")


@pytest.fixture(scope="session", autouse=True)
def _libofx_built():
    """The C-ABI library is built in-tree before any test needs it (a no-op when it is up to date)."""
    from occlusionfusion_b200 import build
    build.build()


@pytest.fixture(scope="session")
def oracle():
    import oracle as O
    O.build()
    return O


@pytest.fixture(scope="session", params=["ref_k4", "ref_k8", "ref_m3"])
def golden(request):
    return request.param, np.load(os.path.join(GOLDEN, request.param + ".npz"))


def load_golden(name):
    return np.load(os.path.join(GOLDEN, name + ".npz"))
