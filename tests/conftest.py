import os
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (ROOT, os.path.join(ROOT, "oracle")):
    if p not in sys.path:
        sys.path.insert(0, p)

GOLDEN = os.path.join(ROOT, "tests", "golden")


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box with -m gpu)")


@pytest.fixture(scope="session")
def lib():
    """The built C-ABI library (builds it once if the .so is missing)."""
    from rti_b200 import _capi
    if not os.path.isfile(_capi.LIB_PATH):
        import __graft_entry__
        __graft_entry__.build()
    return _capi.load()


@pytest.fixture(scope="session")
def ref_small():
    return dict(np.load(os.path.join(GOLDEN, "ref_small.npz")))


@pytest.fixture(scope="session")
def ref_rbf_n96():
    return dict(np.load(os.path.join(GOLDEN, "ref_rbf_n96.npz")))


@pytest.fixture(scope="session")
def relight_golden():
    return dict(np.load(os.path.join(GOLDEN, "relight.npz")))


@pytest.fixture(scope="session")
def relight_rows_golden():
    return dict(np.load(os.path.join(GOLDEN, "relight_rows.npz")))
