import os
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

GOLDEN = os.path.join(ROOT, "tests", "golden")


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box with -m gpu)")


@pytest.fixture(scope="session")
def oracle():
    from oracle import Oracle
    return Oracle()


@pytest.fixture(scope="session")
def dtw_golden():
    return np.load(os.path.join(GOLDEN, "dtw_cases.npz"))


@pytest.fixture(scope="session")
def event_golden():
    return np.load(os.path.join(GOLDEN, "event_cases.npz"))


@pytest.fixture(scope="session")
def ctx():
    """One CUDA context for the whole GPU session; fails loudly if the library or the GPU is missing."""
    import rawalign_b200 as ra
    c = ra.Context(0)
    yield c
    c.close()


def unpack(flat, off, k):
    return flat[int(off[k]):int(off[k + 1])]


def bits(x):
    return np.ascontiguousarray(x, np.float32).view(np.uint32)
