import os
import sys

import numpy as np
import pytest

REPO = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if REPO not in sys.path:
    sys.path.insert(0, REPO)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box)")


@pytest.fixture(scope="session")
def golden():
    """Values produced by running the reference itself (tests/golden/make_golden.py)."""
    path = os.path.join(REPO, "tests", "golden", "ref_golden.npz")
    return dict(np.load(path))
