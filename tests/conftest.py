import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box)")


@pytest.fixture(scope="session")
def wav():
    """The reference's bundled speech clip, committed as a fixture by
    tests/golden/make_golden.py (the GPU box has no /root/reference)."""
    import numpy as np
    p = os.path.join(ROOT, "tests", "golden", "f2bjrop1.npz")
    return np.load(p)["samples"]
