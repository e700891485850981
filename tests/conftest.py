import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (ROOT, os.path.join(ROOT, "tests"), os.path.join(ROOT, "tests", "golden")):
    if p not in sys.path:
        sys.path.insert(0, p)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box with -m gpu)")


@pytest.fixture(scope="session")
def c1():
    """BASELINE config C1: matrix_export.py inputs as the reference loads them + its golden outputs."""
    import numpy as np
    from make_golden import c1_inputs
    W, H, X = c1_inputs()
    g = np.load(os.path.join(ROOT, "tests", "golden", "golden_c1.npz"))
    return dict(W=W, H=H, X=X, Wtest=g["Wtest"], Htest=g["Htest"])
