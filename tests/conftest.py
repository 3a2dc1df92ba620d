import os
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box with -m gpu)")


@pytest.fixture(scope="session")
def datasets():
    """The reference's load_data() series (see tests/golden/make_datasets.py)."""
    d = np.load(os.path.join(ROOT, "pyhgf_b200", "data", "datasets.npz"))
    return {"u": d["binary_u"], "y": d["binary_y"], "x": d["continuous"]}
