import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box)")


@pytest.fixture(scope="session")
def golden():
    import numpy as np
    return np.load(os.path.join(ROOT, "tests", "golden", "reference_outputs.npz"))


@pytest.fixture(scope="session")
def fdsm():
    import fdsm_b200
    return fdsm_b200


@pytest.fixture(autouse=True)
def _seed_everything():
    """Every test starts from the same torch RNG state (CPU and CUDA): no test outcome depends on test order."""
    import torch
    torch.manual_seed(20260101)
    yield
