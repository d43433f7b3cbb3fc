import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

GOLDEN = os.path.join(ROOT, "tests", "golden")


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box)")


@pytest.fixture(scope="session")
def golden():
    import numpy as np

    def load(name):
        return dict(np.load(os.path.join(GOLDEN, name + ".npz")))

    return load


AUG_CASES = ["aug_84", "aug_small", "aug_rect", "aug_odd"]
REPLAY_CASES = ["replay_small", "replay_84"]
# tolerance contract: float kernels <= 1e-5 abs on the /255 scale
TOL_NORM = 1e-5
