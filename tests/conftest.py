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
    path = os.path.join(ROOT, "tests", "golden", "reference_outputs.npz")
    return dict(np.load(path))


@pytest.fixture(scope="session")
def golden_train():
    import numpy as np
    path = os.path.join(ROOT, "tests", "golden", "reference_train.npz")
    return dict(np.load(path))


@pytest.fixture(scope="session")
def golden_encode():
    import numpy as np
    path = os.path.join(ROOT, "tests", "golden", "reference_vae_encode.npz")
    return dict(np.load(path))
