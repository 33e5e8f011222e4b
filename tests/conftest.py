import os
import sys

import numpy as np
import pytest
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)
GOLDEN = os.path.join(ROOT, "tests", "golden")


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box)")


def pytest_collection_modifyitems(config, items):
    if torch.cuda.is_available():
        return
    skip = pytest.mark.skip(reason="no CUDA device")
    for item in items:
        if "gpu" in item.keywords:
            item.add_marker(skip)


def load_golden(name):
    return np.load(os.path.join(GOLDEN, name))


@pytest.fixture(scope="session")
def nerf_pth_state_dict():
    """examples/nerf.pth rebuilt from the committed fixture (key order kept)."""
    z = load_golden("nerf_pth.npz")
    from collections import OrderedDict
    return OrderedDict((str(k), torch.from_numpy(z[str(k)])) for k in z["__keys__"])


@pytest.fixture(scope="session")
def g1_oracle(nerf_pth_state_dict):
    from oracle import g1
    m = g1.G1NeRF(normalize_position=6.0)
    m.load_state_dict(nerf_pth_state_dict, strict=True)
    return m.eval()


@pytest.fixture(scope="session")
def g1_ref():
    return load_golden("g1_ref.npz")


@pytest.fixture(scope="session")
def head_ref():
    return load_golden("head_ref.npz")
