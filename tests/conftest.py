import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (ROOT, os.path.join(ROOT, "tests", "golden")):
    if p not in sys.path:
        sys.path.insert(0, p)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (B200); run with -m gpu")


@pytest.fixture(scope="session")
def golden():
    import numpy as np
    return np.load(os.path.join(ROOT, "tests", "golden", "golden_small.npz"))


@pytest.fixture(scope="session")
def golden_full():
    import json
    with open(os.path.join(ROOT, "tests", "golden", "golden_full.json")) as f:
        return json.load(f)


@pytest.fixture(scope="session")
def sv():
    """The product package with its CUDA library built and loaded (fails loudly if it cannot)."""
    import torch
    assert torch.cuda.is_available(), "gpu tests need a CUDA device"
    from stereo_vision_b200 import build as _b
    _b.build()
    import stereo_vision_b200 as m
    from stereo_vision_b200 import _lib
    _lib.load()
    return m
