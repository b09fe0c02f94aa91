import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "raytracing-gpu_b200"))
sys.path.insert(0, os.path.join(ROOT, "tests"))
sys.path.insert(0, ROOT)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a B200 (run with -m gpu on the GPU box)")
    config.addinivalue_line("markers", "slow: long CPU test (still part of the default CPU suite unless deselected)")


def pytest_collection_modifyitems(config, items):
    # GPU tests must fail loudly on a GPU box if the extension is missing; on a box
    # without a GPU they are simply not selected (-m "not gpu").
    pass


@pytest.fixture(scope="session")
def rtb():
    import pyrtb
    pyrtb.lib()
    return pyrtb


@pytest.fixture(scope="session")
def gpu(rtb):
    n = rtb.device_count()
    assert n > 0, "no CUDA device: GPU tests need a B200 (there is no CPU fallback)"
    return rtb
