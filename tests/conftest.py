import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box with -m gpu)")


@pytest.fixture(scope="session")
def cuda_dev():
    import torch
    if not torch.cuda.is_available():
        pytest.skip("no CUDA device")
    return torch.device("cuda:0")


@pytest.fixture
def torch_backend_default():
    """Tests written as "reference = the stock torch layers, candidate = explicit conv_backend('tc')" run with the
    TEST-ONLY strict backend as the ambient default; the product default is 'tc' (adaptpoint_b200/tcconv.py)."""
    from adaptpoint_b200 import tcconv
    prev = tcconv.get_conv_backend()
    tcconv.set_conv_backend("torch")
    try:
        yield
    finally:
        tcconv.set_conv_backend(prev)
