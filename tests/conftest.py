import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box with -m gpu)")


@pytest.fixture(scope="session")
def orc():
    from oracle import oracle
    oracle.lib()
    return oracle


@pytest.fixture(scope="session")
def cuda_api():
    """The product's host mirror over the C ABI; fails loudly when the .so is missing."""
    import torch
    assert torch.cuda.is_available(), "gpu test without a CUDA device"
    from rtx_nerf_b200 import _lib, api
    _lib.load()
    return api
