import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a B200 (run with -m gpu on the GPU box)")


@pytest.fixture(scope="session")
def cuda_lib():
    """Build (if stale) and load the CUDA library; GPU tests fail loudly when it is missing."""
    import torch
    from taboo_brittleness_b200 import _lib, build
    assert torch.cuda.is_available(), "GPU test collected on a machine without CUDA"
    if os.path.exists("/usr/local/cuda/bin/nvcc"):
        build.build()
    lib = _lib.load()
    from taboo_brittleness_b200 import ops
    ops.device_check()
    return lib
