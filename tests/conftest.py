import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (B200)")


@pytest.fixture(scope="session")
def cuda_lib():
    """Loads the product library; GPU tests must go through it (no fallback)."""
    import torch
    import physicsbasedfwi2_b200 as pkg
    if not torch.cuda.is_available():
        pytest.skip("no CUDA device")
    L = pkg.lib()
    assert L.fwi_device_ok(0) == 0, L.fwi_last_error()
    return L
