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


@pytest.fixture(autouse=True)
def _poison_free_device_memory(request):
    """Before every GPU test the caching allocator's free memory is filled with NaN: a kernel that reads a workspace it
    was supposed to initialise (torch.empty blocks recycled from an earlier test) then fails in every test order instead
    of in some (that is how a timing-dependent result of the elastic kernels showed up in round 2).
    FWI_TEST_POISON_MB sets the size (default 3072, 0 = off)."""
    if request.node.get_closest_marker("gpu") is None:
        yield
        return
    import torch
    mb = int(os.environ.get("FWI_TEST_POISON_MB", "3072"))
    if mb > 0 and torch.cuda.is_available():
        for dev in range(min(torch.cuda.device_count(), 2)):
            try:
                x = torch.full((mb * 262144,), float("nan"), dtype=torch.float32, device=f"cuda:{dev}")
                del x
            except RuntimeError:          # not enough free memory left by an earlier test: nothing to poison
                pass
    yield
