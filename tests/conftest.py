import os
import sys

import pytest

# the virtual-world tests run several ranks' streams on one GPU, some of them parked in a
# wait-for-flag kernel: give every stream its own hardware queue so that none of them ends
# up queued behind another rank's wait (must be set before the CUDA context exists)
os.environ.setdefault("CUDA_DEVICE_MAX_CONNECTIONS", "32")

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box)")


def pytest_collection_modifyitems(config, items):
    try:
        import torch

        has_gpu = torch.cuda.is_available()
    except Exception:
        has_gpu = False
    if has_gpu:
        return
    skip = pytest.mark.skip(reason="no CUDA device")
    for item in items:
        if "gpu" in item.keywords:
            item.add_marker(skip)


@pytest.fixture(params=[pytest.param("host-double"), pytest.param("b200", marks=pytest.mark.gpu)])
def engine(request):
    """Runs a test body against (a) the host-memory test double of the C-ABI -- host
    logic only, no GPU needed -- and (b) the real CUDA library on a B200."""
    from cluster_anywhere_b200 import _cabi, device

    if request.param == "host-double":
        from _fake_cabi import FakeLib

        _cabi._install_test_double(FakeLib())
        device.reset_default_runtime()
        yield request.param
        _cabi._install_test_double(None)
        device.reset_default_runtime()
    else:
        _cabi._install_test_double(None)
        device.reset_default_runtime()
        yield request.param
