import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: test needs a CUDA device (run on the B200 box)")


@pytest.fixture(scope="session", autouse=True)
def _built_library():
    """Make sure librendercompare.so and the oracle's native pieces exist (nvcc/gcc
    cross-compile without a GPU; on the GPU box the prebuilt files travel with the snapshot)."""
    import __graft_entry__ as entry

    entry.build()
    yield


def has_cuda():
    try:
        import torch

        return torch.cuda.is_available()
    except Exception:
        return False


@pytest.fixture(scope="session")
def cuda_device():
    if not has_cuda():
        pytest.skip("no CUDA device")
    import torch

    return torch.device("cuda:0")
