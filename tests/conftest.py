import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box)")


@pytest.fixture(scope="session")
def built():
    """Both native pieces compiled (nvcc cross-compiles without a GPU)."""
    import __graft_entry__ as g

    g.build()
    return True


@pytest.fixture(scope="session")
def skidpad():
    from ihm2_b200.track import load_packaged_track

    return load_packaged_track("skidpad")


@pytest.fixture(scope="session")
def fsds():
    from ihm2_b200.track import load_packaged_track

    return load_packaged_track("fsds_default")


@pytest.fixture(scope="session")
def torch_cuda(built):
    import torch
    assert torch.cuda.is_available(), "these tests need the B200"
    return torch
