import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run with -m gpu on the B200 box)")


@pytest.fixture(scope="session")
def sl():
    """The product package, with a device selected.  Fails (not skips) when the CUDA library is
    missing: GPU tests must never pass on a fallback."""
    import sparse_linear_b200 as pkg
    pkg._capi.lib()
    assert pkg.device_count() > 0, "no CUDA device visible"
    pkg.set_device(0)
    return pkg
