import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box)")


@pytest.fixture(scope="session")
def capi():
    from pir_b200 import _capi
    return _capi


@pytest.fixture(scope="session")
def gpu_lib(capi):
    L = capi.lib()
    if L.pir_device_count() <= 0:
        pytest.fail("GPU test selected but no CUDA device is visible; the product has no CPU path")
    return L
