import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box)")
    config.addinivalue_line("markers", "fast_mode: the library's default blur arithmetic (test_gpu_sharded.py)")


def _have_gpu():
    try:
        from sift3d_b200 import cuda_device_count
        return cuda_device_count() > 0
    except Exception:
        return False


@pytest.fixture(scope="session")
def gpu_available():
    return _have_gpu()


def pytest_collection_modifyitems(config, items):
    # `-m gpu` on a box without a GPU must fail loudly rather than skip: the product has
    # no CPU path. Without `-m gpu`, GPU tests are simply deselected by the marker
    # expression the driver passes (-m "not gpu").
    pass
