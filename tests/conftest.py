import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)
os.environ.setdefault("HF_HUB_OFFLINE", "1")


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box with -m gpu)")


@pytest.fixture(scope="session")
def golden_dir():
    return os.path.join(ROOT, "tests", "golden")


@pytest.fixture(scope="session")
def native_lib():
    """libddp_b200.so, built on demand (nvcc cross-compiles sm_100a without a GPU)."""
    import ddp_b200
    if not os.path.exists(ddp_b200._lib.LIB_PATH):
        ddp_b200._lib.build()
    return ddp_b200._lib.lib()
