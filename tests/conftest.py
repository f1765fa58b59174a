import os
import sys

import pytest

# caller-side setting of the library's stream count: before the process creates its CUDA context
os.environ.setdefault("CUDA_DEVICE_MAX_CONNECTIONS", "32")
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a real B200 (run with -m gpu on the GPU box)")


@pytest.fixture(scope="session")
def oracle():
    from tests import oracle_lib
    return oracle_lib.load()


@pytest.fixture(scope="session")
def ctx():
    from madpose_b200 import capi
    c = capi.Context(0)  # raises loudly if the CUDA library or the device is missing
    yield c
    c.close()
