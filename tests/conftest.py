import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box with -m gpu)")


@pytest.fixture(scope="session")
def native_lib():
    """Build (if stale) and load libpathnet_b200.so; loading needs no GPU."""
    from pathnet_gym_b200 import build as _b
    _b.build_library()
    from pathnet_gym_b200 import _native
    return _native.load()
