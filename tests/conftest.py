import os
import subprocess
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box)")


@pytest.fixture(scope="session", autouse=True)
def _built():
    """Engine library (nvcc cross-compiles without a GPU) and the oracle are built once per session."""
    import __graft_entry__ as g
    g.build()
    yield


def has_gpu():
    try:
        from bramble_b200 import capi
        return capi.lib().brb_device_count() > 0
    except Exception:
        return False
