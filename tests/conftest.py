import os
import sys

import pytest

os.environ.setdefault("CUDA_DEVICE_MAX_CONNECTIONS", "32")

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box with -m gpu)")


@pytest.fixture(scope="session", autouse=True)
def _built():
    """Everything native is built once per session (the .so files are git-ignored)."""
    import __graft_entry__ as ge
    ge.build()
    yield


@pytest.fixture(scope="session")
def ctx():
    """One device context for the whole GPU session.  Fails loudly (no fallback) if the library or GPU is missing."""
    from minimap2_chaindp_b200 import api
    c = api.ChainContext(device=0, profile=False)
    yield c
    c.close()
