import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA (sm_100a) device; run with -m gpu on the B200 box")


@pytest.fixture(scope="session")
def built_lib():
    """Build librvcdec.so once per session if it is not there (nvcc cross-compiles without a GPU)."""
    import __graft_entry__ as ge
    from rvc_b200 import _lib

    if not os.path.exists(_lib.LIB_PATH):
        ge.build()
    return _lib.lib()
