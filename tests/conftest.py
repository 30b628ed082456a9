import os
import subprocess
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box with -m gpu)")


@pytest.fixture(scope="session")
def oracle():
    """The CPU oracle (oracle/libgt_oracle.so), built on demand.  Checker only."""
    import oracle_lib
    return oracle_lib.load()


@pytest.fixture(scope="session")
def gt():
    """The product library through its C-ABI (ctypes)."""
    import _pkg
    return _pkg.load_package()
