import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box)")


@pytest.fixture(scope="session", autouse=True)
def _built():
    """Build the CUDA library (nvcc cross-compiles without a GPU) and the oracle once per session."""
    from fqtool_b200 import build
    build.build()
    build.build_host()
    from oracle import pyoracle
    pyoracle.build()
