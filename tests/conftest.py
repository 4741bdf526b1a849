import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box)")


@pytest.fixture(scope="session")
def ctx():
    """One StitchContext for the whole GPU session; fails loudly if the CUDA library is missing."""
    from imagestitching_b200 import StitchContext
    c = StitchContext(0)
    yield c
    c.close()
