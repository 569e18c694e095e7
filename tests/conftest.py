import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box)")


@pytest.fixture(scope="session")
def built_library():
    """libnsb200.so, built in-tree by __graft_entry__.build()."""
    import __graft_entry__
    __graft_entry__.build()
    import nsb200
    return nsb200._lib.load()
