import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box)")


@pytest.fixture(scope="session")
def built_lib():
    """libvs_tracker.so must exist (built by __graft_entry__.build()); never silently skipped."""
    from mobile_slam_b200 import _abi
    if not os.path.exists(_abi.LIB_PATH):
        import __graft_entry__ as g
        g.build()
    return _abi.load()
