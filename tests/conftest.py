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
    """Build (if stale) and return the path of libdlwp_b200.so."""
    import __graft_entry__ as entry
    entry.build()
    from dlwp_benchmark_b200 import native
    return native.LIB_PATH
