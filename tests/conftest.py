import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (ROOT, os.path.join(ROOT, "tests")):
    if p not in sys.path:
        sys.path.insert(0, p)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box with -m gpu)")


@pytest.fixture(scope="session")
def built():
    """The in-tree libraries; build them if the tree is fresh (no GPU needed to build)."""
    import __graft_entry__ as g

    from motionplanning_b200 import _capi as capi

    if not os.path.exists(capi.LIB_PATH) or not os.path.exists(os.path.join(ROOT, "oracle", "libmppi_oracle.so")):
        g.build()
    return capi.load()
