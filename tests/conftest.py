import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box)")


@pytest.fixture(scope="session")
def oracle():
    """CPU oracle (plain-C restatement).  Test infrastructure only."""
    from oracle import oracle as O
    O.build(ref=False)
    return O.Oracle()


@pytest.fixture(scope="session")
def oracle_mod():
    """The oracle's Python module (its ctypes structures).  Test infrastructure only."""
    from oracle import oracle as O
    return O


@pytest.fixture(scope="session")
def h3d():
    """The product: libh3d.so through its C-ABI.  Fails loudly if it is not built."""
    from pointcloud_segmentation_b200 import capi
    capi.lib()
    return capi
