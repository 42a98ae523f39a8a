import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)
GOLDEN = os.path.join(ROOT, "tests", "golden")


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box with -m gpu)")


@pytest.fixture(scope="session")
def built():
    """Build (or reuse) csrc/libgradavg.so and the C oracle; returns the loaded ctypes library."""
    import __graft_entry__ as g
    g.build()
    import pytorch_distributed_example_b200 as ga
    return ga.load_library()


@pytest.fixture(scope="session")
def golden_dir():
    return GOLDEN
