import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def pytest_configure(config):
    # The shared library is a build product (git-ignored).  In a checkout where it has not been built yet,
    # build it once here (nvcc cross-compiles without a GPU) instead of failing every test that loads it.
    if not os.path.exists(os.path.join(ROOT, "hmc_b200", "libhmcb.so")):
        import __graft_entry__
        __graft_entry__.build()
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box with -m gpu)")
    config.addinivalue_line("markers", "refonly: needs the reference tree at /root/reference (build container only)")


@pytest.fixture(scope="session")
def golden_dir():
    return os.path.join(ROOT, "tests", "golden")
