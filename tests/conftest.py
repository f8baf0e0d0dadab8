import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
PKG_DIR = os.path.join(ROOT, "02689-advancednumericalalgorithmp3_b200")
for p in (ROOT, PKG_DIR):
    if p not in sys.path:
        sys.path.insert(0, p)

GOLDEN = os.path.join(ROOT, "tests", "golden")


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box)")


@pytest.fixture(scope="session")
def golden():
    import numpy as np

    class G:
        def __getattr__(self, name):
            return np.load(os.path.join(GOLDEN, f"{name}.npz"))

    return G()


@pytest.fixture(scope="session")
def built_lib():
    """Build (if stale) and load the CUDA library; every test that touches it goes through here."""
    import __graft_entry__ as ge

    ge.build()
    import ldc_b200

    return ldc_b200
