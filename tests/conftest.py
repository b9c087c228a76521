import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
PKG = os.path.join(ROOT, "medical-image-analysis_b200")
for p in (ROOT, PKG):
    if p not in sys.path:
        sys.path.insert(0, p)

GOLDEN = os.path.join(ROOT, "tests", "golden")


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a B200 (run with -m gpu on the GPU box)")
    config.addinivalue_line("markers", "reference: needs /root/reference (container only)")


@pytest.fixture(scope="session", autouse=True)
def _built_library():
    """The C-ABI library must exist for every test session (CPU tests check its
    symbols; GPU tests call through it)."""
    from b200al import build
    build.build()


@pytest.fixture(scope="session")
def golden():
    import numpy as np
    return {name: np.load(os.path.join(GOLDEN, f"{name}.npz")) for name in ("model", "pool_K2", "pool_K3", "mc_K3")}
