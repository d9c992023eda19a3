import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
PKG = os.path.join(ROOT, "gp-quadrature_b200")
for p in (PKG, ROOT):
    if p not in sys.path:
        sys.path.insert(0, p)
GOLDEN = os.path.join(ROOT, "tests", "golden")


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box)")


def oracle():
    """Import the CPU oracle (tests only)."""
    op = os.path.join(ROOT, "oracle")
    if op not in sys.path:
        sys.path.append(op)
    import efgp_oracle
    return efgp_oracle


@pytest.fixture(scope="session")
def golden_dir():
    return GOLDEN
