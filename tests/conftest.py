import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box with -m gpu)")


@pytest.fixture(scope="session")
def oracle():
    """CPU checker (plain-C restatement, bit-identical to the reference functors)."""
    from oracle.oracle import Oracle
    return Oracle("port")


@pytest.fixture(scope="session")
def oracle_ref():
    from oracle.oracle import Oracle, have_ref
    if not have_ref():
        pytest.skip("oracle/_ref not built (needs /root/reference)")
    return Oracle("ref")
