import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (ROOT, os.path.join(ROOT, "tests")):
    if p not in sys.path:
        sys.path.insert(0, p)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box)")


@pytest.fixture(scope="session")
def oracle():
    from oracle_backend import OracleBackend
    return OracleBackend(threads=0)


@pytest.fixture(scope="session")
def cuda():
    """The product backend. Fails loudly (no fallback) if the library or the device is missing."""
    import rtw_b200
    be = rtw_b200.get_default_backend()
    assert be.device_count() > 0, "no CUDA device visible"
    return be
