import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a B200 (run by `pytest -m gpu` on the GPU box)")


@pytest.fixture(scope="session")
def oracle_port():
    import oracle
    return oracle.Oracle("port")


@pytest.fixture(scope="session")
def oracle_ref():
    import oracle
    if oracle.load_ref() is None:
        pytest.skip("oracle/_ref not built (no /root/reference here)")
    return oracle.Oracle("reference")
