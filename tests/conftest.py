import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a B200 (sm_100a) GPU; run with -m gpu")


@pytest.fixture(scope="session")
def oracle_port():
    import oracle
    return oracle.load_oracle()


@pytest.fixture(scope="session")
def oracle_ref():
    """The reference's own functions (oracle/_ref), or None where it was not built/shipped."""
    import oracle
    return oracle.load_ref()


@pytest.fixture(scope="session")
def ctx():
    import visnav_b200
    c = visnav_b200.Context(0)
    yield c
    c.close()
