import os
import sys
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box with -m gpu)")


@pytest.fixture(scope="session")
def orc():
    from oracle import orc as o
    o.build()
    return o


@pytest.fixture(scope="session")
def gs():
    """the CUDA library through its C ABI; fails loudly if it is missing"""
    from gaussiansplatting_b200 import capi
    capi.lib()
    return capi


@pytest.fixture(scope="session")
def ctx(gs):
    c = gs.Context(0)
    yield c
    c.close()
