import importlib
import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (ROOT, os.path.join(ROOT, "tests")):
    if p not in sys.path:
        sys.path.insert(0, p)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA sm_100 device (run on the B200 box)")


@pytest.fixture(scope="session")
def pkg():
    """The product package; builds liblz4e_b200.so when nvcc is here and it is missing."""
    p = importlib.import_module("lz4-sgori_b200")
    if not os.path.exists(p.LIB_PATH):
        p.build()
    p.lib()
    return p


@pytest.fixture(scope="session")
def ref_lib():
    from oracle import oracle
    return oracle.load("auto")


@pytest.fixture(scope="session")
def port_lib():
    from oracle import oracle
    return oracle.load("port")


@pytest.fixture(scope="session")
def ctx(pkg):
    c = pkg.Context(0)  # raises loudly without a GPU: there is no CPU path
    yield c
    c.close()
