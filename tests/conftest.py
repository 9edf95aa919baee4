import importlib
import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (ROOT, os.path.join(ROOT, "oracle")):
    if p not in sys.path:
        sys.path.insert(0, p)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box)")


@pytest.fixture(scope="session")
def zk():
    return importlib.import_module("node-zkstream_b200")


@pytest.fixture(scope="session")
def synth():
    return importlib.import_module("node-zkstream_b200.synth")


@pytest.fixture(scope="session")
def oracle():
    import zkoracle
    zkoracle.load()
    return zkoracle


@pytest.fixture(scope="session")
def ctx(zk):
    """One codec context on cuda:0 for the whole GPU test session; fails loudly without a GPU."""
    c = zk.Context(0)
    yield c
    c.close()
