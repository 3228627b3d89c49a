import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box)")


@pytest.fixture(scope="session")
def built():
    """Build (or reuse) libptx.so, the oracle and the host-emulation library."""
    import __graft_entry__ as ge
    ge.build()
    return ge


@pytest.fixture(scope="session")
def pkg(built):
    return built.load_package()


@pytest.fixture(scope="session")
def oracle(built):
    import oracle_binding
    return oracle_binding.Oracle()


@pytest.fixture(scope="session")
def ctx(pkg):
    c = pkg.Context(0)
    yield c
    c.close()


SCENES = os.path.join(ROOT, "scenes")
GOLDEN = os.path.join(ROOT, "tests", "golden")
