import os
import sys

import pytest

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)
for p in (HERE, ROOT):
    if p not in sys.path:
        sys.path.insert(0, p)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box)")


@pytest.fixture(scope="session")
def oracle():
    """Strongest CPU checker available (compiled reference if present, else the port)."""
    import oracle_lib
    return oracle_lib.checker()


@pytest.fixture(scope="session")
def pkg():
    """The product package (compression-accelerator_b200/), loaded by path."""
    import __graft_entry__ as g
    if not os.path.exists(os.path.join(g.PKG_DIR, "libsnappy_b200.so")):
        g.build()
    return g.load_package()


@pytest.fixture(scope="session")
def ctx(pkg):
    """A device context on cuda:0; GPU tests only."""
    c = pkg.Context(0)
    yield c
    c.close()
