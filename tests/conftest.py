import sys
from pathlib import Path

import pytest

ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT))


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a B200 (run with -m gpu on the GPU box)")


@pytest.fixture(scope="session")
def oracle():
    from oracle.pyoracle import Oracle
    return Oracle()


@pytest.fixture(scope="session")
def refq():
    from oracle.pyoracle import RefQuant, have_ref
    if not have_ref():
        pytest.skip("oracle/_ref not built")
    return RefQuant()


@pytest.fixture(scope="session")
def harness():
    from oracle.pyoracle import RefHarness, have_ref
    if not have_ref():
        pytest.skip("oracle/_ref not built")
    return RefHarness.get()


@pytest.fixture(scope="session")
def pkg():
    import b200_pkg
    return b200_pkg.load()


@pytest.fixture(scope="session")
def ctx(pkg):
    c = pkg.Context(0)
    yield c
    c.close()
