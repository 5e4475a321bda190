import importlib
import os
import sys

import pytest

TESTS = os.path.dirname(os.path.abspath(__file__))
if TESTS not in sys.path:
    sys.path.insert(0, TESTS)
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

ORACLE_LIB = os.path.join(ROOT, "oracle", "libba_oracle.so")


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box with -m gpu)")


@pytest.fixture(scope="session", autouse=True)
def _built():
    import __graft_entry__ as g
    g.build()


@pytest.fixture(scope="session")
def ba(_built):
    return importlib.import_module("3dreconstruction_b200")


@pytest.fixture(scope="session")
def oracle(ba):
    """The CPU oracle, bound through the same ctypes wrapper as the engine. Checker only."""
    return ba.Library(ORACLE_LIB)


@pytest.fixture(scope="session")
def engine_lib(ba):
    """The CUDA engine library; fails loudly if it is missing (no fallback)."""
    return ba.engine_library()
