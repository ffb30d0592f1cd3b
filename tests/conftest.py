import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a B200 (run with -m gpu on the GPU box)")


@pytest.fixture(scope="session")
def port():
    from oracle import Port
    return Port()


@pytest.fixture(scope="session")
def ref():
    from oracle import have_ref
    if not have_ref():
        pytest.skip("oracle/_ref/libref_shim.so not built (needs /root/reference at build time)")
    from oracle.ref import Ref
    return Ref()


@pytest.fixture(scope="session")
def ctx():
    from lazac_copy_number_b200.api import Context
    c = Context(0)  # raises without a GPU or without the built library: no CPU fallback
    yield c
    c.close()
