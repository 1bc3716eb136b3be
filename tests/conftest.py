import os, sys
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box)")
    config.addinivalue_line("markers", "ref: needs oracle/_ref (the compiled reference)")


@pytest.fixture(scope="session")
def oracle():
    from tests import helpers
    return helpers.load_oracle()


@pytest.fixture(scope="session")
def ref():
    from tests import helpers
    r = helpers.load_ref()
    if r is None:
        pytest.skip("oracle/_ref/libde265_ref.so not built (reference sources absent)")
    return r


@pytest.fixture(scope="session")
def ctx():
    from libde265_b200 import lib
    c = lib.Context(0)   # raises (test error, not skip) if the CUDA library/device is missing
    yield c
    c.close()
