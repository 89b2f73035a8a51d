import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box)")


@pytest.fixture(scope="session")
def oracle():
    from oracle import pyoracle
    return pyoracle.port()


@pytest.fixture(scope="session")
def ref_oracle():
    from oracle import pyoracle
    pyoracle.build()
    r = pyoracle.ref()
    if r is None:
        pytest.skip("oracle/_ref not built (needs /root/reference)")
    return r
