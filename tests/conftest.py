import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "oracle"))


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box)")


@pytest.fixture(scope="session")
def oracle():
    import bm_oracle
    bm_oracle.build()
    return bm_oracle.Oracle()


@pytest.fixture(scope="session")
def bm():
    import bucketmap_b200
    return bucketmap_b200


@pytest.fixture(scope="session")
def ctx(bm):
    c = bm.Context(0)
    yield c
    c.close()
