import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box)")


@pytest.fixture(scope="session")
def fsq():
    from __graft_entry__ import load_package
    return load_package()


@pytest.fixture(scope="session")
def synth():
    from __graft_entry__ import load_synth
    return load_synth()


@pytest.fixture(scope="session")
def oracle():
    import oraclelib
    return oraclelib.oracle()


@pytest.fixture()
def ctx(fsq):
    c = fsq.Context(0)
    yield c
    c.close()
