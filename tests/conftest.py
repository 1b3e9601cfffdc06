import os
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

GOLDEN = os.path.join(ROOT, "tests", "golden")


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box)")


def load_golden(name):
    return dict(np.load(os.path.join(GOLDEN, name + ".npz")))


@pytest.fixture(scope="session")
def golden():
    return load_golden


@pytest.fixture(scope="session")
def oracle():
    """CPU oracle (test infrastructure): C restatement of the reference's numpy path."""
    from oracle import oracle as orc
    orc.build()
    return orc


@pytest.fixture(autouse=True)
def _default_config():
    """every test starts from (and leaves behind) the default hot-path configuration"""
    from mast3r_slam_b200 import config
    config.set_config({})
    yield
    config.set_config({})
