import os
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (ROOT, os.path.join(ROOT, "oracle")):
    if p not in sys.path:
        sys.path.insert(0, p)

GOLD = os.path.join(ROOT, "tests", "golden")


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box)")


def load_golden(name):
    return np.load(os.path.join(GOLD, name))


def golden_args(d):
    return (int(d["n_cam"]), int(d["n_pts"]), d["ci"], d["pi"], d["uv"], d["K"])


@pytest.fixture(scope="session")
def engine():
    import b200ba
    eng = b200ba.Engine(0)
    yield eng
    eng.close()
