import gzip
import json
import os
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)
GOLD = os.path.join(ROOT, "tests", "golden")


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box with -m gpu)")


def jload(name):
    with gzip.open(os.path.join(GOLD, name), "rt") as f:
        return json.load(f)


def npz(name):
    return np.load(os.path.join(GOLD, name))


@pytest.fixture(scope="session")
def vocab43():
    v = {0: "blank", 1: "UNK"}
    for line in open(os.path.join(GOLD, "assets", "units")):
        w = line.strip()
        if w:
            v[len(v)] = w
    assert len(v) == 43
    return v


@pytest.fixture(scope="session")
def arpa():
    return os.path.join(GOLD, "assets", "lm_phone_bg.arpa")
