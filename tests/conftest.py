import json
import os
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))

GOLDEN_DIR = os.path.join(ROOT, "tests", "golden")
FIX_DIR = os.path.join(GOLDEN_DIR, "fixtures")


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box)")


@pytest.fixture(scope="session")
def golden():
    return np.load(os.path.join(GOLDEN_DIR, "golden.npz"))


@pytest.fixture(scope="session")
def fixture_meta():
    with open(os.path.join(GOLDEN_DIR, "fixtures.json")) as f:
        return json.load(f)


def fixture_names():
    with open(os.path.join(GOLDEN_DIR, "fixtures.json")) as f:
        return sorted(json.load(f).keys())


def fixture_path(name):
    return os.path.join(FIX_DIR, name + ".pgen")
