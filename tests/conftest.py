import importlib
import json
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


def pkg():
    """The product package; its directory name has a dash, so it is imported by string."""
    return importlib.import_module("image-stitching_b200")


def workloads():
    return importlib.import_module("image-stitching_b200.workloads")


def _load_cases(fname, prefix):
    z = np.load(os.path.join(GOLDEN, fname))
    meta = json.loads(bytes(z["meta"]).decode())
    return z, meta


@pytest.fixture(scope="session")
def small_cases():
    return _load_cases("small_cases.npz", "c")


@pytest.fixture(scope="session")
def undistort_cases():
    return _load_cases("undistort_cases.npz", "u")


@pytest.fixture(scope="session")
def border_constant_cases():
    return _load_cases("border_constant_cases.npz", "b")


@pytest.fixture(scope="session")
def known_answers():
    with open(os.path.join(GOLDEN, "known_answers.json")) as f:
        return json.load(f)
