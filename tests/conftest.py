import importlib
import os
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box)")


@pytest.fixture(scope="session")
def synth():
    return importlib.import_module("3dposeestimation_b200.synth")


@pytest.fixture(scope="session")
def golden():
    return np.load(os.path.join(ROOT, "tests", "golden", "golden.npz"))


@pytest.fixture(scope="session")
def po():
    from oracle import pyoracle
    pyoracle.build()
    return pyoracle


@pytest.fixture(scope="session")
def capi():
    mod = importlib.import_module("3dposeestimation_b200.capi")
    mod.build()
    return mod


@pytest.fixture(scope="session")
def gpu(capi):
    if capi.device_count() < 1:
        pytest.fail("GPU test selected but no CUDA device is visible (there is no CPU fallback)")
    return capi
