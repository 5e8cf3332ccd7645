import os
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (ROOT, os.path.join(ROOT, "sq-learn_b200")):
    if p not in sys.path:
        sys.path.insert(0, p)
GOLDEN = os.path.join(ROOT, "tests", "golden")


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (B200); run with -m gpu")


@pytest.fixture(scope="session")
def golden_det():
    import json
    with open(os.path.join(GOLDEN, "deterministic.json")) as f:
        return json.load(f)


@pytest.fixture(scope="session")
def golden_qmeans():
    return np.load(os.path.join(GOLDEN, "qmeans_noiseless.npz"))


@pytest.fixture(scope="session")
def golden_stoch():
    return np.load(os.path.join(GOLDEN, "stochastic_reference.npz"))


@pytest.fixture(scope="session")
def lib():
    from sqlearn import _cabi
    return _cabi


@pytest.fixture()
def cuda():
    import torch
    if not torch.cuda.is_available():
        pytest.skip("no CUDA device")
    return torch.device("cuda", 0)
