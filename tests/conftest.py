import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box with -m gpu)")


@pytest.fixture(scope="session")
def golden():
    import json

    import numpy as np
    d = os.path.join(ROOT, "tests", "golden")
    return np.load(os.path.join(d, "reference_golden.npz")), json.load(open(os.path.join(d, "reference_golden.json")))


@pytest.fixture(scope="session")
def pkg():
    import importlib
    return importlib.import_module("gpt-oss-vllm_b200")


@pytest.fixture(scope="session")
def cuda_device():
    import torch
    if not torch.cuda.is_available():
        pytest.skip("no CUDA device")
    return torch.device("cuda:0")
