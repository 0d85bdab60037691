import importlib
import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (ROOT, os.path.join(ROOT, "tests")):
    if p not in sys.path:
        sys.path.insert(0, p)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a B200 (sm_100a); run with -m gpu on the GPU box")
    config.addinivalue_line("markers", "slow: takes more than a few seconds")


@pytest.fixture(scope="session")
def pkg():
    return importlib.import_module("multimodal-reid_b200")


@pytest.fixture(scope="session")
def synth():
    return importlib.import_module("multimodal-reid_b200.synth")


@pytest.fixture(scope="session")
def oracle():
    from oracle import reid_oracle
    return reid_oracle


@pytest.fixture(scope="session")
def eng():
    import torch
    if not torch.cuda.is_available():
        pytest.skip("no CUDA device")
    engine = importlib.import_module("multimodal-reid_b200.engine")
    return engine.Engine(engine="auto")


@pytest.fixture(scope="session")
def engine_mod():
    return importlib.import_module("multimodal-reid_b200.engine")
