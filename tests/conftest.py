import os
import sys

import pytest
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (B200, sm_100a); run with -m gpu")


def pytest_collection_modifyitems(config, items):
    if torch.cuda.is_available():
        return
    skip = pytest.mark.skip(reason="no CUDA device")
    for item in items:
        if "gpu" in item.keywords:
            item.add_marker(skip)


@pytest.fixture(scope="session")
def cuda():
    return torch.device("cuda:0")


@pytest.fixture(autouse=True)
def _seed_global_rng():
    """The reference's host logic draws from torch's global RNG (e.g. the multinomial restart selection of
    initialize_q_batch, initializers.py:946; prior resampling in fit retries), and so does the mirror here: seed it per test so
    that every test sees the same stream regardless of test order.  With the stream fixed the GPU path is bit-reproducible
    within and across processes (tools/determinism_check.py)."""
    torch.manual_seed(0)
    if torch.cuda.is_available():
        torch.cuda.manual_seed_all(0)
    yield
