"""Shared pytest configuration.

* ``gpu`` marker: tests that need a real B200.  Everything else runs on CPU.
* Seeding mirrors the reference's autouse fixture (tests/conftest.py:47-64).
* ``/root/reference`` is never read here: golden vectors live in tests/golden/.
"""

from __future__ import annotations

import os
import random
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

GOLDEN_DIR = os.path.join(ROOT, "tests", "golden")


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box)")


@pytest.fixture(autouse=True)
def _seed():
    seed = int(os.environ.get("VLLM_METAL_TEST_SEED", "0"))
    random.seed(seed)
    np.random.seed(seed)
    try:
        import torch

        torch.manual_seed(seed)
    except Exception:  # noqa: BLE001
        pass
    yield


@pytest.fixture(scope="session")
def ops():
    """The native operator object; GPU tests fail loudly if it cannot load."""
    import torch

    assert torch.cuda.is_available(), "gpu-marked test running without CUDA"
    from vllm_metal_b200.metal import get_ops

    return get_ops()


@pytest.fixture(scope="session")
def golden_attention():
    return np.load(os.path.join(GOLDEN_DIR, "attention_golden.npz"))


@pytest.fixture(scope="session")
def golden_metadata():
    import json

    with open(os.path.join(GOLDEN_DIR, "metadata_golden.json")) as f:
        return json.load(f)
