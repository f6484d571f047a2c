import os
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)
GOLDEN = os.path.join(ROOT, "tests", "golden")


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box with -m gpu)")


@pytest.fixture(scope="session")
def golden():
    def load(name):
        return dict(np.load(os.path.join(GOLDEN, name)))
    return load


@pytest.fixture(scope="session")
def state_dict():
    import torch
    return torch.load(os.path.join(GOLDEN, "puflow-x4-pu1k.pt"), map_location="cpu")


def _ensure_native_built():
    """The package refuses to import without its CUDA library; nvcc cross-compiles it here (no GPU
    needed).  On the GPU box the prebuilt .so travels with the snapshot."""
    so = os.path.join(ROOT, "puflow_b200", "_lib", "libpuflow_b200.so")
    if not os.path.exists(so):
        import __graft_entry__
        __graft_entry__.build()


_ensure_native_built()
