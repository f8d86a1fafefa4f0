import os
import sys

import numpy as np
import pytest
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)
GOLDEN = os.path.join(ROOT, "tests", "golden")


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (B200); run with -m gpu")


def pytest_collection_modifyitems(config, items):
    if torch.cuda.is_available():
        return
    skip = pytest.mark.skip(reason="no CUDA device")
    for item in items:
        if "gpu" in item.keywords:
            item.add_marker(skip)


def load_golden(name):
    """tests/golden/<name>.npz as a dict of torch tensors / numpy scalars."""
    out = {}
    with np.load(os.path.join(GOLDEN, name + ".npz")) as z:
        for k in z.files:
            a = z[k]
            out[k] = torch.from_numpy(a) if a.ndim > 0 else a.item()
    return out


def randn(shape, seed, scale=1.0):
    """Same seeded-input helper as oracle/make_golden.py."""
    return torch.randn(shape, generator=torch.Generator().manual_seed(seed)) * scale


def cosine(a, b):
    a, b = a.detach().double().flatten(), b.detach().double().flatten()
    return float((a @ b) / (a.norm() * b.norm()))


@pytest.fixture(scope="session")
def reference_tree():
    """`import comfy` works (the staged copy under baseline/_ref on the GPU box; --cpu mode on machines without a GPU)."""
    from comfy_gen_server_b200.reference import ensure_importable
    where = ensure_importable(cpu=not torch.cuda.is_available())
    import comfy.utils
    comfy.utils.PROGRESS_BAR_ENABLED = False
    return where


@pytest.fixture
def installed(reference_tree):
    """integration.install() for the duration of a test (undone afterwards)."""
    from comfy_gen_server_b200 import integration
    integration.install()
    yield
    integration.uninstall()
