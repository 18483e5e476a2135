"""pytest wiring: registers the `gpu` marker, puts the package container dir and oracle/ on
sys.path, and offers golden-fixture loading helpers."""
import os
import sys

import numpy as np
import pytest
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
PKG_DIR = os.path.join(ROOT, "robot-manipulation-vq-vae_b200")
for p in (PKG_DIR, os.path.join(ROOT, "oracle"), ROOT):
    if p not in sys.path:
        sys.path.insert(0, p)

GOLDEN = os.path.join(ROOT, "tests", "golden")


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box with -m gpu)")


def load_golden(name):
    """-> dict[str, torch.Tensor] from tests/golden/<name>.npz"""
    with np.load(os.path.join(GOLDEN, name + ".npz")) as f:
        return {k: torch.from_numpy(np.array(f[k])) for k in f.files}


def sub(d, prefix):
    """entries of d under 'prefix.' with the prefix stripped"""
    n = len(prefix) + 1
    return {k[n:]: v for k, v in d.items() if k.startswith(prefix + ".")}


@pytest.fixture(scope="session")
def golden():
    return load_golden


def pytest_collection_modifyitems(config, items):
    if torch.cuda.is_available():
        return
    skip = pytest.mark.skip(reason="no CUDA device")
    for it in items:
        if "gpu" in it.keywords:
            it.add_marker(skip)
