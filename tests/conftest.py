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
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box)")


def pytest_collection_modifyitems(config, items):
    if torch.cuda.is_available():
        return
    skip = pytest.mark.skip(reason="no CUDA device")
    for item in items:
        if "gpu" in item.keywords:
            item.add_marker(skip)


@pytest.fixture(scope="session")
def golden():
    def load(name):
        return np.load(os.path.join(GOLDEN, name + ".npz"), allow_pickle=False)
    return load


def close_rms(x, ref, rtol=1e-5):
    """Parity norm of SURVEY 7.3-2: |x-ref| <= rtol * max(|ref|, rms(ref))."""
    x = torch.as_tensor(x).double().cpu()
    ref = torch.as_tensor(ref).double().cpu()
    rms = ref.pow(2).mean().sqrt() if ref.numel() else torch.tensor(0.0)
    bound = rtol * torch.maximum(ref.abs(), rms.expand_as(ref)) + 1e-30
    err = (x - ref).abs()
    ok = bool((err <= bound).all())
    worst = float((err / bound).max()) if ref.numel() else 0.0
    return ok, worst
