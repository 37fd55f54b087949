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


def close_rms(x, ref, rtol=1e-5, support=None):
    """Parity norm of SURVEY 7.3-2: |x-ref| <= rtol * max(|ref|, rms(ref)).

    ``support`` (bool tensor): the entries that carry signal, e.g. the unmasked logits of a
    gradient.  rms is then taken over the support only (a tensor that is 95 % structural zeros
    would otherwise shrink the rms term to nothing and turn the test into a pure relative one
    on entries that are themselves differences of nearly equal fp32 numbers); entries outside
    the support must be checked for exact zeros by the caller.
    """
    x = torch.as_tensor(x).double().cpu()
    ref = torch.as_tensor(ref).double().cpu()
    if support is not None:
        sel = torch.as_tensor(support).cpu().bool()
        x, ref = x[sel], ref[sel]
    if ref.numel() == 0:
        return True, 0.0
    rms = ref.pow(2).mean().sqrt()
    bound = rtol * torch.maximum(ref.abs(), rms.expand_as(ref)) + 1e-30
    err = (x - ref).abs()
    return bool((err <= bound).all()), float((err / bound).max())
