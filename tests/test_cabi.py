"""The C-ABI shared library: loads on a CPU-only box, exports every symbol include/mbeast.h
declares, and validates arguments before touching the GPU.  No compute calls here."""
import ctypes
import os
import re

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _declared():
    text = open(os.path.join(ROOT, "include", "mbeast.h")).read()
    return sorted(set(re.findall(r"MB_API\s+[\w\s\*]+?\b(mb_\w+)\s*\(", text)))


def test_header_declares_the_expected_surface():
    names = _declared()
    for n in ("mb_version", "mb_last_error", "mb_head_eval_fwd", "mb_head_eval_bwd", "mb_head_sample",
              "mb_vtrace_workspace_bytes", "mb_vtrace_loss", "mb_set_tuning"):
        assert n in names


def test_library_loads_and_exports_every_declared_symbol():
    from microbeast_b200 import _lib
    assert os.path.isfile(_lib.LIB_PATH)
    raw = ctypes.CDLL(_lib.LIB_PATH)
    for name in _declared():
        assert hasattr(raw, name), f"{name} declared in include/mbeast.h but not exported"
        assert name in _lib.SIGNATURES, f"{name} has no ctypes signature in _lib.py"
    assert _lib.ABI_VERSION == 100


def test_argument_validation_needs_no_gpu():
    from microbeast_b200 import _lib
    lib = _lib.lib
    null = ctypes.c_void_p(0)
    rc = lib.mb_head_eval_fwd(null, null, null, null, null, -1, 64, null, null)
    assert rc == 1 and "n_frames" in _lib.last_error()
    rc = lib.mb_head_eval_fwd(null, null, null, null, null, 4, 0, null, null)
    assert rc == 1
    rc = lib.mb_head_eval_fwd(null, null, null, null, null, 4, 64, null, null)
    assert rc == 1 and "null" in _lib.last_error()
    assert lib.mb_head_eval_fwd(null, null, null, null, null, 0, 64, null, null) == 0     # empty batch
    assert lib.mb_head_sample(null, null, null, null, null, null, 0, 64, null, null) == 0
    assert lib.mb_head_eval_bwd(null, null, null, null, null, null, 0, 64, null, null) == 0
    rc = lib.mb_vtrace_loss(null, null, null, null, null, null, null, 8, 0, 0, 0,
                            0.99, 1.0, 1.0, 1.0, 0.01, 0.0, null, null, null, null, null, null, 0, null)
    assert rc == 1 and "T=" in _lib.last_error()
    assert lib.mb_vtrace_workspace_bytes(0) >= 256
    assert lib.mb_vtrace_workspace_bytes(65536) >= 256 + 2048 * 24
    with pytest.raises(ValueError):
        _lib.set_tuning("no_such_knob", 1)
    _lib.set_tuning("vtrace_variant", -1)


def test_ops_refuse_cpu_tensors():
    import torch
    from microbeast_b200 import ops
    z = torch.zeros(2, 78 * 4); m = torch.ones(2, 78 * 4, dtype=torch.uint8); a = torch.zeros(2, 28, dtype=torch.int64)
    with pytest.raises(RuntimeError, match="no CPU fallback"):
        ops.head_eval_fwd(z, m, a)
    with pytest.raises(RuntimeError, match="no CPU fallback"):
        ops.head_sample(z, m, torch.zeros(2, 28))
    x = torch.zeros(4, 16)
    with pytest.raises(RuntimeError, match="no CPU fallback"):
        ops.vtrace_loss(x, x, x, torch.zeros(4), x, x.bool(), x)


def test_every_entry_point_has_a_python_wrapper():
    """Each compute entry point of the C ABI is called from microbeast_b200/ops.py (a wrapper dropped by
    accident would otherwise only show up on the GPU box)."""
    import inspect
    from microbeast_b200 import _lib, ops
    src = inspect.getsource(ops)
    housekeeping = {"mb_version", "mb_last_error", "mb_set_tuning", "mb_vtrace_loss_static"}
    missing = [name for name in _lib.SIGNATURES if name not in housekeeping and f"lib.{name}" not in src]
    assert not missing, missing
