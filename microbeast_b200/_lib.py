"""ctypes binding of ``libmbeast_sm100.so`` (``include/mbeast.h``).

There is no CPU fallback: if the shared library is missing this module raises
at import time, and every op refuses non-CUDA tensors.
"""
from __future__ import annotations

import ctypes
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "libmbeast_sm100.so")

_c_float_p = ctypes.c_void_p      # device pointers travel as integers
_P = ctypes.c_void_p

SIGNATURES = {
    "mb_version": (ctypes.c_int, []),
    "mb_last_error": (ctypes.c_char_p, []),
    "mb_head_eval_fwd": (ctypes.c_int, [_P, _P, _P, _P, _P, ctypes.c_int64, ctypes.c_int32, _P, _P]),
    "mb_head_eval_bwd": (ctypes.c_int, [_P, _P, _P, _P, _P, _P, ctypes.c_int64, ctypes.c_int32, _P, _P]),
    "mb_head_sample": (ctypes.c_int, [_P, _P, _P, _P, _P, _P, ctypes.c_int64, ctypes.c_int32, _P, _P]),
    "mb_head_sample_exp": (ctypes.c_int, [_P, _P, _P, _P, ctypes.c_int64, ctypes.c_int32, _P]),
    "mb_head_eval_cells_fwd": (ctypes.c_int, [_P, _P, _P, _P, ctypes.c_int64, _P, _P]),
    "mb_head_eval_cells_bwd": (ctypes.c_int, [_P, _P, _P, _P, _P, _P, ctypes.c_int64, _P, _P]),
    "mb_vtrace_workspace_bytes": (ctypes.c_size_t, [ctypes.c_int64]),
    "mb_vtrace_loss": (ctypes.c_int, [_P, _P, _P, _P, _P, _P, _P, ctypes.c_int64, ctypes.c_int32,
                                      ctypes.c_int64, ctypes.c_int64,
                                      ctypes.c_float, ctypes.c_float, ctypes.c_float, ctypes.c_float,
                                      ctypes.c_float, ctypes.c_float,
                                      _P, _P, _P, _P, _P, _P, ctypes.c_size_t, _P]),
    "mb_vtrace_loss_tmajor": (ctypes.c_int, [_P, _P, _P, _P, _P, _P, _P, ctypes.c_int64, ctypes.c_int32,
                                             ctypes.c_int64, ctypes.c_int64,
                                             ctypes.c_float, ctypes.c_float, ctypes.c_float, ctypes.c_float,
                                             ctypes.c_float, ctypes.c_float,
                                             _P, _P, _P, _P, _P, _P, ctypes.c_size_t, _P]),
    "mb_vtrace_loss_static": (ctypes.c_int, [_P, _P, _P, _P, _P, _P, _P, ctypes.c_int64, ctypes.c_int32,
                                             ctypes.c_int64, ctypes.c_int64,
                                             ctypes.c_float, ctypes.c_float, ctypes.c_float, ctypes.c_float,
                                             ctypes.c_float, ctypes.c_float,
                                             _P, _P, _P, _P, _P, _P]),
    "mb_unpack_rollout": (ctypes.c_int, [_P, _P, _P, _P, ctypes.c_int64, ctypes.c_int32, _P]),
    "mb_unpack_obs": (ctypes.c_int, [_P, _P, ctypes.c_int64, ctypes.c_int32, ctypes.c_int32, _P]),
    "mb_active_cells": (ctypes.c_int, [_P, _P, ctypes.c_int64, _P]),
    "mb_conv3x3_fwd_supported": (ctypes.c_int, [ctypes.c_int32, ctypes.c_int32, ctypes.c_int32, ctypes.c_int32]),
    "mb_conv3x3_fwd": (ctypes.c_int, [_P, _P, _P, _P, ctypes.c_int64, ctypes.c_int32, ctypes.c_int32, ctypes.c_int32,
                                      ctypes.c_int32, ctypes.c_int32, _P]),
    "mb_conv3x3_fwd_fused": (ctypes.c_int, [_P, _P, _P, _P, ctypes.c_int64, ctypes.c_int32, ctypes.c_int32, ctypes.c_int32,
                                            ctypes.c_int32, ctypes.c_int32, ctypes.c_int32, _P, _P, _P]),
    "mb_conv3x3_wgrad_relu": (ctypes.c_int, [_P, _P, _P, ctypes.c_int32, ctypes.c_int64, ctypes.c_int32, ctypes.c_int32,
                                             ctypes.c_int32, ctypes.c_int32, ctypes.c_int32, _P, _P]),
    "mb_conv1_bits_fwd": (ctypes.c_int, [_P, _P, _P, _P, ctypes.c_int64, ctypes.c_int32, ctypes.c_int32, ctypes.c_int32,
                                         ctypes.c_int32, _P]),
    "mb_conv1_bits_wgrad": (ctypes.c_int, [_P, _P, _P, ctypes.c_int64, ctypes.c_int32, ctypes.c_int32, ctypes.c_int32,
                                           ctypes.c_int32, _P]),
    "mb_split_tf32": (ctypes.c_int, [_P, _P, _P, ctypes.c_int64, _P]),
    "mb_linear_tf32x3": (ctypes.c_int, [_P, _P, _P, _P, _P, _P, ctypes.c_int64, ctypes.c_int64, ctypes.c_int32, _P, _P]),
    "mb_pack_cells": (ctypes.c_int, [_P, _P, _P, _P, ctypes.c_int64, ctypes.c_int32, _P, _P]),
    "mb_pack_actor_weight": (ctypes.c_int, [_P, _P, _P, _P, _P, ctypes.c_int32, ctypes.c_int32, _P]),
    "mb_actor_head_fwd": (ctypes.c_int, [_P, _P, _P, _P, _P, _P, _P, _P, ctypes.c_int32, _P, _P, ctypes.c_int64,
                                         ctypes.c_int32, ctypes.c_int32, _P, _P]),
    "mb_actor_head_list_workspace_bytes": (ctypes.c_size_t, [ctypes.c_int64, ctypes.c_int32]),
    "mb_actor_head_list_fwd": (ctypes.c_int, [_P, _P, _P, _P, _P, _P, _P, _P, _P, _P, _P, _P, ctypes.c_size_t, ctypes.c_int64,
                                              ctypes.c_int32, ctypes.c_int32, _P, _P]),
    "mb_cells_bwd_workspace_bytes": (ctypes.c_size_t, [ctypes.c_int64, ctypes.c_int32]),
    "mb_cells_bwd": (ctypes.c_int, [_P, _P, _P, _P, _P, _P, _P, _P, _P, _P, _P, ctypes.c_size_t, ctypes.c_int64, ctypes.c_int32,
                                    ctypes.c_int32, _P]),
    "mb_maxpool3s2_bwd": (ctypes.c_int, [_P, _P, _P, ctypes.c_int64, ctypes.c_int32, ctypes.c_int32, _P]),
    "mb_maxpool3s2_fwd": (ctypes.c_int, [_P, _P, _P, ctypes.c_int64, ctypes.c_int32, ctypes.c_int32, _P]),
    "mb_maxpool3s2_bwd_code": (ctypes.c_int, [_P, _P, _P, ctypes.c_int64, ctypes.c_int32, ctypes.c_int32, _P]),
    "mb_conv3x3_wgrad_blocks": (ctypes.c_int, [ctypes.c_int64, ctypes.c_int32, ctypes.c_int32, ctypes.c_int32,
                                               ctypes.c_int32]),
    "mb_conv3x3_wgrad": (ctypes.c_int, [_P, _P, _P, ctypes.c_int32, ctypes.c_int64, ctypes.c_int32, ctypes.c_int32,
                                        ctypes.c_int32, ctypes.c_int32, _P]),
    "mb_set_tuning": (ctypes.c_int, [ctypes.c_char_p, ctypes.c_int32]),
}


def _load() -> ctypes.CDLL:
    if not os.path.isfile(LIB_PATH):
        raise RuntimeError(
            f"microbeast_b200: CUDA extension {LIB_PATH} is missing. Build it with "
            "`python -c 'import __graft_entry__ as g; g.build()'` or microbeast_b200/csrc/build.sh "
            "(nvcc, sm_100a). There is no CPU fallback.")
    lib = ctypes.CDLL(LIB_PATH)
    for name, (restype, argtypes) in SIGNATURES.items():
        fn = getattr(lib, name)          # AttributeError here = stale .so: fail loudly
        fn.restype = restype
        fn.argtypes = argtypes
    return lib


lib = _load()
ABI_VERSION = lib.mb_version()


def last_error() -> str:
    msg = lib.mb_last_error()
    return msg.decode("utf-8", "replace") if msg else ""


def set_tuning(key: str, value: int) -> None:
    if lib.mb_set_tuning(key.encode(), int(value)) != 0:
        raise ValueError(last_error())
