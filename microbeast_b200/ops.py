"""Tensor-level wrappers of the C ABI and their autograd glue.

Every function takes CUDA tensors, launches on ``torch.cuda.current_stream()`` and
returns without synchronising.  CPU tensors are rejected: there is no fallback path.
"""
from __future__ import annotations

import ctypes
from typing import Dict, Optional, Tuple

import torch

from . import _lib
from ._lib import lib

LOGITS_PER_CELL = 78
COMPONENTS_PER_CELL = 7
STATUS_BAD_ACTION = 1
STATUS_TIMEOUT = 2

# launches issued through this module, by kernel name (bench.py reports the count)
launch_counts: Dict[str, int] = {}

_status_words: Dict[int, torch.Tensor] = {}
_workspaces: Dict[Tuple[int, int], torch.Tensor] = {}
_retired_workspaces: list = []


_timing = None      # list of (name, start_event, end_event) while kernel timing is on


_timing_only = None


def start_kernel_timing(only=None) -> None:
    """Record a CUDA-event pair around every kernel launch (on the launching stream); ``only``: an iterable of launch
    names to restrict it to (an event pair costs ~3 us of stream time, which adds up over the ~120 launches of a step)."""
    global _timing, _timing_only
    _timing = []
    _timing_only = set(only) if only is not None else None


def stop_kernel_timing() -> Dict[str, list]:
    """Stops recording and returns ``{kernel: [ms, ...]}`` (synchronises)."""
    global _timing
    rec, _timing = _timing or [], None
    torch.cuda.synchronize()
    out: Dict[str, list] = {}
    for name, s, e in rec:
        out.setdefault(name, []).append(s.elapsed_time(e))
    return out


launch_notes: Dict[str, dict] = {}


class _Launch:
    """Counts a launch and, when timing is on, brackets it with CUDA events."""

    def __init__(self, name: str, note=None):
        self.name = name
        if note is not None:
            launch_notes[name] = note        # problem size of the latest launch under this name (bench.py reads it)

    def __enter__(self):
        self.timed = _timing is not None and (_timing_only is None or self.name in _timing_only)
        if self.timed:
            self.start = torch.cuda.Event(enable_timing=True)
            self.start.record()
        return self

    def __exit__(self, exc_type, exc, tb):
        if exc_type is None:
            launch_counts[self.name] = launch_counts.get(self.name, 0) + 1
            if self.timed and _timing is not None:
                end = torch.cuda.Event(enable_timing=True)
                end.record()
                _timing.append((self.name, self.start, end))
        return False


def total_launches() -> int:
    return sum(launch_counts.values())


def _ptr(t: Optional[torch.Tensor]):
    return ctypes.c_void_p(0 if t is None else t.data_ptr())


def _stream():
    return ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)


def _check(rc: int) -> None:
    if rc == 0:
        return
    msg = _lib.last_error()
    if rc == 1:
        raise ValueError(msg)
    raise RuntimeError(msg)


def _need_cuda(**tensors) -> torch.device:
    dev = None
    for name, t in tensors.items():
        if t is None:
            continue
        if not t.is_cuda:
            raise RuntimeError(f"microbeast_b200: `{name}` is on {t.device}; the kernels are CUDA-only "
                               "(sm_100a) and there is no CPU fallback")
        if dev is None:
            dev = t.device
        elif t.device != dev:
            raise RuntimeError(f"microbeast_b200: `{name}` is on {t.device}, expected {dev}")
    return dev


def _as(t: torch.Tensor, dtype, name: str) -> torch.Tensor:
    if t.dtype != dtype:
        raise TypeError(f"`{name}` must be {dtype}, got {t.dtype}")
    return t if t.is_contiguous() else t.contiguous()


def _mask_u8(mask: torch.Tensor) -> torch.Tensor:
    """Any mask the reference accepts (it casts with ``masks.type(torch.BoolTensor)``, ``model.py:40``: uint8,
    bool, the int32 of gym-microRTS ``get_action_mask()``, floats) as contiguous uint8, non-zero = valid."""
    if mask.dtype == torch.bool:
        mask = mask.contiguous().view(torch.uint8)
    elif mask.dtype != torch.uint8:
        mask = (mask != 0).view(torch.uint8)
    return _as(mask, torch.uint8, "mask")


def status_word(device: torch.device) -> torch.Tensor:
    """Per-device int32 word the kernels OR error bits into (zeroed at creation)."""
    idx = device.index if device.index is not None else torch.cuda.current_device()
    w = _status_words.get(idx)
    if w is None:
        w = torch.zeros(1, dtype=torch.int32, device=torch.device("cuda", idx))
        _status_words[idx] = w
    return w


def check_status(device: Optional[torch.device] = None) -> None:
    """Synchronising check of the status word; raises like torch's ``validate_args`` (Q13)."""
    device = device or torch.device("cuda", torch.cuda.current_device())
    w = status_word(device)
    bits = int(w.item())
    if bits:
        w.zero_()
    if bits & STATUS_TIMEOUT:
        raise RuntimeError("microbeast_b200: a kernel pipeline wait timed out (internal error)")
    if bits & STATUS_BAD_ACTION:
        raise ValueError("microbeast_b200: action index outside the support of its categorical "
                         "(reference: torch Categorical.log_prob validate_args)")


def _cells(width: int, per_cell: int, name: str) -> int:
    if width % per_cell:
        raise ValueError(f"`{name}` last dim {width} is not a multiple of {per_cell}")
    return width // per_cell


def head_eval_fwd(logits: torch.Tensor, mask: torch.Tensor, action: torch.Tensor):
    """``(logprobs [N], entropy [N])`` of the taken actions -- reference ``model.py:180-214``."""
    dev = _need_cuda(logits=logits, mask=mask, action=action)
    logits = _as(logits, torch.float32, "logits")
    mask = _mask_u8(mask)
    action = _as(action, torch.int64, "action")
    n = logits.shape[0]
    cells = _cells(logits.shape[1], LOGITS_PER_CELL, "logits")
    if mask.shape != logits.shape or tuple(action.shape) != (n, cells * COMPONENTS_PER_CELL):
        raise ValueError(f"shape mismatch: logits {tuple(logits.shape)} mask {tuple(mask.shape)} "
                         f"action {tuple(action.shape)}")
    logp = torch.empty(n, dtype=torch.float32, device=dev)
    ent = torch.empty(n, dtype=torch.float32, device=dev)
    with torch.cuda.device(dev), _Launch("head_eval_fwd"):
        _check(lib.mb_head_eval_fwd(_ptr(logits), _ptr(mask), _ptr(action), _ptr(logp), _ptr(ent),
                                    n, cells, _ptr(status_word(dev)), _stream()))
    return logp, ent


def head_eval_bwd(logits, mask, action, g_logp, g_ent, out: Optional[torch.Tensor] = None):
    """``dlogits [N,78C]`` for upstream grads ``g_logp [N]``, ``g_ent [N]``."""
    dev = _need_cuda(logits=logits, mask=mask, action=action, g_logp=g_logp, g_ent=g_ent, out=out)
    logits = _as(logits, torch.float32, "logits")
    mask = _mask_u8(mask)
    action = _as(action, torch.int64, "action")
    g_logp = _as(g_logp, torch.float32, "g_logp")
    g_ent = _as(g_ent, torch.float32, "g_ent")
    n = logits.shape[0]
    cells = _cells(logits.shape[1], LOGITS_PER_CELL, "logits")
    if g_logp.numel() != n or g_ent.numel() != n:
        raise ValueError("g_logp / g_ent must have one entry per frame")
    if out is None:
        out = torch.empty_like(logits)
    elif out.shape != logits.shape or out.dtype != torch.float32 or not out.is_contiguous():
        raise ValueError("`out` must be a contiguous float32 tensor shaped like logits")
    with torch.cuda.device(dev), _Launch("head_eval_bwd"):
        _check(lib.mb_head_eval_bwd(_ptr(logits), _ptr(mask), _ptr(action), _ptr(g_logp), _ptr(g_ent),
                                    _ptr(out), n, cells, _ptr(status_word(dev)), _stream()))
    return out


def head_sample(logits, mask, uniforms, want_entropy: bool = True):
    """Inverse-CDF sampling from caller-supplied uniforms (definition: ``oracle/head.py``).

    Returns ``(action [N,7C] i64, logprobs [N], entropy [N] or None)`` -- the actor branch
    of reference ``model.py:171-177,194-210``.
    """
    dev = _need_cuda(logits=logits, mask=mask, uniforms=uniforms)
    logits = _as(logits, torch.float32, "logits")
    mask = _mask_u8(mask)
    uniforms = _as(uniforms, torch.float32, "uniforms")
    n = logits.shape[0]
    cells = _cells(logits.shape[1], LOGITS_PER_CELL, "logits")
    if mask.shape != logits.shape or tuple(uniforms.shape) != (n, cells * COMPONENTS_PER_CELL):
        raise ValueError("shape mismatch between logits / mask / uniforms")
    action = torch.empty((n, cells * COMPONENTS_PER_CELL), dtype=torch.int64, device=dev)
    logp = torch.empty(n, dtype=torch.float32, device=dev)
    ent = torch.empty(n, dtype=torch.float32, device=dev) if want_entropy else None
    with torch.cuda.device(dev), _Launch("head_sample"):
        _check(lib.mb_head_sample(_ptr(logits), _ptr(mask), _ptr(uniforms), _ptr(action), _ptr(logp),
                                  _ptr(ent), n, cells, _ptr(status_word(dev)), _stream()))
    return action, logp, ent


def head_sample_exp(logits, mask, exponentials: Optional[torch.Tensor] = None, generator=None, want_entropy: bool = False):
    """The reference's sampling rule (``model.py:177``: ``Categorical.sample`` -> ``torch.multinomial`` =
    ``argmax(probs / Exp(1))``) with the exponentials as an input: ``exponentials [N, 78C]`` f32, drawn with
    ``torch.empty_like(logits).exponential_(generator=...)`` when omitted.  Returns
    ``(action [N,7C] i64, logprobs [N], entropy [N] or None)``."""
    dev = _need_cuda(logits=logits, mask=mask, exponentials=exponentials)
    logits = _as(logits, torch.float32, "logits")
    mask = _mask_u8(mask)
    n = logits.shape[0]
    cells = _cells(logits.shape[1], LOGITS_PER_CELL, "logits")
    if mask.shape != logits.shape:
        raise ValueError("shape mismatch between logits / mask")
    if exponentials is None:
        exponentials = torch.empty_like(logits).exponential_(generator=generator)
    exponentials = _as(exponentials, torch.float32, "exponentials")
    if exponentials.shape != logits.shape:
        raise ValueError("exponentials must be shaped like logits (one draw per category)")
    action = torch.empty((n, cells * COMPONENTS_PER_CELL), dtype=torch.int64, device=dev)
    with torch.cuda.device(dev), _Launch("head_sample_exp"):
        _check(lib.mb_head_sample_exp(_ptr(logits), _ptr(mask), _ptr(exponentials), _ptr(action), n, cells, _stream()))
    logp, ent = head_eval_fwd(logits, mask, action)
    return action, logp, (ent if want_entropy else None)


def unpack_rollout(mask_bits: Optional[torch.Tensor], action_u8: Optional[torch.Tensor], cells: int,
                   mask_out: Optional[torch.Tensor] = None, action_out: Optional[torch.Tensor] = None):
    """Device-side expansion of the rollout store's wire format: bit-packed masks
    ``[N, ceil(78C/8)]`` -> ``[N,78C]`` u8 and uint8 actions ``[N,7C]`` -> i64 (the dtypes of
    reference ``libs/utils.py:43-44``).  Either input may be ``None``."""
    dev = _need_cuda(mask_bits=mask_bits, action_u8=action_u8, mask_out=mask_out, action_out=action_out)
    n = (mask_bits if mask_bits is not None else action_u8).shape[0]
    if mask_bits is not None:
        mask_bits = _as(mask_bits, torch.uint8, "mask_bits")
        if mask_bits.shape[1] != (cells * LOGITS_PER_CELL + 7) // 8:
            raise ValueError("mask_bits must be [N, ceil(78*cells/8)]")
        if mask_out is None:
            mask_out = torch.empty((n, cells * LOGITS_PER_CELL), dtype=torch.uint8, device=dev)
    if action_u8 is not None:
        action_u8 = _as(action_u8, torch.uint8, "action_u8")
        if tuple(action_u8.shape) != (n, cells * COMPONENTS_PER_CELL):
            raise ValueError("action_u8 must be [N, 7*cells]")
        if action_out is None:
            action_out = torch.empty((n, cells * COMPONENTS_PER_CELL), dtype=torch.int64, device=dev)
    with torch.cuda.device(dev), _Launch("unpack_rollout"):
        _check(lib.mb_unpack_rollout(_ptr(mask_bits), _ptr(action_u8), _ptr(mask_out if mask_bits is not None else None),
                                     _ptr(action_out if action_u8 is not None else None), n, cells, _stream()))
    return mask_out, action_out


def unpack_obs(obs_bits: torch.Tensor, planes: int = 27, out: Optional[torch.Tensor] = None) -> torch.Tensor:
    """Bit-packed observations ``[N,h,w]`` int32 (bit c = plane c) -> ``[N,planes,h,w]`` fp32 NCHW,
    the layout the encoder's convolutions read."""
    dev = _need_cuda(obs_bits=obs_bits, out=out)
    obs_bits = _as(obs_bits, torch.int32, "obs_bits")
    if obs_bits.dim() != 3:
        raise ValueError("obs_bits must be [N, h, w]")
    n, h, w = obs_bits.shape
    if out is None:
        out = torch.empty((n, planes, h, w), dtype=torch.float32, device=dev)
    with torch.cuda.device(dev), _Launch("unpack_obs"):
        _check(lib.mb_unpack_obs(_ptr(obs_bits), _ptr(out), n, h * w, planes, _stream()))
    return out


def active_cells(action_mask: torch.Tensor, cells: int) -> torch.Tensor:
    """``[N, cells]`` bool: cells of ``action_mask [N, 78*cells]`` with at least one valid action."""
    dev = _need_cuda(action_mask=action_mask)
    action_mask = _mask_u8(action_mask)
    n = action_mask.shape[0]
    if action_mask.dim() != 2 or action_mask.shape[1] != cells * LOGITS_PER_CELL:
        raise ValueError("action_mask must be [N, 78*cells]")
    out = torch.empty((n, cells), dtype=torch.uint8, device=dev)
    with torch.cuda.device(dev), _Launch("active_cells", note={"cells": n * cells}):
        _check(lib.mb_active_cells(_ptr(action_mask), _ptr(out), n * cells, _stream()))
    return out.view(torch.bool)


def conv3x3_wgrad(x: torch.Tensor, dy: torch.Tensor, relu_x: bool = False):
    """``(dW [cout, cin, 3, 3], dbias [cout])`` of a 3x3 / stride 1 / pad 1 convolution from its input
    ``x [n,cin,h,w]`` and output gradient ``dy [n,cout,h,w]`` (fp32, NCHW contiguous), or ``None``
    when the shape is not one ``csrc/convgrad.cu`` instantiates (the caller then uses cuDNN).  ``relu_x``: the
    convolution was applied to ``relu(x)`` (``x`` is rectified as the kernel reads it)."""
    dev = _need_cuda(x=x, dy=dy)
    if x.dtype != torch.float32 or dy.dtype != torch.float32 or x.dim() != 4 or dy.dim() != 4:
        return None
    n, cin, h, w = x.shape
    cout = dy.shape[1]
    if dy.shape[0] != n or tuple(dy.shape[2:]) != (h, w):
        raise ValueError("conv3x3_wgrad: x and dy disagree on batch or map size")
    with torch.cuda.device(dev):
        blocks = lib.mb_conv3x3_wgrad_blocks(n, cin, cout, h, w)
        if blocks < 0:
            raise RuntimeError(_lib.last_error())
        if blocks == 0 or n == 0:
            return None
        x, dy = x.contiguous(), dy.contiguous()
        if (x.data_ptr() | dy.data_ptr()) & 15:
            return None
        if DETERMINISTIC_WGRAD:
            partials = torch.empty((blocks, cout * cin * 9 + cout), dtype=torch.float32, device=dev)
            with _Launch(f"conv3x3_wgrad_{cin}x{cout}x{h}", note={"n": n, "cin": cin, "cout": cout, "h": h}):
                _check(lib.mb_conv3x3_wgrad_relu(_ptr(x), _ptr(dy), _ptr(partials), blocks, n, cin, cout, h, w,
                                                 1 if relu_x else 0, None, _stream()))
            total = partials.sum(dim=0)           # fixed order: deterministic
        else:
            total = torch.zeros(cout * cin * 9 + cout, dtype=torch.float32, device=dev)
            with _Launch(f"conv3x3_wgrad_{cin}x{cout}x{h}", note={"n": n, "cin": cin, "cout": cout, "h": h}):
                _check(lib.mb_conv3x3_wgrad_relu(_ptr(x), _ptr(dy), None, blocks, n, cin, cout, h, w,
                                                 1 if relu_x else 0, _ptr(total), _stream()))
    return total[: cout * cin * 9].view(cout, cin, 3, 3), total[cout * cin * 9:]


def conv3x3_fwd(x: torch.Tensor, wk: torch.Tensor, bias: Optional[torch.Tensor], wmode: int = 0, relu_in: bool = False,
                mask_src: Optional[torch.Tensor] = None, add_src: Optional[torch.Tensor] = None) -> Optional[torch.Tensor]:
    """``out [n,cout,h,w]`` of a 3x3 / stride 1 / pad 1 convolution of ``x [n,cin,h,w]`` (``mb_conv3x3_fwd``), or
    ``None`` for a shape that is not covered.  ``wmode`` 0: ``wk [cin, cout, 9]``; 1: ``wk`` is a torch Conv2d
    weight ``[cout, cin, 3, 3]`` (forward); 2: ``x`` is an output gradient and ``wk`` the torch weight
    ``[cin, cout, 3, 3]`` of the convolution it came from -- the result is that convolution's input gradient.
    Fused element-wise neighbours (``mb_conv3x3_fwd_fused``): ``relu_in`` convolves ``relu(x)``; ``mask_src`` (shaped like
    the output) zeroes the output where ``mask_src <= 0``; ``add_src`` (shaped like the output) is added last."""
    dev = _need_cuda(x=x, wk=wk, bias=bias)
    if x.dtype != torch.float32 or x.dim() != 4 or wk.dim() not in (3, 4):
        return None
    n, cin, h, w = x.shape
    cout = wk.shape[0] if wmode == 1 else wk.shape[1]
    if wk.numel() != cin * cout * 9 or (wk.shape[1] if wmode == 1 else wk.shape[0]) != cin or \
            not lib.mb_conv3x3_fwd_supported(cin, cout, h, w):
        return None
    x = x.contiguous()
    wk = _as(wk, torch.float32, "wk")
    if bias is not None:
        bias = _as(bias, torch.float32, "bias")
    if x.data_ptr() & 15:
        return None
    out = torch.empty((n, cout, h, w), dtype=torch.float32, device=dev)
    for name, t in (("mask_src", mask_src), ("add_src", add_src)):
        if t is not None and (tuple(t.shape) != tuple(out.shape) or t.dtype != torch.float32 or not t.is_contiguous() or not t.is_cuda):
            raise ValueError(f"conv3x3_fwd: `{name}` must be a contiguous float32 CUDA tensor shaped like the output")
    with torch.cuda.device(dev), _Launch(f"conv3x3_fwd_{cin}x{cout}x{h}", note={"n": n, "cin": cin, "cout": cout, "h": h}):
        _check(lib.mb_conv3x3_fwd_fused(_ptr(x), _ptr(wk), _ptr(bias), _ptr(out), n, cin, cout, h, w, int(wmode),
                                        1 if relu_in else 0, _ptr(mask_src), _ptr(add_src), _stream()))
    return out


def conv1_bits_fwd(obs_bits: torch.Tensor, weight: torch.Tensor, bias: Optional[torch.Tensor]) -> Optional[torch.Tensor]:
    """The encoder's first convolution from bit-packed observations (``mb_conv1_bits_fwd``): ``obs_bits [N,h,w]`` int32,
    ``weight [16,27,3,3]`` -> ``[N,16,h,w]`` fp32; ``None`` when the shape is not covered."""
    dev = _need_cuda(obs_bits=obs_bits, weight=weight, bias=bias)
    if obs_bits.dtype != torch.int32 or obs_bits.dim() != 3 or tuple(weight.shape) != (16, 27, 3, 3):
        return None
    n, h, w = obs_bits.shape
    if h != w or h not in (8, 16) or n == 0:
        return None
    obs_bits = obs_bits.contiguous()
    weight = _as(weight, torch.float32, "weight")
    if bias is not None:
        bias = _as(bias, torch.float32, "bias")
    out = torch.empty((n, 16, h, w), dtype=torch.float32, device=dev)
    with torch.cuda.device(dev), _Launch(f"conv1_bits_fwd_{h}", note={"n": n, "h": h}):
        _check(lib.mb_conv1_bits_fwd(_ptr(obs_bits), _ptr(weight), _ptr(bias), _ptr(out), n, 27, 16, h, w, _stream()))
    return out


def conv1_bits_wgrad(obs_bits: torch.Tensor, dy: torch.Tensor) -> Optional[Tuple[torch.Tensor, torch.Tensor]]:
    """``(dW [16,27,3,3], db [16])`` of that first convolution from the bits (``mb_conv1_bits_wgrad``: work in
    proportion to the set -- or, for mostly-set planes, the unset -- cells); ``None`` when the shape is not covered."""
    dev = _need_cuda(obs_bits=obs_bits, dy=dy)
    if obs_bits.dtype != torch.int32 or obs_bits.dim() != 3 or dy.dim() != 4 or dy.shape[1] != 16:
        return None
    n, h, w = obs_bits.shape
    if h != w or h not in (8, 16) or n == 0 or tuple(dy.shape) != (n, 16, h, w):
        return None
    obs_bits = obs_bits.contiguous()
    dy = _as(dy, torch.float32, "dy")
    total = torch.zeros(16 * 27 * 9 + 16, dtype=torch.float32, device=dev)
    with torch.cuda.device(dev), _Launch(f"conv1_bits_wgrad_{h}", note={"n": n, "h": h}):
        _check(lib.mb_conv1_bits_wgrad(_ptr(obs_bits), _ptr(dy), _ptr(total), n, 27, 16, h, w, _stream()))
    return total[:16 * 27 * 9].view(16, 27, 3, 3), total[16 * 27 * 9:]


def split_tf32(x: torch.Tensor) -> Tuple[torch.Tensor, torch.Tensor]:
    """``(hi, lo)``: ``hi`` = ``x`` rounded to TF32, ``lo`` = the remainder rounded to TF32 (``x = hi + lo`` up
    to 2^-22 |x|) -- the operand form ``linear_tf32x3`` takes."""
    dev = _need_cuda(x=x)
    x = _as(x, torch.float32, "x")
    hi, lo = torch.empty_like(x), torch.empty_like(x)
    with torch.cuda.device(dev), _Launch("split_tf32"):
        _check(lib.mb_split_tf32(_ptr(x), _ptr(hi), _ptr(lo), x.numel(), _stream()))
    return hi, lo


def linear_tf32x3(x: torch.Tensor, weight: torch.Tensor, bias: Optional[torch.Tensor]) -> Optional[torch.Tensor]:
    """``x @ weight.T + bias`` (``x [M,K]``, ``weight [N,K]``) on the tcgen05 tensor cores as three TF32 products
    of hi/lo-split operands with fp32 accumulation (``mb_linear_tf32x3``): fp32-level accuracy.  ``None`` when
    the shape is not covered (``K % 32``, ``N % 4``)."""
    dev = _need_cuda(x=x, weight=weight, bias=bias)
    if x.dtype != torch.float32 or weight.dtype != torch.float32 or x.dim() != 2 or weight.dim() != 2:
        return None
    m, k = x.shape
    n = weight.shape[0]
    if weight.shape[1] != k or k % 32 or n % 4 or m == 0:
        return None
    a_hi, a_lo = split_tf32(x)
    b_hi, b_lo = split_tf32(weight)
    if bias is not None:
        bias = _as(bias, torch.float32, "bias")
    out = torch.empty((m, n), dtype=torch.float32, device=dev)
    with torch.cuda.device(dev), _Launch("linear_tf32x3", note={"m": m, "n": n, "k": k}):
        _check(lib.mb_linear_tf32x3(_ptr(a_hi), _ptr(a_lo), _ptr(b_hi), _ptr(b_lo), _ptr(bias), _ptr(out), m, n, k,
                                    _ptr(status_word(dev)), _stream()))
    return out


def pack_cells(action_mask: torch.Tensor, action: torch.Tensor, cells: int, out: Optional[torch.Tensor] = None,
               out_alive: Optional[torch.Tensor] = None) -> torch.Tensor:
    """``cell_meta [cells, N, 4]`` int32: one 16-byte record per (cell, frame) -- 78 mask bits and the seven action
    indices as 6-bit fields -- the operand form of the fused actor-head kernels (``mb_pack_cells``).  ``out`` /
    ``out_alive`` (``[cells, N, 4]`` / ``[cells, ceil(N/32)]`` int32) let a caller keep the records in buffers of its
    own (``learner.BatchPrefetcher`` packs the next batch on its side stream)."""
    dev = _need_cuda(action_mask=action_mask, action=action)
    action_mask = _mask_u8(action_mask)
    action = _as(action, torch.int64, "action")
    n = action_mask.shape[0]
    if tuple(action_mask.shape) != (n, cells * LOGITS_PER_CELL) or tuple(action.shape) != (n, cells * COMPONENTS_PER_CELL):
        raise ValueError("pack_cells: action_mask must be [N, 78*cells] and action [N, 7*cells]")
    meta = out if out is not None else torch.empty((cells, n, 4), dtype=torch.int32, device=dev)
    alive = out_alive if out_alive is not None else torch.empty((cells, (n + 31) // 32), dtype=torch.int32, device=dev)
    if tuple(meta.shape) != (cells, n, 4) or tuple(alive.shape) != (cells, (n + 31) // 32) or meta.dtype != torch.int32 or \
            alive.dtype != torch.int32 or not (meta.is_contiguous() and alive.is_contiguous()) or meta.device != dev or alive.device != dev:
        raise ValueError("pack_cells: `out` must be a contiguous int32 [cells, N, 4] and `out_alive` [cells, ceil(N/32)] on the inputs' device")
    with torch.cuda.device(dev), _Launch("pack_cells", note={"cells": n * cells}):
        _check(lib.mb_pack_cells(_ptr(action_mask), _ptr(action), _ptr(meta), _ptr(alive), n, cells,
                                 _ptr(status_word(dev)), _stream()))
    meta.alive_bits = alive          # [cells, ceil(N/32)]: bit i of word w = frame 32w+i has a valid action (mb_cells_bwd)
    return meta


def pack_actor_weight(weight: torch.Tensor, bias: Optional[torch.Tensor], cells: int):
    """``(w_hi, w_lo [80*cells, K], bias_p [80*cells])``: the actor Linear's weight as TF32 hi / lo parts with each
    cell's 78 rows padded to 80 (``mb_pack_actor_weight``)."""
    dev = _need_cuda(weight=weight, bias=bias)
    weight = _as(weight, torch.float32, "weight")
    if weight.dim() != 2 or weight.shape[0] != cells * LOGITS_PER_CELL:
        raise ValueError("pack_actor_weight: weight must be [78*cells, K]")
    k = weight.shape[1]
    if bias is not None:
        bias = _as(bias, torch.float32, "bias")
    w_hi = torch.empty((cells * 80, k), dtype=torch.float32, device=dev)
    w_lo = torch.empty_like(w_hi)
    bias_p = torch.empty(cells * 80, dtype=torch.float32, device=dev)
    with torch.cuda.device(dev), _Launch("pack_actor_weight"):
        _check(lib.mb_pack_actor_weight(_ptr(weight), _ptr(bias), _ptr(w_hi), _ptr(w_lo), _ptr(bias_p), cells, k, _stream()))
    return w_hi, w_lo, bias_p


def actor_head_fwd(hidden: torch.Tensor, packed_weight, cell_meta: torch.Tensor, want_logits=False):
    """Actor Linear + policy head in one tcgen05 kernel (``mb_actor_head_fwd``, reference ``model.py:167-214``).

    ``hidden [N, K]`` fp32, ``packed_weight`` from :func:`pack_actor_weight`, ``cell_meta`` from :func:`pack_cells`.
    ``want_logits``: False, True (``policy_logits`` of every cell) or ``"alive"`` (only cells with a valid action are
    written into an otherwise uninitialised buffer: what a backward pass needs).
    Returns ``(logprobs [N], entropy [N], cell_out [cells, N, 2], policy_logits [N, 78*cells] or None)``; ``None`` when
    the shape is not covered (``K % 32``)."""
    dev = _need_cuda(hidden=hidden, cell_meta=cell_meta)
    w_hi, w_lo, bias_p = packed_weight
    hidden = _as(hidden, torch.float32, "hidden")
    n, k = hidden.shape
    cells = cell_meta.shape[0]
    if k % 32 or w_hi.shape != (cells * 80, k) or tuple(cell_meta.shape) != (cells, n, 4) or n == 0:
        return None
    a_hi, a_lo = split_tf32(hidden)
    cell_out = torch.empty((cells, n, 2), dtype=torch.float32, device=dev)
    logp = torch.empty(n, dtype=torch.float32, device=dev)
    ent = torch.empty(n, dtype=torch.float32, device=dev)
    logits = torch.empty((n, cells * LOGITS_PER_CELL), dtype=torch.float32, device=dev) if want_logits else None
    with torch.cuda.device(dev), _Launch("actor_head_fwd", note={"m": n, "cells": cells, "k": k, "logits": str(want_logits)}):
        _check(lib.mb_actor_head_fwd(_ptr(a_hi), _ptr(a_lo), _ptr(w_hi), _ptr(w_lo), _ptr(bias_p), _ptr(cell_meta),
                                     _ptr(cell_out), _ptr(logits), 1 if want_logits == "alive" else 0, _ptr(logp),
                                     _ptr(ent), n, cells, k,
                                     _ptr(status_word(dev)), _stream()))
    return logp, ent, cell_out, logits


_cells_ws: Dict[Tuple[int, int], torch.Tensor] = {}


def actor_head_list_fwd(hidden: torch.Tensor, packed_weight, cell_meta: torch.Tensor, want_logits=False):
    """:func:`actor_head_fwd` over the (frame, cell) pairs that have a valid action only (``mb_actor_head_list_fwd``):
    cell-ordered lists built on the device, 128 gathered rows of ``hidden`` per tile against one cell's weight rows.
    Same outputs, bit for bit; ``want_logits``: False or ``"alive"`` (logits of the live cells into an otherwise
    uninitialised ``[N, 78*cells]`` buffer).  ``None`` when the shape is not covered."""
    dev = _need_cuda(hidden=hidden, cell_meta=cell_meta)
    w_hi, w_lo, bias_p = packed_weight
    hidden = _as(hidden, torch.float32, "hidden")
    n, k = hidden.shape
    cells = cell_meta.shape[0]
    alive_bits = getattr(cell_meta, "alive_bits", None)
    if k % 32 or w_hi.shape != (cells * 80, k) or tuple(cell_meta.shape) != (cells, n, 4) or n == 0 or cells > 1024 or \
            alive_bits is None or want_logits is True:
        return None
    a_hi, a_lo = split_tf32(hidden)
    need = int(lib.mb_actor_head_list_workspace_bytes(n, cells))
    idx = dev.index if dev.index is not None else torch.cuda.current_device()
    ws = _cells_ws.get((idx, 1))
    if ws is None or ws.numel() < need:
        if ws is not None:
            _retired_workspaces.append(ws)          # a captured graph may still address it
        ws = torch.empty(need, dtype=torch.uint8, device=dev)
        _cells_ws[(idx, 1)] = ws
    cell_out = torch.empty((cells, n, 2), dtype=torch.float32, device=dev)
    logp = torch.empty(n, dtype=torch.float32, device=dev)
    ent = torch.empty(n, dtype=torch.float32, device=dev)
    logits = torch.empty((n, cells * LOGITS_PER_CELL), dtype=torch.float32, device=dev) if want_logits else None
    with torch.cuda.device(dev), _Launch("actor_head_list_fwd", note={"m": n, "cells": cells, "k": k}):
        _check(lib.mb_actor_head_list_fwd(_ptr(a_hi), _ptr(a_lo), _ptr(w_hi), _ptr(w_lo), _ptr(bias_p), _ptr(cell_meta),
                                          _ptr(alive_bits), _ptr(cell_out), _ptr(logits), _ptr(logp), _ptr(ent), _ptr(ws),
                                          ws.numel(), n, cells, k, _ptr(status_word(dev)), _stream()))
    return logp, ent, cell_out, logits
# True: every CTA of mb_conv3x3_wgrad writes a partial result and torch sums them in a fixed order (one more pass per
# convolution); False: the CTAs add into the result with atomics (the cell-list backward of the actor head uses atomics
# anyway, so a learner step is not bit-reproducible from run to run either way)
DETERMINISTIC_WGRAD = False


def cells_bwd(hidden: torch.Tensor, weight: torch.Tensor, logits: torch.Tensor, cell_meta: torch.Tensor,
              g_logp: torch.Tensor, g_ent: torch.Tensor, alive_bits: Optional[torch.Tensor] = None):
    """``(d_hidden [N,256], d_weight [78C,256], d_bias [78C])`` of actor Linear + policy head over the cells with a valid
    action, entirely on the device and without a host synchronisation (``mb_cells_bwd``).  ``logits`` needs valid
    values only in those cells (``actor_head_fwd(..., want_logits="alive")``)."""
    dev = _need_cuda(hidden=hidden, weight=weight, logits=logits, cell_meta=cell_meta, g_logp=g_logp, g_ent=g_ent)
    hidden = _as(hidden, torch.float32, "hidden")
    weight = _as(weight, torch.float32, "weight")
    logits = _as(logits, torch.float32, "logits")
    g_logp = _as(g_logp, torch.float32, "g_logp")
    g_ent = _as(g_ent, torch.float32, "g_ent")
    n, k = hidden.shape
    cells = cell_meta.shape[0]
    if alive_bits is None:
        alive_bits = getattr(cell_meta, "alive_bits", None)
    if alive_bits is None or tuple(alive_bits.shape) != (cells, (n + 31) // 32):
        raise ValueError("cells_bwd: alive_bits [cells, ceil(N/32)] int32 is required (ops.pack_cells attaches it to cell_meta)")
    if k != 256 or tuple(weight.shape) != (cells * LOGITS_PER_CELL, k) or tuple(logits.shape) != (n, cells * LOGITS_PER_CELL) \
            or tuple(cell_meta.shape) != (cells, n, 4) or g_logp.numel() != n or g_ent.numel() != n:
        raise ValueError("cells_bwd: shape mismatch (hidden [N,256], weight [78C,256], logits [N,78C], cell_meta [C,N,4])")
    need = int(lib.mb_cells_bwd_workspace_bytes(n, cells))
    idx = dev.index if dev.index is not None else torch.cuda.current_device()
    ws = _cells_ws.get((idx, 0))
    if ws is None or ws.numel() < need:
        if ws is not None:
            _retired_workspaces.append(ws)          # a captured graph may still address it
        ws = torch.empty(need, dtype=torch.uint8, device=dev)
        _cells_ws[(idx, 0)] = ws
    d_hidden = torch.empty_like(hidden)
    d_weight = torch.empty_like(weight)
    d_bias = torch.empty(cells * LOGITS_PER_CELL, dtype=torch.float32, device=dev)
    with torch.cuda.device(dev), _Launch("cells_bwd", note={"n": n, "cells": cells}):
        _check(lib.mb_cells_bwd(_ptr(hidden), _ptr(weight), _ptr(logits), _ptr(cell_meta), _ptr(alive_bits), _ptr(g_logp), _ptr(g_ent),
                                _ptr(d_hidden), _ptr(d_weight), _ptr(d_bias), _ptr(ws), ws.numel(), n, cells, k, _stream()))
    return d_hidden, d_weight, d_bias


def maxpool3s2_bwd(dy: torch.Tensor, idx: torch.Tensor, h: int, w: int) -> torch.Tensor:
    """``dx [n,c,h,w]`` of ``max_pool2d(x, 3, stride=2, padding=1)`` from ``dy`` and the forward's argmax
    indices (both ``[n,c,(h+1)//2,(w+1)//2]``)."""
    dev = _need_cuda(dy=dy, idx=idx)
    dy = _as(dy, torch.float32, "dy")
    idx = _as(idx, torch.int64, "idx")
    if dy.dim() != 4 or dy.shape != idx.shape or tuple(dy.shape[2:]) != ((h + 1) // 2, (w + 1) // 2):
        raise ValueError("maxpool3s2_bwd: dy / idx must be [n, c, (h+1)//2, (w+1)//2]")
    n, c = dy.shape[:2]
    dx = torch.empty((n, c, h, w), dtype=torch.float32, device=dev)
    with torch.cuda.device(dev), _Launch("maxpool3s2_bwd"):
        _check(lib.mb_maxpool3s2_bwd(_ptr(dy), _ptr(idx), _ptr(dx), n * c, h, w, _stream()))
    return dx


def maxpool3s2_fwd(x: torch.Tensor):
    """``(y, code)`` of ``max_pool2d(x, 3, stride=2, padding=1)``: ``code`` uint8 = position of the winner inside its 3x3
    window (``ky * 3 + kx``, ATen's first-maximum tie rule) -- ``mb_maxpool3s2_fwd``."""
    dev = _need_cuda(x=x)
    x = _as(x, torch.float32, "x")
    n, c, h, w = x.shape
    y = torch.empty((n, c, (h + 1) // 2, (w + 1) // 2), dtype=torch.float32, device=dev)
    code = torch.empty(y.shape, dtype=torch.uint8, device=dev)
    with torch.cuda.device(dev), _Launch("maxpool3s2_fwd"):
        _check(lib.mb_maxpool3s2_fwd(_ptr(x), _ptr(y), _ptr(code), n * c, h, w, _stream()))
    return y, code


def maxpool3s2_bwd_code(dy: torch.Tensor, code: torch.Tensor, h: int, w: int) -> torch.Tensor:
    """Backward of :func:`maxpool3s2_fwd` from its one-byte codes (``mb_maxpool3s2_bwd_code``)."""
    dev = _need_cuda(dy=dy, code=code)
    dy = _as(dy, torch.float32, "dy")
    code = _as(code, torch.uint8, "code")
    if dy.dim() != 4 or dy.shape != code.shape or tuple(dy.shape[2:]) != ((h + 1) // 2, (w + 1) // 2):
        raise ValueError("maxpool3s2_bwd_code: dy / code must be [n, c, (h+1)//2, (w+1)//2]")
    n, c = dy.shape[:2]
    dx = torch.empty((n, c, h, w), dtype=torch.float32, device=dev)
    with torch.cuda.device(dev), _Launch("maxpool3s2_bwd"):
        _check(lib.mb_maxpool3s2_bwd_code(_ptr(dy), _ptr(code), _ptr(dx), n * c, h, w, _stream()))
    return dx


def vtrace_workspace(device: torch.device, rows: int) -> torch.Tensor:
    """Zero-initialised scratch for ``mb_vtrace_loss`` (cached per device; the kernel keeps
    its ticket word at zero between calls).  Calls on different streams that may overlap
    must not share it -- pass their own ``workspace=``."""
    need = int(lib.mb_vtrace_workspace_bytes(int(rows)))
    idx = device.index if device.index is not None else torch.cuda.current_device()
    key = (idx, 0)
    ws = _workspaces.get(key)
    if ws is None or ws.numel() < need:
        if ws is not None:
            # an outgrown workspace may still be addressed by a captured CUDA graph (GraphedLearner records the
            # raw pointer): keep it alive for the life of the process instead of handing it back to the allocator
            _retired_workspaces.append(ws)
        ws = torch.zeros(max(need, 1 << 16), dtype=torch.uint8, device=torch.device("cuda", idx))
        _workspaces[key] = ws
    return ws


def vtrace_loss(lp_old, lp_new, values, bootstrap, rewards, done, entropy, *, gamma=0.99, rho_bar=1.0,
                c_bar=1.0, pg_rho_bar=1.0, entropy_coef=0.01, inv_n: Optional[float] = None,
                want_grads: bool = True, workspace: Optional[torch.Tensor] = None, time_major: bool = False):
    """V-trace + losses (+ gradients) -- reference ``libs/utils.py:277-329``.

    All inputs ``[R,T]`` (time contiguous, the reference's layout) or, with ``time_major=True``,
    ``[T,R]`` (rows contiguous, the layout the rollout buffers are written in; outputs are then
    ``[T,R]`` too); ``bootstrap [R]``.  Returns a dict with
    ``vs``, ``pg_advantages`` ``[R,T]``, ``loss4`` ``[4]`` = pg, value, entropy, total and,
    with ``want_grads``, ``d_logp``, ``d_values`` ``[R,T]`` and the scalar ``d_entropy``.
    """
    dev = _need_cuda(lp_old=lp_old, lp_new=lp_new, values=values, bootstrap=bootstrap, rewards=rewards,
                     done=done, entropy=entropy, workspace=workspace)
    lp_old = _as(lp_old, torch.float32, "lp_old")
    lp_new = _as(lp_new, torch.float32, "lp_new")
    values = _as(values, torch.float32, "values")
    rewards = _as(rewards, torch.float32, "rewards")
    entropy = _as(entropy, torch.float32, "entropy")
    bootstrap = _as(bootstrap, torch.float32, "bootstrap")
    if done.dtype == torch.bool:
        done = done.view(torch.uint8) if done.is_contiguous() else done.contiguous().view(torch.uint8)
    done = _as(done, torch.uint8, "done")
    if values.dim() != 2:
        raise ValueError("values must be [R, T] (or [T, R] with time_major)")
    rows, T = (values.shape[1], values.shape[0]) if time_major else values.shape
    shape = tuple(values.shape)
    for name, t in (("lp_old", lp_old), ("lp_new", lp_new), ("rewards", rewards), ("done", done),
                    ("entropy", entropy)):
        if tuple(t.shape) != shape:
            raise ValueError(f"`{name}` has shape {tuple(t.shape)}, expected {shape}")
    if bootstrap.numel() != rows:
        raise ValueError("bootstrap must have one entry per row")
    if inv_n is None:
        inv_n = 1.0 / float(max(rows * T, 1))
    vs = torch.empty(shape, dtype=torch.float32, device=dev)
    adv = torch.empty(shape, dtype=torch.float32, device=dev)
    loss4 = torch.empty(4, dtype=torch.float32, device=dev)
    dlp = torch.empty(shape, dtype=torch.float32, device=dev) if want_grads else None
    dv = torch.empty(shape, dtype=torch.float32, device=dev) if want_grads else None
    ws = workspace if workspace is not None else vtrace_workspace(dev, rows)
    entry, pitch = (lib.mb_vtrace_loss_tmajor, rows) if time_major else (lib.mb_vtrace_loss, T)
    with torch.cuda.device(dev), _Launch("vtrace_loss_tmajor" if time_major else "vtrace_loss"):
        _check(entry(_ptr(lp_old), _ptr(lp_new), _ptr(values), _ptr(bootstrap), _ptr(rewards),
                                  _ptr(done), _ptr(entropy), rows, T, pitch, pitch,
                                  float(gamma), float(rho_bar), float(c_bar), float(pg_rho_bar),
                                  float(entropy_coef), float(inv_n),
                                  _ptr(vs), _ptr(adv), _ptr(loss4), _ptr(dlp), _ptr(dv),
                                  _ptr(ws), ws.numel(), _stream()))
    out = dict(vs=vs, pg_advantages=adv, loss4=loss4)
    if want_grads:
        out.update(d_logp=dlp, d_values=dv, d_entropy=-float(entropy_coef) * float(inv_n))
    return out


class HeadEval(torch.autograd.Function):
    """``(logits, mask, action) -> (logprobs, entropy)`` with the hand-written backward."""

    @staticmethod
    def forward(ctx, logits, mask, action):
        logp, ent = head_eval_fwd(logits, mask, action)
        ctx.save_for_backward(logits, mask, action)
        return logp, ent

    @staticmethod
    def backward(ctx, g_logp, g_ent):
        logits, mask, action = ctx.saved_tensors
        n = logits.shape[0]
        if g_logp is None:
            g_logp = torch.zeros(n, dtype=torch.float32, device=logits.device)
        if g_ent is None:
            g_ent = torch.zeros(n, dtype=torch.float32, device=logits.device)
        return head_eval_bwd(logits, mask, action, g_logp.contiguous(), g_ent.contiguous()), None, None


class HeadEvalCells(torch.autograd.Function):
    """Cell-list form of :class:`HeadEval`: ``(logits [M,78], mask [M,78], action [M,7]) ->
    (logp_cell [M], ent_cell [M])``; ``M`` a multiple of 64 (pad with fully masked cells)."""

    @staticmethod
    def forward(ctx, logits, mask, action):
        dev = _need_cuda(logits=logits, mask=mask, action=action)
        logits = _as(logits, torch.float32, "logits")
        mask = _mask_u8(mask)
        action = _as(action, torch.int64, "action")
        m = logits.shape[0]
        if logits.shape[1] != LOGITS_PER_CELL or tuple(mask.shape) != (m, LOGITS_PER_CELL) or \
                tuple(action.shape) != (m, COMPONENTS_PER_CELL):
            raise ValueError("cell lists must be logits/mask [M,78] and action [M,7]")
        out4 = torch.empty((4, m), dtype=torch.float32, device=dev)
        with torch.cuda.device(dev), _Launch("head_eval_cells_fwd", note={"cells": m}):
            _check(lib.mb_head_eval_cells_fwd(_ptr(logits), _ptr(mask), _ptr(action), _ptr(out4), m,
                                              _ptr(status_word(dev)), _stream()))
        ctx.save_for_backward(logits, mask, action)
        return out4[0] + out4[2], out4[1] + out4[3]

    @staticmethod
    def backward(ctx, g_logp, g_ent):
        logits, mask, action = ctx.saved_tensors
        return head_eval_cells_bwd(logits, mask, action, g_logp, g_ent), None, None


def head_eval_cells_bwd(logits, mask, action, g_logp, g_ent) -> torch.Tensor:
    """``d loss / d logits [M,78]`` of a cell list from the per-CELL upstream gradients ``g_logp [M]``,
    ``g_ent [M]`` (the frame's gradients repeated for each of its cells); ``M`` a multiple of 64."""
    dev = _need_cuda(logits=logits, mask=mask, action=action)
    logits = _as(logits, torch.float32, "logits")
    mask = _mask_u8(mask)
    action = _as(action, torch.int64, "action")
    m = logits.shape[0]
    if logits.shape[1] != LOGITS_PER_CELL or tuple(mask.shape) != (m, LOGITS_PER_CELL) or \
            tuple(action.shape) != (m, COMPONENTS_PER_CELL):
        raise ValueError("cell lists must be logits/mask [M,78] and action [M,7]")
    g_logp = torch.zeros(m, device=dev) if g_logp is None else _as(g_logp, torch.float32, "g_logp")
    g_ent = torch.zeros(m, device=dev) if g_ent is None else _as(g_ent, torch.float32, "g_ent")
    out = torch.empty_like(logits)
    with torch.cuda.device(dev), _Launch("head_eval_cells_bwd", note={"cells": m}):
        _check(lib.mb_head_eval_cells_bwd(_ptr(logits), _ptr(mask), _ptr(action), _ptr(g_logp), _ptr(g_ent),
                                          _ptr(out), m, _ptr(status_word(dev)), _stream()))
    return out


class VTraceLoss(torch.autograd.Function):
    """Differentiable wrt ``lp_new``, ``values``, ``entropy``; everything V-trace is
    ``no_grad`` exactly as in the reference (``utils.py:289``).  Returns
    ``(loss4, vs, pg_advantages)``; ``loss4 = [pg, value, entropy, total]``."""

    @staticmethod
    def forward(ctx, lp_new, values, entropy, lp_old, bootstrap, rewards, done, gamma, rho_bar, c_bar,
                pg_rho_bar, entropy_coef, inv_n, time_major=False):
        out = vtrace_loss(lp_old, lp_new, values, bootstrap, rewards, done, entropy, gamma=gamma,
                          rho_bar=rho_bar, c_bar=c_bar, pg_rho_bar=pg_rho_bar, entropy_coef=entropy_coef,
                          inv_n=inv_n, want_grads=True, time_major=time_major)
        ctx.save_for_backward(out["d_logp"], out["d_values"])
        ctx.entropy_coef = float(entropy_coef)
        ctx.inv_n = float(inv_n) if inv_n is not None else 1.0 / float(max(values.numel(), 1))
        ctx.mark_non_differentiable(out["vs"], out["pg_advantages"])
        return out["loss4"], out["vs"], out["pg_advantages"]

    @staticmethod
    def backward(ctx, g_loss4, _g_vs, _g_adv):
        d_logp, d_values = ctx.saved_tensors
        # loss4 = [pg, value, entropy, total]; total = pg + value - coef * entropy
        g_pg = g_loss4[0] + g_loss4[3]
        g_v = g_loss4[1] + g_loss4[3]
        g_en = (g_loss4[2] - ctx.entropy_coef * g_loss4[3]) * ctx.inv_n
        return (d_logp * g_pg, d_values * g_v, g_en.expand_as(d_logp), None, None, None, None,
                None, None, None, None, None, None, None)
