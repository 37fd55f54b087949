"""Drop-in ``Agent`` / ``CategoricalMasked`` for reference ``model.py``.

The conv/linear encoder (reference ``model.py:56-137``) stays on PyTorch -- it is the
only dense contraction on the path and runs on cuDNN / cuBLAS.  The action head
(``model.py:165-216``: 7*C ``CategoricalMasked`` objects per call) is replaced by the
fused CUDA kernels of ``csrc/head.cu``.  Parameter names equal the reference's, so a
reference ``state_dict`` loads unchanged.
"""
from __future__ import annotations

from typing import Optional

import numpy as np
import torch
import torch.nn as nn
import torch.nn.functional as F

from . import ops

NVEC_CELL = (6, 4, 4, 4, 4, 7, 49)
LOGITS_PER_CELL = 78
COMPONENTS_PER_CELL = 7


def layer_init(layer, std=np.sqrt(2), bias_const=0.0):
    """Orthogonal weight / constant bias init -- reference ``model.py:24-27``."""
    torch.nn.init.orthogonal_(layer.weight, std)
    torch.nn.init.constant_(layer.bias, bias_const)
    return layer


class CategoricalMasked:
    """All ``7*C`` masked categoricals of a batch of frames as ONE object.

    The reference builds one ``CategoricalMasked(logits=, masks=)`` per component
    (``model.py:33-52,176,190``) and stacks / sums their ``log_prob`` and ``entropy``
    over the frame (``model.py:194-214``).  This class takes the un-split tensors --
    ``logits [N,78C]``, ``masks [N,78C]`` -- and returns the per-frame sums directly.
    Differentiable wrt ``logits``.
    """

    def __init__(self, probs=None, logits=None, validate_args=None, masks=None):
        if probs is not None:
            raise ValueError("CategoricalMasked: only `logits=` is supported (as used by the reference)")
        if logits is None or masks is None or len(masks) == 0:
            raise ValueError("CategoricalMasked: `logits` and `masks` are required")
        self.logits = logits
        self.masks = masks
        self.validate_args = validate_args
        self._cache = None          # (action, logp, ent)

    def _eval(self, action):
        if self._cache is None or self._cache[0] is not action:
            logp, ent = ops.HeadEval.apply(self.logits, self.masks, action)
            self._cache = (action, logp, ent)
        return self._cache

    def sample(self, uniforms: Optional[torch.Tensor] = None, exponentials: Optional[torch.Tensor] = None,
               mode: str = "invcdf") -> torch.Tensor:
        """``action [N,7C]``.  ``mode="invcdf"``: inverse-CDF from ``uniforms [N,7C]`` (``torch.rand`` when omitted).
        ``mode="multinomial"`` (or ``exponentials=`` given): the reference's own rule, ``torch.multinomial`` =
        ``argmax(probs / Exp(1))`` (``model.py:177``), from ``exponentials [N,78C]`` (torch's generator when omitted)."""
        if mode == "multinomial" or exponentials is not None:
            with torch.no_grad():
                action, logp, ent = ops.head_sample_exp(self.logits.detach(), self.masks, exponentials, want_entropy=True)
            self._cache = (action, logp, ent)
            return action
        if uniforms is None:
            n, width = self.logits.shape
            uniforms = torch.rand((n, width // LOGITS_PER_CELL * COMPONENTS_PER_CELL),
                                  device=self.logits.device)
        with torch.no_grad():
            action, logp, ent = ops.head_sample(self.logits.detach(), self.masks, uniforms)
        self._cache = (action, logp, ent)
        return action

    def log_prob(self, action: torch.Tensor) -> torch.Tensor:
        """Per-frame sum of the component log-probs (``model.py:194,199,214``)."""
        out = self._eval(action)[1]
        if self.validate_args:
            ops.check_status(self.logits.device)
        return out

    def entropy(self, action: Optional[torch.Tensor] = None) -> torch.Tensor:
        """Per-frame sum of the masked component entropies (``model.py:46-52,196,214``)."""
        if action is None:
            if self._cache is None:
                n, width = self.logits.shape
                action = torch.zeros((n, width // LOGITS_PER_CELL * COMPONENTS_PER_CELL),
                                     dtype=torch.int64, device=self.logits.device)
            else:
                action = self._cache[0]
        return self._eval(action)[2]


class _Conv3x3(torch.autograd.Function):
    """3x3 / stride 1 / pad 1 convolution whose backward-data pass is itself a forward convolution
    (with the kernel flipped and its channel axes swapped).  Same mathematics as autograd's
    ``conv2d`` backward; the point is which cuDNN kernel runs: for these small-channel fp32 shapes
    cuDNN's data-gradient kernel (``dgrad2d_grouped_direct``) is ~3x slower than its forward
    implicit-GEMM kernel (``profiles/r01_step_launches_c2.csv``), and it was 37 % of the step."""

    @staticmethod
    def forward(ctx, x, weight, bias):
        ctx.save_for_backward(x, weight)
        if FAST_CONV_FORWARD:
            y = ops.conv3x3_fwd(x, weight, bias, wmode=1)          # the kernel reads the torch layout itself
            if y is not None:
                return y
        return F.conv2d(x, weight, bias, padding=1)

    @staticmethod
    def backward(ctx, g):
        x, weight = ctx.saved_tensors
        g = g.contiguous()
        gx = None
        if ctx.needs_input_grad[0]:
            cout, cin = weight.shape[:2]
            if FAST_CONV_FORWARD:           # the data gradient is a forward-type convolution with the kernel flipped
                gx = ops.conv3x3_fwd(g, weight, None, wmode=2)
            if gx is None:
                gx = F.conv2d(g, weight.flip(2, 3).transpose(0, 1).contiguous(), None, padding=1)
        both = ops.conv3x3_wgrad(x, g) if FAST_CONV_WGRAD else None    # csrc/convgrad.cu; None: shape not covered
        if both is not None:
            return gx, both[0], both[1]
        _, gw, gb = torch.ops.aten.convolution_backward(
            g, x, weight, [weight.shape[0]], [1, 1], [1, 1], [1, 1], False, [0, 0], 1, [False, True, True])
        return gx, gw, gb


class _Conv1Bits(torch.autograd.Function):
    """The encoder's first convolution (reference ``model.py:88,121``) evaluated straight from bit-packed observations
    (``mb_conv1_bits_fwd``: sums of weight columns over the set bits, no ``[N, 27, h, w]`` fp32 tensor in the forward
    pass); the weight and bias gradients come from the bits as well (``mb_conv1_bits_wgrad``).  Observations carry no
    gradient."""

    @staticmethod
    def forward(ctx, obs_bits, weight, bias):
        ctx.save_for_backward(obs_bits)
        ctx.planes = weight.shape[1]
        return ops.conv1_bits_fwd(obs_bits, weight, bias)

    @staticmethod
    def backward(ctx, g):
        (obs_bits,) = ctx.saved_tensors
        g = g.contiguous()
        both = ops.conv1_bits_wgrad(obs_bits, g) if FAST_CONV1_BITS_WGRAD else None
        if both is None:                                   # the planes as numbers, for the dense weight-gradient kernel
            both = ops.conv3x3_wgrad(ops.unpack_obs(obs_bits, planes=ctx.planes), g)
        return None, both[0], both[1]


def conv1_from_bits(conv: nn.Conv2d, obs_bits: torch.Tensor):
    """``conv(planes(obs_bits))`` through :class:`_Conv1Bits`, or ``None`` when that kernel does not cover the shape."""
    if not (FAST_CONV1_BITS and obs_bits.is_cuda and obs_bits.dtype == torch.int32 and obs_bits.dim() == 3):
        return None
    n, h, w = obs_bits.shape
    if tuple(conv.weight.shape) != (16, 27, 3, 3) or h != w or h not in (8, 16) or n == 0:
        return None
    return _Conv1Bits.apply(obs_bits, conv.weight, conv.bias)


class _MaxPool3s2(torch.autograd.Function):
    """``max_pool2d(x, 3, stride=2, padding=1)`` (reference model.py:96) with the backward as a gather over
    the saved argmax indices (``mb_maxpool3s2_bwd``) instead of ATen's ``max_pool_backward_nchw``."""

    @staticmethod
    def forward(ctx, x):
        ctx.hw = (x.shape[2], x.shape[3])
        if FAST_MAXPOOL_FORWARD:
            y, code = ops.maxpool3s2_fwd(x.contiguous())          # one-byte argmax instead of ATen's int64 index
            ctx.save_for_backward(code)
            ctx.coded = True
            return y
        y, idx = F.max_pool2d(x, kernel_size=3, stride=2, padding=1, return_indices=True)
        ctx.save_for_backward(idx)
        ctx.coded = False
        return y

    @staticmethod
    def backward(ctx, g):
        (idx,) = ctx.saved_tensors
        if ctx.coded:
            return ops.maxpool3s2_bwd_code(g.contiguous(), idx, *ctx.hw)
        return ops.maxpool3s2_bwd(g, idx, *ctx.hw)


def maxpool3s2(x: torch.Tensor) -> torch.Tensor:
    if x.is_cuda and FAST_CONV_BACKWARD and torch.is_grad_enabled() and x.dtype == torch.float32:
        return _MaxPool3s2.apply(x)
    return F.max_pool2d(x, kernel_size=3, stride=2, padding=1)


def conv3x3(conv: nn.Conv2d, x: torch.Tensor) -> torch.Tensor:
    """``conv(x)`` for the encoder's 3x3 convolutions; on CUDA it goes through :class:`_Conv3x3`."""
    if x.is_cuda and FAST_CONV_BACKWARD and torch.is_grad_enabled():
        return _Conv3x3.apply(x, conv.weight, conv.bias)
    return conv(x)


FAST_CONV_BACKWARD = True
# first convolution straight from the bit-packed observations of the compact rollout store (mb_conv1_bits_fwd)
FAST_CONV1_BITS = True
# ... and its weight gradient too (mb_conv1_bits_wgrad); False: mb_unpack_obs + mb_conv3x3_wgrad in the backward
FAST_CONV1_BITS_WGRAD = True
# max-pool forward by mb_maxpool3s2_fwd (value + one-byte argmax) instead of ATen's kernel with int64 indices
FAST_MAXPOOL_FORWARD = True
# weight gradients of the 3x3 convolutions by mb_conv3x3_wgrad (fp32 FMA, deterministic) instead of cuDNN's
# wgrad kernels, which were a third of the learner step (profiles/r01_encoder_layers.txt)
FAST_CONV_WGRAD = True
# forward convolutions and their data gradients by mb_conv3x3_fwd (fp32 FMA, register tiled) instead of cuDNN's
# implicit-GEMM / winograd kernels
FAST_CONV_FORWARD = True
# a whole residual block (relu-conv-relu-conv + skip, model.py:67-73) as one autograd node whose element-wise steps are
# folded into the convolution kernels (no ReLU / add / ReLU-backward passes over the activations)
FUSED_RES_BLOCK = True
# actor Linear forward (learner mode) by mb_linear_tf32x3: tcgen05 tensor cores, three TF32 products of hi/lo-split
# operands accumulated in fp32 (error 5.5e-7 of max|logit|; cuBLAS fp32: 5.0e-7) instead of a SIMT SGEMM
TENSOR_CORE_ACTOR = True
# learner mode: actor Linear and policy head as ONE tcgen05 kernel (mb_actor_head_fwd): the logits stay in tensor memory,
# only the cells that have a valid action are written out (for the backward pass); off: mb_linear_tf32x3 + mb_head_eval_fwd
# ... over the pairs with a valid action only when policy_logits is not asked for (mb_actor_head_list_fwd)
LIST_ACTOR_HEAD = True
FUSED_ACTOR_HEAD = True
# backward of the actor head by mb_cells_bwd (device-side cell lists, no host sync) whenever the fused forward ran and
# Agent.sparse_backward is not False; off: the round-1 torch cell-list path (nonzero + bmm + index_add_)
CELL_KERNEL_BACKWARD = True


def cell_list(action_mask: torch.Tensor, cells: int, auto: bool):
    """Active (frame, cell) pairs grouped by cell: ``(fidx [cells, lmax], cidx [cells, lmax])`` with
    ``fidx`` the frame index, padded with ``n`` (one past the last frame) up to the longest group
    rounded to 64.  ``None`` with ``auto`` when the padded list would be at least half the dense grid.
    Costs two host synchronisations (``nonzero`` and the longest group's length)."""
    return cell_list_from_active(ops.active_cells(action_mask, cells), auto)


def cell_list_from_active(active: torch.Tensor, auto: bool):
    """``cell_list`` from the ``[N, C]`` bool map of cells with a valid action (pure index bookkeeping; runs on
    any device, which is how ``tests/test_host_logic.py`` covers it without a GPU)."""
    n, cells = active.shape
    counts = active.sum(dim=0)                                         # active frames per cell
    pairs = active.t().nonzero()                                       # [P, 2] = (cell, frame), grouped by cell
    lmax = int(counts.max().item()) if pairs.shape[0] else 0
    lmax = max(-(-max(lmax, 1) // 64) * 64, 64)                        # groups of 64 cells for the kernel
    if auto and 2 * lmax > n:
        return None
    dev = active.device
    starts = torch.cumsum(counts, 0) - counts
    pos = torch.arange(pairs.shape[0], device=dev) - starts[pairs[:, 0]]
    fidx = torch.full((cells, lmax), n, dtype=torch.int64, device=dev)
    fidx[pairs[:, 0], pos] = pairs[:, 1]
    cidx = torch.arange(cells, device=dev).unsqueeze(1).expand(cells, lmax)
    return fidx, cidx


_PAD_ROWS = 1024
_SPARSE_BACKWARD_MIN_PAIRS = 1 << 20


def spread_padding(fidx: torch.Tensor, n: int) -> torch.Tensor:
    """Row indices for gathering / scatter-adding ``[n + _PAD_ROWS, width]`` tensors through a cell list: the
    padding entries (``fidx == n``) are spread over ``_PAD_ROWS`` dummy rows instead of all hitting row ``n`` --
    an ``index_add`` of ~10 % padding rows into ONE row serialises its atomics."""
    lmax = fidx.shape[1]
    dummy = n + (torch.arange(lmax, device=fidx.device) % _PAD_ROWS)
    return torch.where(fidx < n, fidx, dummy.unsqueeze(0).expand_as(fidx)).reshape(-1)


def gather_cells(fidx, cidx, n, action_mask, action, cells):
    """Mask / action rows of a cell list; padding entries get an all-zero mask (a dead cell: log-prob 0,
    entropy 0, zero gradient -- quirk Q1) and any in-range action."""
    valid = fidx < n
    safe = fidx.clamp(max=n - 1)
    mask3 = action_mask.view(n, cells, LOGITS_PER_CELL)
    mask_c = torch.where(valid.unsqueeze(2), mask3[safe, cidx], torch.zeros((), dtype=mask3.dtype, device=mask3.device))
    act_c = action.view(n, cells, COMPONENTS_PER_CELL)[safe, cidx]
    return safe, mask_c.reshape(-1, LOGITS_PER_CELL), act_c.reshape(-1, COMPONENTS_PER_CELL)


def _linear_bias_in_k(x: torch.Tensor, weight: torch.Tensor, bias: torch.Tensor) -> torch.Tensor:
    """``x @ weight.T + bias`` as ONE matmul: the bias rides along as an extra column of ``weight`` against a
    column of ones in ``x`` (K padded by 8 to keep rows 32-byte aligned).  For the [N, 78C] logits cuBLASLt
    otherwise adds the bias in a separate pass over the output, 1.4 ms of a 25 ms step at 16x16."""
    n, k = x.shape
    xa = x.new_zeros((n, k + 8))
    xa[:, :k] = x
    xa[:, k] = 1.0
    wa = weight.new_zeros((weight.shape[0], k + 8))
    wa[:, :k] = weight
    wa[:, k] = bias
    return xa @ wa.t()


class _ActorHead(torch.autograd.Function):
    """Actor ``Linear`` (model.py:136,167) + policy head with a DENSE forward -- ``policy_logits`` is an
    output, as in the reference -- and a backward restricted to the cells that can have a gradient.

    ``d loss / d logits`` is exactly zero on every cell without a valid action (quirk Q1), which is 94 %
    of the rows at microRTS densities, so the two backward GEMMs of the Linear (``dlogits^T @ hidden`` and
    ``dlogits @ W``, 7.5 ms of the 30 ms learner step at 16x16) multiply mostly zeros.  The backward
    builds the list of active (frame, cell) pairs, runs the cell-list head-backward kernel on it and does
    the two products as batched matmuls over the list: same fp32 arithmetic on the non-zero rows, only
    the summation order differs.  ``mode="auto"`` falls back to the dense kernels when the padded list
    would be at least half the grid; ``False`` always uses them (needed under CUDA-graph capture: the
    list is sized from the data)."""

    @staticmethod
    def forward(ctx, hidden, weight, bias, action_mask, action, cells, mode, want_logits=True, meta=None):
        fused = None
        if TENSOR_CORE_ACTOR and FUSED_ACTOR_HEAD and hidden.shape[1] % 32 == 0 and hidden.shape[0] > 0:
            # model.py:167-214 in one kernel; ``logits`` then holds the cells with a valid action only (enough for
            # the backward: every other cell has exactly zero gradient) unless the caller wants policy_logits
            if meta is None or tuple(meta.shape) != (cells, hidden.shape[0], 4) or not hasattr(meta, "alive_bits"):
                meta = ops.pack_cells(action_mask, action, cells)      # else: packed ahead by learner.BatchPrefetcher
            packed = ops.pack_actor_weight(weight, bias, cells)
            if LIST_ACTOR_HEAD and not want_logits:
                # nobody reads policy_logits: form the logits of the (frame, cell) pairs with a valid action only
                fused = ops.actor_head_list_fwd(hidden, packed, meta, want_logits="alive")
            if fused is None:
                fused = ops.actor_head_fwd(hidden, packed, meta, want_logits=True if want_logits else "alive")
        ctx.meta = None
        if fused is not None:
            logprobs, entropy, _cell_out, logits = fused
            ctx.meta = meta
            ctx.alive_bits = meta.alive_bits
        else:
            logits = ops.linear_tf32x3(hidden, weight, bias) if TENSOR_CORE_ACTOR else None    # model.py:167
            if logits is None:
                logits = _linear_bias_in_k(hidden, weight, bias)
            logprobs, entropy = ops.head_eval_fwd(logits, action_mask, action)
        ctx.save_for_backward(hidden, weight, logits, action_mask, action)
        ctx.cells, ctx.mode = cells, mode
        ctx.mark_non_differentiable(logits)
        # undefined output gradients arrive as None: autograd would otherwise zero-fill an [N, 78C] fp32 gradient for
        # `logits` before every backward (1.66 GB, 0.22 ms of the config-2 step)
        ctx.set_materialize_grads(False)
        return logits, logprobs, entropy

    @staticmethod
    def backward(ctx, _g_logits, g_lp, g_ent):
        hidden, weight, logits, action_mask, action = ctx.saved_tensors
        n, cells = hidden.shape[0], ctx.cells
        g_lp = torch.zeros(n, device=hidden.device) if g_lp is None else g_lp.contiguous()
        g_ent = torch.zeros(n, device=hidden.device) if g_ent is None else g_ent.contiguous()
        if ctx.meta is not None and ctx.mode and CELL_KERNEL_BACKWARD and hidden.shape[1] == 256 and cells <= 1024:
            # the device-side cell lists + the fp32 FMA products of csrc/cells.cu: no host synchronisation, capturable
            g_hidden, g_weight, g_bias = ops.cells_bwd(hidden, weight, logits, ctx.meta, g_lp, g_ent, ctx.alive_bits)
            return g_hidden, g_weight, g_bias, None, None, None, None, None, None
        # "auto": below ~1e6 (frame, cell) pairs the dense GEMMs cost less than the list's bookkeeping and its two
        # host synchronisations (8x8 maps are launch-bound, profiles/r01_bench_c1_n1.json)
        small = ctx.mode == "auto" and n * cells < _SPARSE_BACKWARD_MIN_PAIRS
        use_list = ctx.mode and ctx.mode != "kernel" and not small          # "kernel" without the kernel path: dense
        cl = cell_list(action_mask, cells, auto=(ctx.mode == "auto")) if use_list else None
        if cl is None:
            dlogits = ops.head_eval_bwd(logits, action_mask, action, g_lp, g_ent)
            return dlogits @ weight, dlogits.t() @ hidden, dlogits.sum(dim=0), None, None, None, None, None, None
        fidx, cidx = cl
        lmax, width = fidx.shape[1], hidden.shape[1]
        safe, mask_c, act_c = gather_cells(fidx, cidx, n, action_mask, action, cells)
        logits_c = logits.view(n, cells, LOGITS_PER_CELL)[safe, cidx].reshape(-1, LOGITS_PER_CELL)
        dl = ops.head_eval_cells_bwd(logits_c, mask_c, act_c, g_lp[safe].reshape(-1), g_ent[safe].reshape(-1))
        dl = dl.view(cells, lmax, LOGITS_PER_CELL)
        flat = spread_padding(fidx, n)
        hid = torch.cat([hidden, hidden.new_zeros(_PAD_ROWS, width)]).index_select(0, flat).view(cells, lmax, width)
        g_weight = torch.bmm(dl.transpose(1, 2), hid).reshape(cells * LOGITS_PER_CELL, width)
        g_bias = dl.sum(dim=1).reshape(-1)
        g_hid = torch.bmm(dl, weight.view(cells, LOGITS_PER_CELL, width))                   # [C, lmax, 256]
        g_hidden = hidden.new_zeros((n + _PAD_ROWS, width)).index_add_(0, flat, g_hid.reshape(-1, width))[:n]
        return g_hidden, g_weight, g_bias, None, None, None, None, None, None


class _ResBlock(torch.autograd.Function):
    """``conv1(relu(conv0(relu(x)))) + x`` (reference ``model.py:67-73``) in four kernel launches forward + backward
    pairs, with every element-wise step folded into a convolution kernel (``mb_conv3x3_fwd_fused`` /
    ``mb_conv3x3_wgrad_relu``): the ReLUs are applied as the convolutions read their input, the skip connection is
    added in conv1's epilogue, and in the backward the ReLU masks and the bypassing gradient are applied in the
    epilogues of the data-gradient convolutions.  Same fp32 arithmetic as the unfused composition (the element-wise
    steps are exact); only ``x`` and ``conv0``'s output are kept for the backward (no ReLU outputs)."""

    @staticmethod
    def forward(ctx, x, w0, b0, w1, b1):
        y0 = ops.conv3x3_fwd(x, w0, b0, wmode=1, relu_in=True)
        out = ops.conv3x3_fwd(y0, w1, b1, wmode=1, relu_in=True, add_src=x)
        ctx.save_for_backward(x, y0, w0, w1)
        return out

    @staticmethod
    def backward(ctx, g):
        x, y0, w0, w1 = ctx.saved_tensors
        g = g.contiguous()
        gw1, gb1 = ops.conv3x3_wgrad(y0, g, relu_x=True)
        g0 = ops.conv3x3_fwd(g, w1, None, wmode=2, mask_src=y0)                  # through conv1, then relu'(y0)
        gw0, gb0 = ops.conv3x3_wgrad(x, g0, relu_x=True)
        gx = ops.conv3x3_fwd(g0, w0, None, wmode=2, mask_src=x, add_src=g)       # through conv0, relu'(x), + skip
        return gx, gw0, gb0, gw1, gb1


def _res_block_fusable(x: torch.Tensor, channels: int) -> bool:
    if not (FUSED_RES_BLOCK and FAST_CONV_BACKWARD and FAST_CONV_FORWARD and FAST_CONV_WGRAD):
        return False
    if not (x.is_cuda and x.dtype == torch.float32 and x.dim() == 4 and torch.is_grad_enabled() and x.is_contiguous()):
        return False
    n, c, h, w = x.shape
    if c != channels or h != w or n == 0 or x.data_ptr() & 15:
        return False
    from ._lib import lib
    return bool(lib.mb_conv3x3_fwd_supported(c, c, h, w)) and lib.mb_conv3x3_wgrad_blocks(n, c, c, h, w) > 0


class Residual_Block(nn.Module):
    """relu-conv-relu-conv + skip -- reference ``model.py:56-73``."""

    def __init__(self, channels):
        super().__init__()
        self.channels = channels
        self.conv0 = nn.Conv2d(channels, channels, kernel_size=3, padding=1)
        self.conv1 = nn.Conv2d(channels, channels, kernel_size=3, padding=1)

    def forward(self, x):
        if _res_block_fusable(x, self.channels):
            return _ResBlock.apply(x, self.conv0.weight, self.conv0.bias, self.conv1.weight, self.conv1.bias)
        y = conv3x3(self.conv0, F.relu(x))
        y = conv3x3(self.conv1, F.relu(y))
        return y + x


class ConvSequence(nn.Module):
    """conv3x3 -> maxpool(3, s2, p1) -> 2 residual blocks -- reference ``model.py:77-109``."""

    def __init__(self, input_shape, out_channels):
        super().__init__()
        self._input_shape = tuple(input_shape)
        self._out_channels = out_channels
        self.conv = nn.Conv2d(self._input_shape[0], out_channels, kernel_size=3, padding=1)
        self.res_block0 = Residual_Block(out_channels)
        self.res_block1 = Residual_Block(out_channels)

    def get_output_shape(self):
        _c, h, w = self._input_shape
        return (self._out_channels, (h + 1) // 2, (w + 1) // 2)

    def forward(self, x):
        if x.dtype == torch.int32:          # bit-packed observations [N, h, w] (first stage only): see Agent._obs
            y = conv1_from_bits(self.conv, x)
            if y is None:
                y = conv3x3(self.conv, ops.unpack_obs(x, planes=self.conv.weight.shape[1]))
        else:
            y = conv3x3(self.conv, x)
        x = maxpool3s2(y)
        return self.res_block1(self.res_block0(x))


class Agent(nn.Module):
    """Reference ``model.py:112-220`` with the same constructor and ``get_action`` contract.

    ``nvec`` must be ``[6,4,4,4,4,7,49] * cells`` (``microbeast.py:145``), which is what the
    kernels are specialised for.
    """

    def __init__(self, obs_space_shape: tuple, nvec: list, mapsize=8 * 8, device="cpu"):
        super().__init__()
        h, w, c = obs_space_shape
        nvec = [int(v) for v in nvec]
        if len(nvec) % COMPONENTS_PER_CELL or nvec != list(NVEC_CELL) * (len(nvec) // COMPONENTS_PER_CELL):
            raise ValueError("Agent: nvec must be [6,4,4,4,4,7,49] repeated once per map cell")
        self.mapsize = mapsize
        self.nvec = nvec
        self.cells = len(nvec) // COMPONENTS_PER_CELL
        self.nchw_copy = True
        self.sparse_head = False           # True / "auto": no policy_logits for fully masked cells (see _sparse_head)
        self.sparse_backward = "auto"      # backward of the actor head over the active cells only (see _ActorHead)
        self.sampler = "invcdf"            # acting mode: "invcdf" (one uniform per component) | "multinomial" (torch's rule)
        self.skips_policy_logits = True    # get_action(..., policy_logits=False) is understood (learner.learner_losses)
        self.obs_planes = c
        shape = (c, h, w)
        layers = []
        for out_channels in (16, 32, 32):
            seq = ConvSequence(shape, out_channels)
            shape = seq.get_output_shape()
            layers.append(seq)
        layers += [nn.Flatten(), nn.ReLU(), nn.Linear(shape[0] * shape[1] * shape[2], 256), nn.ReLU()]
        self.network = nn.Sequential(*layers)
        self.actor = layer_init(nn.Linear(256, sum(nvec)), std=0.0)
        self.critic = layer_init(nn.Linear(256, 1), std=1)

    def initial_state(self, batch_size):
        return tuple()

    def _obs(self, input_dict, learning):
        obs = input_dict["obs"] if learning else input_dict["obs"][0, 0, ...]     # model.py:156-160
        if obs.dtype == torch.int32:     # bit-packed planes from the compact rollout store: [N,h,w]
            if FAST_CONV1_BITS and obs.is_cuda:
                return obs               # the first ConvSequence convolves the bits directly (or expands them itself)
            return ops.unpack_obs(obs, planes=self.obs_planes)
        if obs.dtype != torch.float32:
            obs = obs.float()            # uint8 wire format of the B200 rollout store
        x = obs.permute((0, 3, 1, 2))
        # fp32 cuDNN convolutions run ~12 % faster from a dense NCHW copy than from the
        # channels-last view on B200 (profiles/r01_encoder_formats.txt); values are identical
        return x.contiguous() if (x.is_cuda and self.nchw_copy) else x

    def _sparse_head(self, hidden, action_mask, action):
        """Learner-mode head on the ACTIVE cells only (opt-in: ``self.sparse_head = True``, or ``"auto"``
        to fall back to the dense path -- by returning None -- when at least half the grid is active).

        A cell whose 78 mask bytes are all zero contributes exactly 0 to the frame's log-prob and
        entropy and gets zero gradient whatever its logits are (quirk Q1), so the actor Linear is
        evaluated only for (frame, cell) pairs with a valid action: pairs are grouped by cell, each
        group padded to the longest one, ``logits = hidden[frames] @ W_cell^T + b_cell`` as one batched
        matmul, the cell list goes through the cell-list head kernels and per-cell results are
        scattered back to their frames.  Same values as the dense path up to fp32 summation order;
        ``policy_logits`` is not produced."""
        n, cells = hidden.shape[0], self.cells
        cl = cell_list(action_mask, cells, auto=(self.sparse_head == "auto"))
        if cl is None:
            return None                      # padded cell list >= half the dense grid: the dense path is cheaper
        fidx, cidx = cl
        lmax = fidx.shape[1]
        _, mask_c, act_c = gather_cells(fidx, cidx, n, action_mask, action, cells)
        # index_select: its backward is an index_add (atomics), not the sort-based index_put of x[idx]
        hid = torch.cat([hidden, hidden.new_zeros(_PAD_ROWS, hidden.shape[1])]).index_select(0, spread_padding(fidx, n))
        hid = hid.view(cells, lmax, -1)                    # [C, lmax, 256]
        w = self.actor.weight.view(cells, LOGITS_PER_CELL, -1)
        b = self.actor.bias.view(cells, 1, LOGITS_PER_CELL)
        logits_c = torch.baddbmm(b, hid, w.transpose(1, 2)).reshape(-1, LOGITS_PER_CELL)      # [C*lmax, 78]
        lp_c, en_c = ops.HeadEvalCells.apply(logits_c, mask_c, act_c)
        # back to frames: every (frame, cell) slot is written at most once, then a fixed-order row sum
        # (deterministic, unlike an atomic index_add); padding lands in the dropped row n
        grid = hidden.new_zeros((2, n + 1, cells))
        grid = grid.index_put((torch.arange(2, device=hidden.device).view(2, 1, 1), fidx, cidx),
                              torch.stack([lp_c, en_c]).view(2, cells, lmax))
        per_frame = grid[:, :n].sum(dim=2)
        return per_frame[0], per_frame[1]

    def forward(self, input_dict: dict, agent_state=(), learning=False, inds=()):
        return self.network(self._obs(input_dict, learning))

    def get_value(self, input_dict: dict, agent_state=(), learning=False, inds=()):
        return self.critic(self.forward(input_dict, agent_state, learning, inds))

    def get_action(self, input_dict: dict, learning=False, inds=(), agent_state=(),
                   uniforms: Optional[torch.Tensor] = None, policy_logits: bool = True,
                   exponentials: Optional[torch.Tensor] = None):
        """Same outputs as reference ``model.py:165-216``.

        acting:   ``dict(action [n,7C], policy_logits [n,78C], logprobs [n], baseline [1,n])``
        learning: the same plus ``entropy [N]``; ``action`` is the batch's action.
        One encoder pass feeds both heads (the reference runs the encoder twice, quirk
        Q6; values and summed gradients are identical).  ``policy_logits=False`` (learning mode; what
        ``learner_losses`` passes, since ``PPO_learn`` never reads them, libs/utils.py:237-329) returns ``None`` under
        that key and lets the fused actor-head kernel skip the [N, 78C] store.
        """
        hidden = self.forward(input_dict, agent_state, learning, inds)
        baseline = self.critic(hidden).view(1, -1)                              # model.py:205
        if not learning:
            logits = self.actor(hidden)                                         # model.py:167
            action_mask = input_dict["action_mask"][0, ...]                     # model.py:173
            if exponentials is not None or (self.sampler == "multinomial" and uniforms is None):
                # the reference's sampling rule (model.py:177): argmax(probs / Exp(1)), one draw per category
                action, logprobs, _ = ops.head_sample_exp(logits.detach(), action_mask, exponentials)
                return (dict(action=action, policy_logits=logits, logprobs=logprobs, baseline=baseline), ())
            if uniforms is None:
                uniforms = torch.rand((logits.shape[0], self.cells * COMPONENTS_PER_CELL),
                                      device=logits.device)
            action, logprobs, _ = ops.head_sample(logits.detach(), action_mask, uniforms,
                                                  want_entropy=False)
            return (dict(action=action, policy_logits=logits, logprobs=logprobs, baseline=baseline), ())
        ready = input_dict.get("_ready")      # set by learner.batch_to_device: side-stream H2D of mask / action
        if ready is not None:
            torch.cuda.current_stream().wait_event(ready)
        action = input_dict["action"]                                           # model.py:182
        action_mask = input_dict["action_mask"]                                 # model.py:188
        if self.sparse_head:                  # False (reference-shaped outputs) | True | "auto"
            sparse = self._sparse_head(hidden, action_mask, action)
            if sparse is not None:
                return (dict(action=action, policy_logits=None, logprobs=sparse[0], baseline=baseline,
                             entropy=sparse[1]), ())
        logits, logprobs, entropy = _ActorHead.apply(hidden, self.actor.weight, self.actor.bias, action_mask,
                                                     action, self.cells, self.sparse_backward, bool(policy_logits),
                                                     input_dict.get("_cell_meta"))
        return (dict(action=action, policy_logits=logits if policy_logits else None, logprobs=logprobs,
                     baseline=baseline, entropy=entropy), ())
