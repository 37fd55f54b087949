"""Data-parallel plumbing for the sharded learner.

The reference has no multi-device code.  The learner batch shards by rows: the policy
head is independent per frame and V-trace per row, so the only exchange step is the
gradient sum.  One process per GPU; every rank holds a model replica, normalises its
local loss sums by the GLOBAL ``rows*T`` (``inv_n`` of ``mb_vtrace_loss``) and the
ranks sum one flat fp32 gradient bucket (5.6 / 21.1 / 83 MB for 8x8 / 16x16 / 32x32
maps) with NCCL over NVLink; the four logged loss scalars ride in the same bucket.
"""
from __future__ import annotations

from typing import List, Optional, Tuple

import torch
import torch.distributed as dist


def bind_to_gpu_numa(device_index: int) -> Optional[List[int]]:
    """Restrict this process to the host cores next to GPU ``device_index`` (NVML's ideal CPU affinity).  With one
    process per GPU on a two-socket host this keeps the launching thread, NCCL's proxy thread and -- by first touch --
    the pinned staging buffers allocated afterwards on the GPU's own socket, so that eight ranks do not pull their
    host->device copies across the socket interconnect.  Returns the core list, or ``None`` when NVML or
    ``sched_setaffinity`` is unavailable (nothing is changed then)."""
    try:
        import os
        import pynvml
        pynvml.nvmlInit()
        handle = None
        try:        # CUDA ordinals follow CUDA_VISIBLE_DEVICES, NVML indices do not: identify the GPU by its PCI address
            props = torch.cuda.get_device_properties(device_index)
            bus_id = "%08x:%02x:%02x.0" % (props.pci_domain_id, props.pci_bus_id, props.pci_device_id)
            handle = pynvml.nvmlDeviceGetHandleByPciBusId(bus_id.encode())
        except Exception:
            handle = None
        if handle is None:
            handle = pynvml.nvmlDeviceGetHandleByIndex(device_index)
        words = pynvml.nvmlDeviceGetCpuAffinity(handle, (os.cpu_count() + 63) // 64)
        cores = [64 * i + b for i, w in enumerate(words) for b in range(64) if (int(w) >> b) & 1]
        allowed = sorted(set(cores) & set(os.sched_getaffinity(0)))
        if not allowed:
            return None
        os.sched_setaffinity(0, allowed)
        return allowed
    except Exception:
        return None


def shard_rows(total_rows: int, rank: int, world_size: int) -> Tuple[int, int]:
    """``[start, stop)`` of this rank's rows; whole rows stay on one rank (a V-trace
    recurrence never crosses ranks).  Equal shards when ``world_size`` divides
    ``total_rows``; otherwise the first ranks take one extra row."""
    if not 0 <= rank < world_size:
        raise ValueError(f"rank {rank} outside world of {world_size}")
    base, extra = divmod(total_rows, world_size)
    start = rank * base + min(rank, extra)
    return start, start + base + (1 if rank < extra else 0)


def shard_batch(batch: dict, rows: int, T: int, rank: int, world_size: int, time_major: bool = False) -> dict:
    """Slice a flat ``[rows*(T+1), ...]`` learner batch to this rank's rows.

    With the reference's reshape semantics (quirk Q7) a row is ``T+1`` consecutive flat
    entries; with ``time_major`` the flat order is ``(t, row)``.
    """
    lo, hi = shard_rows(rows, rank, world_size)
    out = {}
    for k, v in batch.items():
        if time_major:
            x = v.reshape((T + 1, rows) + tuple(v.shape[1:]))[:, lo:hi]
            out[k] = x.reshape((-1,) + tuple(v.shape[1:]))
        else:
            out[k] = v[lo * (T + 1): hi * (T + 1)]
    return out


class GradAllReducer:
    """Sums gradients (and a few extra scalars) across ranks through ONE flat bucket."""

    def __init__(self, process_group: Optional[dist.ProcessGroup] = None):
        self.group = process_group
        self.enabled = dist.is_available() and dist.is_initialized()
        self.world_size = dist.get_world_size(process_group) if self.enabled else 1
        self.rank = dist.get_rank(process_group) if self.enabled else 0
        self._flat: Optional[torch.Tensor] = None
        self.calls = 0

    def _bucket(self, numel: int, like: torch.Tensor) -> torch.Tensor:
        if self._flat is None or self._flat.numel() != numel or self._flat.device != like.device:
            self._flat = torch.empty(numel, dtype=torch.float32, device=like.device)
        return self._flat

    # ---- overlapped path: gradients ARE the bucket ----------------------------------------------------------
    # ``attach`` makes every ``p.grad`` a view into one flat buffer, the parameters that finish their backward first
    # (the actor / critic heads: 97 % of the bytes at 16x16) at its start.  ``begin`` zeroes the buffer in place of
    # ``optimizer.zero_grad()`` (autograd then accumulates into the views) and arms a hook on the head parameters: as
    # soon as their gradients exist, that slice is all-reduced asynchronously while the encoder's backward is still
    # running; ``finish`` reduces the rest (plus the logged loss scalars) and waits for both.  No copy in, no copy out.
    def attach(self, model: torch.nn.Module, first: Tuple[str, ...] = ("actor.", "critic.")) -> None:
        named = [(n, p) for n, p in model.named_parameters() if p.requires_grad]
        head = [(n, p) for n, p in named if n.startswith(first)]
        rest = [(n, p) for n, p in named if not n.startswith(first)]
        order = head + rest
        sizes = [p.numel() for _, p in order]
        dev = order[0][1].device
        self._n_extra = 8
        self._grads = torch.zeros(sum(sizes) + self._n_extra, dtype=torch.float32, device=dev)
        off = 0
        for (_, p), n in zip(order, sizes):
            p.grad = self._grads[off: off + n].view_as(p)
            off += n
        self._head_numel = sum(p.numel() for _, p in head)
        self._total_numel = off
        self._head_params = [p for _, p in head]
        self._pending = 0
        self._early = None
        self._attached = model
        for p in self._head_params:
            p.register_post_accumulate_grad_hook(self._head_ready)

    def _mark(self, name: str) -> None:
        """With ``self.marks = {}`` set by a caller (scripts/dp_timeline.py), records a CUDA event on the compute stream
        at the named point of the step; ``None`` (the default) costs nothing."""
        marks = getattr(self, "marks", None)
        if marks is not None:
            ev = torch.cuda.Event(enable_timing=True)
            ev.record()
            marks[name] = ev

    def _head_ready(self, _param) -> None:
        if self._pending <= 0:
            return
        self._pending -= 1
        if self._pending == 0:
            self._mark("head_grads_ready")
        if self._pending == 0 and self.world_size > 1 and self._head_numel:
            # issued behind the kernels that produced these gradients; runs on the backend's own stream while the
            # rest of the backward pass keeps the compute stream busy
            self._early = dist.all_reduce(self._grads[: self._head_numel], op=dist.ReduceOp.SUM, group=self.group,
                                          async_op=True)

    def begin(self, model: torch.nn.Module) -> None:
        """Call instead of ``optimizer.zero_grad()`` right before ``backward()``."""
        if getattr(self, "_attached", None) is not model:
            self.attach(model)
        else:
            for p in model.parameters():            # an optimizer.zero_grad(set_to_none=True) elsewhere drops the views
                if p.requires_grad and p.grad is None:
                    self.attach(model)
                    break
        self._mark("begin")
        self._grads.zero_()
        self._pending = len(self._head_params)
        self._early = None

    def finish(self, extra: Optional[torch.Tensor] = None):
        """After ``backward()``: sums what ``begin``'s hook has not already put on the wire; returns the summed ``extra``."""
        n_extra = extra.numel() if extra is not None else 0
        if n_extra > self._n_extra:
            raise ValueError("GradAllReducer.finish: at most 8 extra scalars")
        tail = self._grads[self._total_numel: self._total_numel + n_extra]
        self._mark("backward_done")
        if n_extra:
            tail.copy_(extra.reshape(-1))
        if self.world_size > 1:
            start = self._head_numel if self._early is not None else 0
            late = dist.all_reduce(self._grads[start: self._total_numel + n_extra], op=dist.ReduceOp.SUM,
                                   group=self.group, async_op=True)
            if self._early is not None:
                self._early.wait()
            late.wait()
            self._early = None
        self._mark("gradients_summed")
        self._pending = 0
        self.calls += 1
        return tail.clone().view_as(extra) if n_extra else extra

    def allreduce(self, model: torch.nn.Module, extra: Optional[torch.Tensor] = None):
        """In-place sum of every ``p.grad`` over the ranks; returns the summed ``extra``."""
        grads: List[torch.Tensor] = [p.grad for p in model.parameters() if p.grad is not None]
        if self.world_size == 1 or not grads:
            return extra
        sizes = [g.numel() for g in grads]
        n_extra = extra.numel() if extra is not None else 0
        flat = self._bucket(sum(sizes) + n_extra, grads[0])
        views = list(flat[: sum(sizes)].split(sizes))
        torch._foreach_copy_(views, [g.reshape(-1) for g in grads])
        if n_extra:
            flat[sum(sizes):].copy_(extra.reshape(-1))
        dist.all_reduce(flat, op=dist.ReduceOp.SUM, group=self.group)
        torch._foreach_copy_([g.view(-1) for g in grads], views)
        self.calls += 1
        if n_extra:
            return flat[sum(sizes):].clone().view_as(extra)
        return extra
