"""Drop-in learner step for reference ``libs/utils.py:223-342`` (``PPO_learn``).

Same signature and return value; the per-component Python loops, the reverse-time
V-trace loop and the autograd graph of ~10*7*C nodes are replaced by four kernel
launches (head forward, V-trace + losses + their gradients, head backward) around the
PyTorch encoder.  Hyper-parameters that are literals in the reference (``gamma=0.99``,
``rho_bar = c_bar = pg_rho_bar = 1``, entropy coefficient ``0.01`` -- quirk Q10) are
keyword arguments defaulting to those literals.
"""
from __future__ import annotations

import threading
from typing import Dict, Optional

import torch
import torch.nn as nn

from . import ops
from .parallel import GradAllReducer

# keys of the rollout batch the learner reads (reference buffers carry 11, libs/utils.py:34-46)
LEARNER_KEYS = ("obs", "action_mask", "action", "reward", "done", "logprobs")


class _Stager:
    """Host -> device staging of a learner batch with the big copies off the critical path.

    ``obs`` is copied on the compute stream (the encoder needs it first); ``action_mask``,
    ``action`` and the small per-frame keys go over a side stream into persistent device
    buffers and are only waited for right before the policy-head kernel, so their DMA
    (0.7 GB at 16x16, T=80, 256 rows) overlaps the encoder forward of the same step.
    """

    def __init__(self, slots: int = 1, obs_on_side_stream: bool = False, prepack_cells: bool = False):
        self.buffers = {}
        self.streams = {}
        self.slots = max(int(slots), 1)
        self.turn = 0
        self.obs_on_side_stream = obs_on_side_stream
        self.prepack_cells = prepack_cells      # also build the fused actor head's cell records on the side stream

    def _buf(self, name, shape, dtype, device, slot=None):
        key = (name, tuple(shape), dtype, device, (self.turn % self.slots) if slot is None else slot)
        buf = self.buffers.get(key)
        if buf is None:
            buf = self.buffers[key] = torch.empty(tuple(shape), dtype=dtype, device=device)
        return buf

    def stage(self, batch, device, keys, slot_free=None, defer=False):
        """Host -> device copies on the side stream, then the device-side widening of the compact wire format (and, for
        a prefetcher, the fused head's cell records).  Returns ``(out, ready_event)``; with ``defer=True`` the widening
        is left to the caller: ``(out, copied_event, finish)`` where ``finish()`` enqueues it (same side stream, after
        the copies), updates ``out`` in place and returns the ready event.
        ``slot_free``: a recorded CUDA event after which the staging buffers of this turn's slot are no longer read
        (``BatchPrefetcher``); without it the side stream waits for everything enqueued on the compute stream so far."""
        compact = "action_mask_bits" in batch        # wire format of buffers.pack_batch / compact store
        keys = [("action_mask_bits" if (k == "action_mask" and compact) else k) for k in keys]
        if all(batch[k].is_cuda for k in keys) and not compact and batch["action"].dtype == torch.int64:
            return {k: batch[k] for k in keys}, None
        side = self.streams.get(device)
        if side is None:
            side = self.streams[device] = torch.cuda.Stream(device=device)
        main = torch.cuda.current_stream(device)
        slot = self.turn % self.slots
        self.turn += 1
        out = {}
        if slot_free is not None:
            side.wait_event(slot_free)          # the step that read this slot's buffers last has finished
        else:
            side.wait_stream(main)              # the staging buffers may still be read by the last step
        for k in keys:
            t = batch[k]
            if t.is_cuda:
                out[k] = t
                continue
            buf = self._buf(k, t.shape, t.dtype, device, slot)
            if k == "obs" and not self.obs_on_side_stream:
                buf.copy_(t, non_blocking=True)
            else:
                with torch.cuda.stream(side):
                    buf.copy_(t, non_blocking=True)
            out[k] = buf

        def finish():
            if compact or out["action"].dtype == torch.uint8:
                # widen bit-packed masks / uint8 actions on the side stream, still off the critical path
                n = out["action"].shape[0]
                cells = out["action"].shape[1] // 7
                bits = out.pop("action_mask_bits", None)
                a8 = out["action"] if out["action"].dtype == torch.uint8 else None
                with torch.cuda.stream(side):
                    mask, act = ops.unpack_rollout(
                        bits, a8, cells,
                        self._buf("wide_mask", (n, cells * 78), torch.uint8, device, slot) if bits is not None else None,
                        self._buf("wide_action", (n, cells * 7), torch.int64, device, slot) if a8 is not None else None)
                if bits is not None:
                    out["action_mask"] = mask
                if a8 is not None:
                    out["action"] = act
            if self.prepack_cells and out["action"].dtype == torch.int64 and out["action_mask"].dim() == 2:
                # the operand form of mb_actor_head_fwd (16-byte (cell, frame) records + alive bitmap): 0.17 ms at config
                # 2 that the step no longer spends on its own stream (model.Agent.get_action picks "_cell_meta" up)
                n = out["action"].shape[0]
                cells = out["action"].shape[1] // 7
                with torch.cuda.stream(side):
                    out["_cell_meta"] = ops.pack_cells(
                        out["action_mask"], out["action"], cells,
                        out=self._buf("cell_meta", (cells, n, 4), torch.int32, device, slot),
                        out_alive=self._buf("cell_alive", (cells, (n + 31) // 32), torch.int32, device, slot))
            ready = torch.cuda.Event()
            ready.record(side)
            return ready

        if defer:
            copied = torch.cuda.Event()
            copied.record(side)
            return out, copied, finish
        return out, finish()


_stager = _Stager()


class BatchPrefetcher:
    """Double-buffered host -> device staging: ``submit(host_batch)`` starts the copies (and the widening of the
    compact wire format) of the NEXT batch on a side stream and returns at once; the dictionary it returns goes to
    ``PPO_learn`` / ``learner_losses`` like any device batch (they wait for its ``"_ready"`` event).  Submitting batch
    k+1 before calling ``PPO_learn`` on batch k hides the transfer behind step k -- what the reference's commented-out
    ``.to(device)`` (``libs/utils.py:216``) would pay in full on every update, and what eight ranks sharing one host's
    memory bandwidth otherwise all pay at the same moment.

        pf = BatchPrefetcher(device); nxt = pf.submit(get_batch(...))
        while training:
            cur, nxt = nxt, pf.submit(get_batch(...))
            PPO_learn(model, model, name, cur, optimizer, B, T, n_envs)
    """

    def __init__(self, device, slots: int = 2):
        self.device = torch.device(device)
        self._stager = _Stager(slots=slots, obs_on_side_stream=True, prepack_cells=True)
        self._last = {}                        # slot -> the dictionary staged into it last

    def submit(self, batch: Dict[str, torch.Tensor], keys=LEARNER_KEYS) -> Dict[str, torch.Tensor]:
        """May be called while the step on the previous batch is still being enqueued or running -- best from inside it:

            nxt = [pf.submit(get_batch(...))]
            while training:
                cur = nxt.pop()
                PPO_learn(model, model, name, cur, optimizer, B, T, n_envs,
                          prefetch=lambda: nxt.append(pf.submit(get_batch(...))))     # runs after step k's forward is enqueued

        ``PPO_learn`` records ``"_release"`` once it no longer reads a staged batch; a slot is overwritten only after
        the release of the batch staged into it before (two slots: batch k+1 waits for step k-1, not for step k)."""
        slot = self._stager.turn % self._stager.slots
        prev = self._last.get(slot)
        free = prev["_release"] if (prev is not None and prev.get("_released")) else None
        # (a batch that was staged but never went through PPO_learn has no release: stage() then waits for the whole
        # compute stream, as it does without a prefetcher)
        staged = self._stager.stage(batch, self.device, keys, slot_free=free, defer=True)
        out = dict(staged[0])
        if len(staged) == 3:
            _, copied, finish = staged
            live = staged[0]                   # finish() updates this dictionary in place

            def _finish():                     # the device-side widening + cell records, enqueued when the consumer says so
                ready = finish()
                out.update(live)
                out.pop("action_mask_bits", None)
                out["_ready"] = ready
            out["_ready_copy"] = copied        # obs too came over the side stream: wait for the copies before the encoder
            out["_ready_all"] = True
            out["_finish"] = _finish
            out["_release"] = torch.cuda.Event()
            out["_released"] = False
            self._last[slot] = out
        return out


def batch_to_device(batch: Dict[str, torch.Tensor], device, keys=LEARNER_KEYS):
    """Host -> device copy of the keys the learner needs (the ``.to(device)`` the reference
    left commented out at ``libs/utils.py:216``).  Returns ``(device_batch, ready_event)``;
    everything but ``obs`` is valid on the compute stream only after ``ready_event``
    (``Agent.get_action`` waits for it under the key ``"_ready"``).  Pinned host tensors copy
    asynchronously."""
    return _stager.stage(batch, torch.device(device), keys)


def learner_losses(learner_model: nn.Module, batch: Dict[str, torch.Tensor], rows: int, T: int, *,
                   gamma=0.99, rho_bar=1.0, c_bar=1.0, pg_rho_bar=1.0, entropy_coef=0.01,
                   global_rows: Optional[int] = None, time_major: bool = False):
    """Forward part of ``PPO_learn`` (``utils.py:234-329``): returns ``(loss4, aux)`` with
    ``loss4 = [pg, value, entropy, total]`` (device tensor, differentiable) and ``aux``
    holding ``vs``, ``pg_advantages``, ``logprobs``, ``baseline``.

    ``global_rows``: total rows over all data-parallel ranks; the means of
    ``utils.py:320-326`` are then taken over the global ``rows*T`` so that a plain
    sum-allreduce of gradients reproduces the single-GPU step.
    ``time_major=False`` keeps the reference's ``reshape((R, T+1))`` of a batch flattened in
    ``(t, b, e)`` order (quirk Q7); ``True`` treats the flat batch as what it is, ``[T+1, rows]``,
    so that a row really is one trajectory, and runs the time-major V-trace kernel on it in place.
    """
    if "_finish" in batch:                   # a BatchPrefetcher batch whose device-side widening is still to be enqueued
        batch.pop("_finish")()
    kw = {}
    if getattr(learner_model, "skips_policy_logits", False):
        kw["policy_logits"] = False          # never read below (utils.py:237-329): the fused kernel need not store them
    out, _ = learner_model.get_action(batch, learning=True, inds=[], agent_state=(), **kw)      # utils.py:234

    ready = batch.get("_ready")
    if ready is not None:
        torch.cuda.current_stream().wait_event(ready)     # reward / done / logprobs staged on the side stream
    inv_n = 1.0 / float((global_rows or rows) * T)
    if time_major:
        # the flat batch is (t, row): views [T+1, rows]; slot T is the bootstrap, batch keys shift by one
        # slot (utils.py:258-265).  No transpose, no copy: mb_vtrace_loss_tmajor reads this layout.
        def tm(x):
            return x.reshape((T + 1, rows) + tuple(x.shape[1:]))
        logp, base, ent = tm(out["logprobs"]), tm(out["baseline"].flatten()), tm(out["entropy"])
        bootstrap = base[-1].detach()
        loss4, vs, adv = ops.VTraceLoss.apply(
            logp[:-1], base[:-1], ent[:-1], tm(batch["logprobs"])[1:], bootstrap, tm(batch["reward"])[1:],
            tm(batch["done"])[1:], gamma, rho_bar, c_bar, pg_rho_bar, entropy_coef, inv_n, True)
        return loss4, dict(vs=vs.transpose(0, 1), pg_advantages=adv.transpose(0, 1), logprobs=logp.transpose(0, 1),
                           baseline=base.transpose(0, 1), entropy=ent.transpose(0, 1))

    def rt(x):                       # utils.py:237-253: the reference's reshape (quirk Q7)
        return x.reshape((rows, T + 1) + tuple(x.shape[1:]))

    logp = rt(out["logprobs"])
    base = rt(out["baseline"].flatten())
    ent = rt(out["entropy"])
    bootstrap = base[:, -1].detach()                                                       # utils.py:258
    loss4, vs, adv = ops.VTraceLoss.apply(
        logp[:, :-1], base[:, :-1], ent[:, :-1],                                           # utils.py:265
        rt(batch["logprobs"])[:, 1:], bootstrap, rt(batch["reward"])[:, 1:], rt(batch["done"])[:, 1:],  # :263
        gamma, rho_bar, c_bar, pg_rho_bar, entropy_coef, inv_n)
    return loss4, dict(vs=vs, pg_advantages=adv, logprobs=logp, baseline=base, entropy=ent)


def PPO_learn(actor_model: nn.Module, learner_model, exp_name: str, batch, optimizer, B: int, T: int,
              n_envs: int, lock=threading.Lock(), *, gamma=0.99, rho_bar=1.0, c_bar=1.0, pg_rho_bar=1.0,
              entropy_coef=0.01, reducer: Optional[GradAllReducer] = None, time_major: bool = False,
              sync: bool = True, sparse_head=None, prefetch=None):
    """Performs a learning (optimization) step -- reference ``libs/utils.py:223-342``.

    ``batch``: flat ``[N, ...]`` tensors, ``N = B*n_envs*(T+1)``, on the host (copied to the
    model's device here, pinned memory makes that asynchronous) or already on the device.
    Returns ``[pg_loss, value_loss, entropy_loss, total_loss]`` as python floats
    (``utils.py:342``); with ``sync=False`` the 4-element device tensor is returned instead
    and nothing synchronises.  ``reducer`` sums gradients across data-parallel ranks; the
    returned losses are then the global ones.  ``sparse_head`` (False / True / "auto") overrides
    ``learner_model.sparse_head`` for this step: evaluate the actor head only on cells with a valid
    action (same losses and gradients; costs two host syncs inside the step).
    ``prefetch``: a callable run once the forward pass and the losses are enqueued, before the backward pass -- the
    place to ``BatchPrefetcher.submit`` the NEXT batch (return what ``submit`` returned): the device then has a forward
    pass of work queued (the host time of the submit is hidden) and the copy has the whole backward pass to overlap
    with; the batch's device-side widening is enqueued at the end of this step.
    """
    if not isinstance(lock, type(threading.Lock())):
        lock = threading.Lock()      # reference quirk Q15: microbeast.py:251 passes exp_name here
    device = next(learner_model.parameters()).device
    rows = B * n_envs
    with lock:
        if isinstance(batch, dict) and "_finish" in batch:      # staged by a BatchPrefetcher, widening not yet enqueued
            batch.pop("_finish")()
        dev_batch, ready = batch_to_device(batch, device)
        if ready is not None:
            dev_batch["_ready"] = ready
        elif "_ready" in batch:                 # staged ahead of time by a BatchPrefetcher
            dev_batch["_ready"] = batch["_ready"]
            if "_cell_meta" in batch:           # the fused actor head's cell records, packed on the prefetcher's stream
                dev_batch["_cell_meta"] = batch["_cell_meta"]
            if batch.get("_ready_all"):
                torch.cuda.current_stream().wait_event(batch.get("_ready_copy", batch["_ready"]))
        world = reducer.world_size if reducer is not None else 1
        keep = getattr(learner_model, "sparse_head", None)
        if sparse_head is not None:
            learner_model.sparse_head = sparse_head            # for this step only: restored below
        try:
            loss4, _aux = learner_losses(learner_model, dev_batch, rows, T, gamma=gamma, rho_bar=rho_bar,
                                         c_bar=c_bar, pg_rho_bar=pg_rho_bar, entropy_coef=entropy_coef,
                                         global_rows=rows * world, time_major=time_major)
        finally:
            if sparse_head is not None:
                learner_model.sparse_head = keep
        staged_next = prefetch() if prefetch is not None else None
        if reducer is not None and reducer.world_size > 1:
            # gradients live in the reducer's flat bucket: zeroed there, the head's slice goes on the wire as soon as
            # the actor-head backward has run and overlaps the encoder's backward (parallel.GradAllReducer)
            reducer.begin(learner_model)                                                   # utils.py:333
            loss4[3].backward()                                                            # utils.py:334
            losses = reducer.finish(loss4.detach())
        else:
            optimizer.zero_grad()                                                          # utils.py:333
            loss4[3].backward()                                                            # utils.py:334
            losses = loss4.detach()
        optimizer.step()                                                                   # utils.py:335
        if batch.get("_release") is not None:   # staged by a BatchPrefetcher: its buffers are free from here on
            batch["_release"].record(torch.cuda.current_stream())
            batch["_released"] = True
        if isinstance(staged_next, dict) and "_finish" in staged_next:
            # the next batch's copies have been running since `prefetch`; its device-side widening is enqueued only now,
            # behind this step's kernels in host order, so that it shares the GPU with the encoder's backward pass
            # rather than with the two persistent actor-head kernels
            staged_next.pop("_finish")()
        if actor_model is not learner_model:
            actor_model.load_state_dict(learner_model.state_dict())                        # utils.py:337
        if not sync:
            return losses
        four = losses.cpu().tolist()                                                       # utils.py:342
        ops.check_status(device)
        return four
