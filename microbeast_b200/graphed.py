"""CUDA-graphed learner step ("CUDA-graphed step").

``PPO_learn`` enqueues ~300 kernels per step (the PyTorch encoder's forward and backward, the four
kernels of this library, fused Adam).  On small maps (8x8: 2.6 k frames per step) the GPU finishes
them faster than one host thread can launch them; with eight ranks on one box the host side also
adds jitter to the max-over-ranks step time.  ``GraphedLearner`` captures the whole step --
forward, V-trace + losses, backward, optimiser -- once into a CUDA graph over static device
buffers and replays it with a single launch per step.  Same kernels, same order, same arithmetic as
``PPO_learn`` (reference ``libs/utils.py:223-342``); only the launch mechanism differs.

With a ``reducer`` (data-parallel ranks) the step is captured as two graphs around the NCCL
allreduce, which stays an ordinary eager call: [forward + backward + bucket fill] -> allreduce ->
[bucket scatter + optimiser].
"""
from __future__ import annotations

from typing import Dict, Optional

import torch
import torch.nn as nn

from . import ops
from .learner import LEARNER_KEYS, learner_losses
from .parallel import GradAllReducer


class GraphedLearner:
    """One learner step as a CUDA graph replay.

    ``step(batch)`` copies the batch (host, pinned or device; plain or the compact wire format of
    ``buffers.pack_batch``) into static device buffers and replays the graph; it returns the four
    losses ``[pg, value, entropy, total]`` like ``PPO_learn`` (python floats, or with ``sync=False``
    the device tensor, valid until the next ``step``).

    The optimiser must tolerate capture (``torch.optim.Adam(..., capturable=True)``, fused or not;
    plain SGD works as is).  The first ``step`` runs ``warmup`` eager steps on its batch (cuDNN
    autotuning, lazy handles, optimiser-state allocation) and then puts the weights and the optimiser
    state back in place -- state tensors the warm-up created are reset to zeros -- before capturing.
    The model's ``sparse_head`` must be False: that (torch) cell-list path sizes its tensors from the data
    and synchronises, which a captured graph cannot do.  The default backward of the actor head -- device-side cell
    lists, ``mb_cells_bwd`` -- has no such limit and is captured as is.
    """

    def __init__(self, learner_model: nn.Module, optimizer: torch.optim.Optimizer, B: int, T: int, n_envs: int = 1, *,
                 gamma=0.99, rho_bar=1.0, c_bar=1.0, pg_rho_bar=1.0, entropy_coef=0.01,
                 reducer: Optional[GradAllReducer] = None, time_major: bool = False,
                 actor_model: Optional[nn.Module] = None, warmup: int = 3):
        if getattr(learner_model, "sparse_head", False):
            raise ValueError("GraphedLearner: Agent.sparse_head synchronises with the host and cannot be captured")
        self.model = learner_model
        self.actor_model = actor_model if actor_model is not learner_model else None
        self.optimizer = optimizer
        self.rows, self.T = B * n_envs, T
        self.reducer = reducer if (reducer is not None and reducer.world_size > 1) else None
        self.world = self.reducer.world_size if self.reducer is not None else 1
        self.kw = dict(gamma=gamma, rho_bar=rho_bar, c_bar=c_bar, pg_rho_bar=pg_rho_bar,
                       entropy_coef=entropy_coef, global_rows=self.rows * self.world, time_major=time_major)
        self.device = next(learner_model.parameters()).device
        self.warmup = max(int(warmup), 1)
        self.static: Dict[str, torch.Tensor] = {}
        self.graphs = None
        self.losses: Optional[torch.Tensor] = None
        self.launches_per_step: Dict[str, int] = {}
        self.steps = 0

    # ---- staging --------------------------------------------------------------------------
    def _load(self, batch: Dict[str, torch.Tensor]) -> None:
        compact = "action_mask_bits" in batch
        keys = [("action_mask_bits" if (k == "action_mask" and compact) else k) for k in LEARNER_KEYS]
        first = not self.static
        for k in keys:
            t = batch[k]
            if first:
                self.static[k] = torch.empty(t.shape, dtype=t.dtype, device=self.device)
            buf = self.static.get(k)
            if buf is None or buf.shape != t.shape or buf.dtype != t.dtype:
                raise ValueError(f"GraphedLearner: batch[{k!r}] changed shape/dtype/format after capture "
                                 f"({tuple(t.shape)} {t.dtype}); build a new GraphedLearner")
            buf.copy_(t, non_blocking=True)

    def _device_batch(self) -> Dict[str, torch.Tensor]:
        """The static buffers as the dict ``learner_losses`` reads; compact formats are widened by
        ``mb_unpack_rollout`` inside the captured region."""
        out = dict(self.static)
        bits = out.pop("action_mask_bits", None)
        a8 = out["action"] if out["action"].dtype == torch.uint8 else None
        if bits is not None or a8 is not None:
            cells = out["action"].shape[1] // 7
            mask, act = ops.unpack_rollout(bits, a8, cells)
            if bits is not None:
                out["action_mask"] = mask
            if a8 is not None:
                out["action"] = act
        return out

    # ---- the step, as PPO_learn runs it ------------------------------------------------------
    def _forward_backward(self):
        # "kernel": the device-side cell-list backward (mb_cells_bwd, no host sync) when the fused forward ran, else the
        # dense kernels; never the torch cell-list path, which sizes tensors from the data and synchronises
        keep = getattr(self.model, "sparse_backward", False)
        self.model.sparse_backward = "kernel" if keep else False
        try:
            loss4, _ = learner_losses(self.model, self._device_batch(), self.rows, self.T, **self.kw)
            loss4[3].backward()                                                            # utils.py:334
        finally:
            self.model.sparse_backward = keep
        return loss4.detach()

    def _capture(self) -> None:
        params = [p for p in self.model.parameters()]
        # the warm-up steps below must leave no trace: keep the weights and the optimiser state to put back
        saved_params = [p.detach().clone() for p in params]
        saved_state = {id(p): {k: (v.detach().clone() if torch.is_tensor(v) else v) for k, v in st.items()}
                       for p, st in self.optimizer.state.items()}
        side = torch.cuda.Stream(device=self.device)
        side.wait_stream(torch.cuda.current_stream(self.device))
        with torch.cuda.stream(side):       # eager warm-up: cuDNN autotuning, lazy handles, optimiser state
            for _ in range(self.warmup):
                self.optimizer.zero_grad(set_to_none=True)
                losses = self._forward_backward()
                if self.reducer is not None:
                    losses = self.reducer.allreduce(self.model, losses)
                self.optimizer.step()
            # restore IN PLACE (the capture binds to these tensors): weights back, optimiser state back to
            # what it was -- or to zeros where the warm-up created it (Adam moments, step counters)
            with torch.no_grad():
                torch._foreach_copy_(params, saved_params)
                for p, st in self.optimizer.state.items():
                    old = saved_state.get(id(p), {})
                    for k, v in st.items():
                        if torch.is_tensor(v):
                            v.copy_(old[k]) if (k in old and torch.is_tensor(old[k])) else v.zero_()
        torch.cuda.current_stream(self.device).wait_stream(side)
        torch.cuda.synchronize(self.device)

        before = dict(ops.launch_counts)
        self.optimizer.zero_grad(set_to_none=True)
        g1 = torch.cuda.CUDAGraph()
        if self.reducer is None:
            with torch.cuda.graph(g1):
                self.losses = self._forward_backward()
                self.optimizer.step()                                                      # utils.py:335
                self._sync_actor()
            self.graphs = (g1,)
        else:
            with torch.cuda.graph(g1):
                local = self._forward_backward()
                self._flat, self._views, self._grads = self._fill_bucket(params, local)
            g2 = torch.cuda.CUDAGraph()
            with torch.cuda.graph(g2, pool=g1.pool()):
                torch._foreach_copy_([g.view(-1) for g in self._grads], self._views)
                self.losses = self._flat[-4:].clone()
                self.optimizer.step()
                self._sync_actor()
            self.graphs = (g1, g2)
        self.launches_per_step = {k: v - before.get(k, 0) for k, v in ops.launch_counts.items()
                                  if v - before.get(k, 0)}

    def _sync_actor(self) -> None:
        if self.actor_model is not None:                                                   # utils.py:337
            with torch.no_grad():
                for dst, src in zip(self.actor_model.parameters(), self.model.parameters()):
                    dst.copy_(src)

    def _fill_bucket(self, params, losses):
        grads = [p.grad for p in params if p.grad is not None]
        sizes = [g.numel() for g in grads]
        flat = torch.empty(sum(sizes) + 4, dtype=torch.float32, device=self.device)
        views = list(flat[: sum(sizes)].split(sizes))
        torch._foreach_copy_(views, [g.reshape(-1) for g in grads])
        flat[sum(sizes):].copy_(losses.reshape(-1))
        return flat, views, grads

    # ---- public ---------------------------------------------------------------------------
    def step(self, batch: Dict[str, torch.Tensor], sync: bool = True):
        self._load(batch)
        if self.graphs is None:
            self._capture()
            # the capture itself does not execute: the first real step is the replay below
        self.graphs[0].replay()
        if self.reducer is not None:
            import torch.distributed as dist
            dist.all_reduce(self._flat, op=dist.ReduceOp.SUM, group=self.reducer.group)
            self.graphs[1].replay()
        for k, v in self.launches_per_step.items():      # replays launch the same kernels; keep the counters honest
            ops.launch_counts[k] = ops.launch_counts.get(k, 0) + v
        self.steps += 1
        if not sync:
            return self.losses
        four = self.losses.cpu().tolist()                                                  # utils.py:342
        ops.check_status(self.device)
        return four
