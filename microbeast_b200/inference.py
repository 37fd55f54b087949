"""Actor-side batched inference.

In the reference every actor process runs its own CPU forward per env step
(``microbeast.py:86-87`` -> ``model.py:165-216``: encoder + 7*cells distribution objects for 6
envs at a time).  Here one GPU model serves all actors: ``InferenceServer`` gathers the
``get_action`` requests that arrive within a short window, evaluates them as ONE batch (one
encoder pass, one ``mb_head_sample`` launch) and hands each actor its slice.  Frames are independent
in the policy head, so an actor gets bit-identical actions whether it is served alone or in a batch
(given the same uniforms).
"""
from __future__ import annotations

import contextlib
import queue
import threading
import time
from typing import Optional

import torch

AGENT_KEYS = ("action", "policy_logits", "logprobs", "baseline")


class _Request:
    __slots__ = ("obs", "mask", "uniforms", "n", "event", "result", "error")

    def __init__(self, obs, mask, uniforms):
        self.obs, self.mask, self.uniforms = obs, mask, uniforms
        self.n = obs.shape[0]
        self.event = threading.Event()
        self.result = None
        self.error = None


class InferenceServer:
    """Dynamic batching of acting-mode ``Agent.get_action`` calls from many actor threads."""

    def __init__(self, agent, max_batch_envs: int = 256, max_wait_ms: float = 0.5, seed: Optional[int] = None,
                 model_lock=None):
        self.agent = agent
        # held while a batch's kernels are enqueued: a learner that updates ``agent``'s weights in place takes the
        # same lock around its step, so inference never reads a half-updated parameter set
        self.model_lock = model_lock if model_lock is not None else contextlib.nullcontext()
        self.device = next(agent.parameters()).device
        self.max_batch_envs = max_batch_envs
        self.max_wait = max_wait_ms * 1e-3
        self.requests: "queue.Queue[Optional[_Request]]" = queue.Queue()
        self.thread: Optional[threading.Thread] = None
        self.gen = None
        if seed is not None:
            self.gen = torch.Generator(device=self.device)
            self.gen.manual_seed(seed)
        self.batches = 0
        self.served = 0
        self.max_seen = 0

    # ---- actor side -----------------------------------------------------------------------
    def infer(self, env_output: dict, uniforms: Optional[torch.Tensor] = None) -> dict:
        """Blocking drop-in for ``agent.get_action(env_output)[0]`` (acting mode): ``env_output`` holds
        ``obs (1,1,n,h,w,27)`` and ``action_mask (1,n,78*cells)`` as produced by ``Env_Packer``;
        returns CPU tensors ``action [n,7C]``, ``policy_logits``, ``logprobs [n]``, ``baseline [1,n]``."""
        req = _Request(env_output["obs"][0, 0], env_output["action_mask"][0], uniforms)
        self.requests.put(req)
        req.event.wait()
        if req.error is not None:
            raise req.error
        return req.result

    # ---- server side ----------------------------------------------------------------------
    def start(self) -> "InferenceServer":
        self.thread = threading.Thread(target=self._loop, daemon=True)
        self.thread.start()
        return self

    def stop(self) -> None:
        self.requests.put(None)
        if self.thread is not None:
            self.thread.join(timeout=30)

    def _gather(self):
        first = self.requests.get()
        if first is None:
            return None
        batch, total = [first], first.n
        deadline = time.perf_counter() + self.max_wait
        while total < self.max_batch_envs:
            left = deadline - time.perf_counter()
            try:
                nxt = self.requests.get(timeout=max(left, 0.0)) if left > 0 else self.requests.get_nowait()
            except queue.Empty:
                break
            if nxt is None:
                self.requests.put(None)          # let the loop see the stop marker after this batch
                break
            batch.append(nxt)
            total += nxt.n
        return batch

    def _loop(self):
        cells = self.agent.cells
        while True:
            batch = self._gather()
            if batch is None:
                return
            try:
                obs = torch.cat([r.obs for r in batch]).to(self.device, non_blocking=True)
                mask = torch.cat([r.mask for r in batch]).to(self.device, non_blocking=True)
                n = obs.shape[0]
                u = torch.rand((n, 7 * cells), device=self.device, generator=self.gen)
                off = 0
                for r in batch:                       # callers may pin their own uniforms (tests, replay)
                    if r.uniforms is not None:
                        u[off:off + r.n] = r.uniforms.to(self.device)
                    off += r.n
                with torch.no_grad(), self.model_lock:
                    out, _ = self.agent.get_action(dict(obs=obs.view((1, 1) + tuple(obs.shape)),
                                                        action_mask=mask.view((1,) + tuple(mask.shape))), uniforms=u)
                host = {k: out[k].cpu() for k in AGENT_KEYS}
                off = 0
                for r in batch:
                    r.result = dict(action=host["action"][off:off + r.n],
                                    policy_logits=host["policy_logits"][off:off + r.n],
                                    logprobs=host["logprobs"][off:off + r.n],
                                    baseline=host["baseline"][:, off:off + r.n])
                    off += r.n
                self.batches += 1
                self.served += len(batch)
                self.max_seen = max(self.max_seen, len(batch))
            except Exception as exc:                  # hand the failure to every waiting actor
                for r in batch:
                    r.error = exc
            for r in batch:
                r.event.set()
