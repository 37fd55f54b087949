"""Learner driver -- thin shim for reference ``microbeast.py:109-264`` (``train`` /
``batch_and_learn``), SURVEY.md section 8f-2.

Same structure: CSV headers (``microbeast.py:130-139``), a rollout store and two index queues
(``:161-175``), ``n_actors`` actors that fill buffers (``:179-191``), a learner model with
``Adam(lr=2.5e-4, eps=1e-5)`` (``:195-200``) and the ``batch_and_learn`` loop that appends one
CSV row per update (``:220-244``).  What differs, deliberately:

* the policy and the learner share ONE model on the GPU (the reference back-propagates through
  the actor model while its optimiser holds another model's parameters, so no weight ever
  changes -- quirk Q12); actors are threads, not processes (the GIL is released inside CUDA
  calls and the env here is injected -- gym-microRTS needs a JVM);
* it terminates: ``total_steps`` bounds the loop, actors get the ``None`` sentinel the reference
  checks for but never sends (``microbeast.py:67-68``) and are joined;
* hyper-parameters are arguments with the reference's values as defaults (``microbeast.py:113-121``).
"""
from __future__ import annotations

import csv
import queue
import threading
from time import perf_counter
from typing import Callable, List, Optional

import torch

from . import buffers as mbuf
from .actor import act
from .env_packer import SyntheticVecEnv
from .learner import PPO_learn
from .model import Agent
from .parallel import GradAllReducer


def train(exp_name: str, *, env_factory: Optional[Callable[[int, int, int], object]] = None, n_actors: int = 10,
          n_envs: int = 6, env_size: int = 8, T: int = 64, B: int = 2, total_steps: int = 10_000_000,
          device: Optional[str] = None, lr: float = 2.5e-4, eps: float = 1e-5, seed: int = 0,
          reducer: Optional[GradAllReducer] = None, compact: bool = True, log: bool = False,
          batched_inference: bool = True, rank_seed: int = 0) -> List[list]:
    """Runs actors + learner until ``total_steps`` env-frames have been learned from
    (``step += n_envs*B*T`` per update, ``microbeast.py:230``).  Returns the loss rows that were
    also appended to ``<exp_name>Losses.csv``.

    ``env_factory(actor_id, n_envs, env_size)`` must return a vec-env with the surface
    ``env_packer.Env_Packer`` uses; the default is ``SyntheticVecEnv`` (no JVM here).
    Data-parallel (``microbeast_b200.launch`` under torchrun): every rank calls this with the same ``seed`` (identical
    model replicas), its own ``rank_seed`` (different environments and sampling streams) and a ``reducer``; every rank
    then performs the same number of updates, which the gradient all-reduce inside ``PPO_learn`` requires.
    """
    if device is None:
        device = "cuda"
    dev = torch.device(device)
    if env_factory is None:
        env_factory = lambda a_id, n, size: SyntheticVecEnv(n, size, seed=seed * 1000 + a_id + 100003 * rank_seed)   # noqa: E731
    n_buffers = max(2 * n_actors, B)                                         # microbeast.py:120

    with open(exp_name + ".csv", "w") as f:                                  # microbeast.py:130-133
        csv.writer(f).writerow(["Return", "steps"])
    with open(exp_name + "Losses.csv", "w") as f:                            # microbeast.py:136-139
        csv.writer(f).writerow(["update", "pg_loss", "value_loss", "entropy_loss", "total_loss", "update time"])

    obs_shape = (env_size, env_size, 27)
    nvec = [6, 4, 4, 4, 4, 7, 49] * (env_size * env_size)                    # microbeast.py:145
    torch.manual_seed(seed)
    model = Agent(obs_shape, nvec, env_size ** 2).to(dev)                    # microbeast.py:146,195
    optimizer = torch.optim.Adam(model.parameters(), lr=lr, eps=eps)         # microbeast.py:200
    buffers = mbuf.create_buffers(n_buffers, n_envs, B, T, obs_shape, share=False, compact=compact)  # :161
    free_queue, full_queue = queue.Queue(), queue.Queue()                    # microbeast.py:169-171
    for index in range(n_buffers):
        free_queue.put(index)                                                # microbeast.py:174-175

    # the learner updates the weights in place while actors / the inference server read them: each side holds this
    # lock while it ENQUEUES its kernels (one default stream per device => enqueue order is execution order)
    model_lock = threading.Lock()
    server = None
    if batched_inference:                 # one batched get_action for all actors (inference.py)
        from .inference import InferenceServer
        server = InferenceServer(model, max_batch_envs=n_actors * n_envs, seed=seed + 15485863 * rank_seed, model_lock=model_lock).start()
    threads = []
    for i in range(n_actors):                                                # microbeast.py:179-191
        th = threading.Thread(target=act, daemon=True,
                              args=(model, n_envs, T, buffers, i, free_queue, full_queue, T, exp_name),
                              kwargs=dict(envs=env_factory(i, n_envs, env_size), seed=seed * 7919 + i + 104729 * rank_seed, server=server,
                                          model_lock=model_lock))
        th.start()
        threads.append(th)

    rows, step, n_update = [], 0, 0
    shapes = dict(batch_envs=(n_envs * B, -1))                               # microbeast.py:152-156 (unused there too)
    while step < total_steps:                                                # microbeast.py:220
        start = perf_counter()
        batch = mbuf.get_batch(B, shapes, 0.99, dev, free_queue, full_queue, buffers)    # microbeast.py:226
        with model_lock:
            losses = PPO_learn(model, model, exp_name, batch, optimizer, B, T, n_envs, reducer=reducer)  # :228
        step += n_envs * B * T                                               # microbeast.py:230
        row = [n_update] + losses + [perf_counter() - start]                 # microbeast.py:233-237
        with open(exp_name + "Losses.csv", "a") as f:
            csv.writer(f).writerow(row)                                      # microbeast.py:239
        rows.append(row)
        if log:
            print("Loss list:", row)
        n_update += 1
    for _ in threads:
        free_queue.put(None)              # the sentinel act() stops on (microbeast.py:67-68)
    for th in threads:
        th.join(timeout=30)
    if server is not None:
        server.stop()
    mbuf.release(buffers)
    return rows
