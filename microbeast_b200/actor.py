"""Actor loop with the policy on the GPU -- reference ``microbeast.py:30-105`` (``act``).

Same buffer protocol: pop a free index, write the carried-over frame into slot 0, run
``unroll_length`` steps writing slot ``t+1`` after each env step, push the index to the full
queue.  What changes: ``agent`` lives on the GPU and evaluates all ``n_envs`` environments of
the actor in one ``get_action`` (encoder on cuDNN + one ``mb_head_sample`` launch) instead of
7*cells CPU distribution objects per step; the env is injected (``envs=``) because gym-microRTS
is not installable here (the reference hard-codes ``create_env(8, ...)``, quirk Q15); and the
loop blocks on the queue instead of spinning on ``free_queue.empty()`` (``microbeast.py:59-61``).
"""
from __future__ import annotations

import contextlib
from typing import Optional

import torch

from .env_packer import Env_Packer

AGENT_KEYS = ("action", "policy_logits", "logprobs", "baseline")


def _to_device(env_output: dict, device) -> dict:
    return {k: (v.to(device, non_blocking=True) if torch.is_tensor(v) else v) for k, v in env_output.items()}


def act(agent, n_envs: int, T: int, buffers, a_id: int, free_queue, full_queue, unroll_length: int,
        exp_name: str, envs=None, max_unrolls: Optional[int] = None, seed: Optional[int] = None, server=None,
        model_lock=None):
    """Fills rollout buffers until ``None`` arrives on ``free_queue`` (or ``max_unrolls``).

    ``server``: an ``inference.InferenceServer``; the actor then sends its ``get_action`` requests
    there (batched with the other actors') instead of running the model itself.  ``model_lock``: held while
    this thread enqueues its ``get_action`` kernels, so that they cannot land between the kernels of a learner
    step that updates the same weights in place (``train()`` passes the lock its learner loop takes; all threads
    launch on the device's default stream, so enqueue order is execution order)."""
    if envs is None:
        raise RuntimeError("act(): pass envs= (gym-microRTS is unavailable here; see env_packer.SyntheticVecEnv)")
    packer = envs if isinstance(envs, Env_Packer) else Env_Packer(envs, a_id, exp_name)
    device = next(agent.parameters()).device
    gen = None
    if seed is not None:
        gen = torch.Generator(device=device)
        gen.manual_seed(seed)
    cells = agent.cells

    def policy(env_output):
        if server is not None:
            return server.infer(env_output)
        dev = _to_device({k: env_output[k] for k in ("obs", "action_mask")}, device)
        u = torch.rand((n_envs, 7 * cells), device=device, generator=gen)
        with torch.no_grad(), (model_lock if model_lock is not None else contextlib.nullcontext()):
            out, _ = agent.get_action(dev, uniforms=u)
        return {k: out[k].cpu() for k in AGENT_KEYS}

    env_output = packer.initial()                                        # microbeast.py:46
    agent_output = policy(env_output)                                    # microbeast.py:47
    done_unrolls = 0
    while max_unrolls is None or done_unrolls < max_unrolls:
        index = free_queue.get()                                         # blocking (microbeast.py:59-65)
        if index is None:                                                # microbeast.py:67-68
            break
        for key in env_output:                                           # microbeast.py:73-75
            buffers[key][index][0, ...] = env_output[key]
        for key in agent_output:                                         # microbeast.py:77-78
            buffers[key][index][0, ...] = agent_output[key]
        for t in range(unroll_length):                                   # microbeast.py:82
            agent_output = policy(env_output)                            # microbeast.py:86-87
            env_output = packer.step(agent_output["action"])             # microbeast.py:91
            for key in env_output:                                       # microbeast.py:95-96
                buffers[key][index][t + 1, ...] = env_output[key]
            for key in agent_output:                                     # microbeast.py:99-100
                buffers[key][index][t + 1, ...] = agent_output[key]
        full_queue.put(index)                                            # microbeast.py:105
        done_unrolls += 1
    return done_unrolls
