"""Tensor-format contract of the environment adaptor -- reference ``env_packer.py``.

Only the packaging is in scope: gym-microRTS itself is a Java engine
(reference ``libs/utils.py:59-76``) that cannot be installed here, so ``Env_Packer`` wraps any
object with the vec-env surface it uses (``num_envs``, ``height``, ``reset()``, ``step(a)``,
``get_action_mask()``) and ``SyntheticVecEnv`` provides microRTS-shaped data for tests and
benchmarks.
"""
from __future__ import annotations

import csv

import numpy as np
import torch

from .synth import LOGITS_PER_CELL, NVEC_CELL, OBS_PLANES, SEG_START


def _format_obs(obs: np.ndarray, dtype=torch.float32) -> torch.Tensor:
    """``(n,h,w,27)`` int array -> ``(1,1,n,h,w,27)`` tensor -- reference ``env_packer.py:8-10``
    (float32 there; ``dtype=torch.uint8`` is the wire format of the B200 rollout store)."""
    obs = torch.from_numpy(np.ascontiguousarray(obs)).to(dtype)
    return obs.view((1, 1) + tuple(obs.shape))


def _format_action_masks(action_mask: np.ndarray) -> torch.Tensor:
    """``(n, 78*cells)`` -> ``(1, n, 78*cells)`` uint8, non-zero = valid -- reference ``env_packer.py:12-14``.
    gym-microRTS returns an int32 array and the reference only casts it inside ``CategoricalMasked``
    (``masks.type(torch.BoolTensor)``, ``model.py:40``); the kernels and the rollout store take one byte per
    entry, so the cast happens here, once, for every integer / bool dtype."""
    action_mask = torch.from_numpy(np.ascontiguousarray(action_mask))
    if action_mask.dtype != torch.uint8:
        action_mask = (action_mask != 0).to(torch.uint8)
    return action_mask.view((1,) + tuple(action_mask.shape))


class Env_Packer:
    """Packages the outputs of a vec-env for the rollout buffers -- reference
    ``env_packer.py:17-111``: same keys, shapes and episode CSV rows
    ``[return, steps, env_idx]`` (``env_packer.py:69-75``).

    Deviation (quirk Q16): the reference's episode accumulators are uint8 tensors, so
    ``ep_step`` wraps at 256 although episodes run 225-375 steps; here they are float32 /
    int32.  They are logged only and never reach the learner.
    """

    def __init__(self, envs, a_id: int, exp_name: str, obs_dtype=torch.float32):
        self.envs = envs
        self.n_envs = self.envs.num_envs
        self.ep_return = None
        self.ep_step = None
        self.a_id = a_id
        self.exp_name = exp_name
        self.obs_dtype = obs_dtype

    def initial(self) -> dict:
        env_h = self.envs.height
        self.ep_return = torch.zeros(1, self.n_envs, dtype=torch.float32)
        self.ep_step = torch.zeros(1, self.n_envs, dtype=torch.int32)
        return dict(
            obs=_format_obs(self.envs.reset(), self.obs_dtype),
            reward=torch.zeros(1, self.n_envs),
            done=torch.zeros(1, self.n_envs, dtype=torch.uint8),
            ep_return=self.ep_return,
            ep_step=self.ep_step,
            last_action=torch.zeros(1, self.n_envs, 7 * (env_h ** 2), dtype=torch.int64),
            action_mask=_format_action_masks(self.envs.get_action_mask().reshape(self.n_envs, -1)),
        )

    def step(self, action: torch.Tensor) -> dict:
        act = action.detach().cpu().numpy() if torch.is_tensor(action) else np.asarray(action)
        obs, reward, done, _info = self.envs.step(act)
        self.ep_step += 1
        self.ep_return += torch.as_tensor(np.asarray(reward), dtype=torch.float32).view(1, self.n_envs)
        ep_step, ep_return = self.ep_step.clone(), self.ep_return.clone()
        finished = np.flatnonzero(np.asarray(done))
        if len(finished) and self.exp_name:
            with open(self.exp_name + ".csv", "a") as f:
                w = csv.writer(f)
                for i in finished:
                    w.writerow([self.ep_return[0, i].item(), self.ep_step[0, i].item(), int(i)])
        for i in finished:
            self.ep_step[0, i] = 0
            self.ep_return[0, i] = 0
        return dict(
            obs=_format_obs(obs, self.obs_dtype),
            reward=torch.as_tensor(np.asarray(reward), dtype=torch.float32).view(1, self.n_envs),
            done=torch.as_tensor(np.asarray(done).astype("uint8")).float().view(1, self.n_envs),
            ep_return=ep_return,
            ep_step=ep_step,
            last_action=action,
            action_mask=_format_action_masks(self.envs.get_action_mask().reshape(self.n_envs, -1)),
        )

    def reset(self) -> dict:
        return self.initial()

    def close(self) -> None:
        if hasattr(self.envs, "close"):
            self.envs.close()


class SyntheticVecEnv:
    """microRTS-shaped stand-in for ``MicroRTSGridModeVecEnv`` (shapes and dtypes of SURVEY 8d):
    obs ``(n,h,w,27)`` int32 one-hot planes, masks ``(n, h*w, 78)`` with sparse active cells,
    sparse rewards, episodes of a few hundred steps.  It checks that submitted actions respect
    the mask it handed out, which is what the Java engine would reject."""

    def __init__(self, num_envs: int, size: int, seed: int = 0, active_p: float = 0.06, episode_len: int = 250):
        self.num_envs = num_envs
        self.height = self.width = size
        self.rs = np.random.RandomState(seed)
        self.active_p = active_p
        self.episode_len = episode_len
        self._mask = None
        self.invalid_actions = 0
        self.steps = 0

    def _obs(self):
        n, h, w = self.num_envs, self.height, self.width
        obs = np.zeros((n, h, w, sum(OBS_PLANES)), np.int32)
        base = 0
        for width in OBS_PLANES:
            idx = np.where(self.rs.rand(n, h, w) < 0.9, 0, self.rs.randint(1, width, (n, h, w))) + base
            np.put_along_axis(obs, idx[..., None], 1, axis=3)
            base += width
        return obs

    def _new_mask(self):
        n, cells = self.num_envs, self.height * self.width
        p = np.empty(LOGITS_PER_CELL)
        for k, (s, wd) in enumerate(zip(SEG_START, NVEC_CELL)):
            p[s:s + wd] = (0.5, 0.5, 0.5, 0.5, 0.5, 0.3, 0.05)[k]
        m = self.rs.rand(n, cells, LOGITS_PER_CELL) < p
        m[:, :, 0] = True
        m &= self.rs.rand(n, cells, 1) < self.active_p
        self._mask = m.astype(np.uint8)

    def reset(self):
        self._new_mask()
        return self._obs()

    def get_action_mask(self):
        return self._mask

    def step(self, action):
        action = np.asarray(action).reshape(self.num_envs, self.height * self.width, 7)
        for k, (s, wd) in enumerate(zip(SEG_START, NVEC_CELL)):
            seg = self._mask[:, :, s:s + wd]
            a = action[:, :, k]
            ok = (a >= 0) & (a < wd)
            picked = np.take_along_axis(seg, np.clip(a, 0, wd - 1)[:, :, None], 2)[:, :, 0] != 0
            self.invalid_actions += int((~ok | (~picked & seg.any(2))).sum())
        self.steps += 1
        reward = np.where(self.rs.rand(self.num_envs) < 0.1, self.rs.standard_normal(self.num_envs), 0.0)
        done = self.rs.rand(self.num_envs) < 1.0 / self.episode_len
        self._new_mask()
        return self._obs(), reward.astype(np.float32), done, [{} for _ in range(self.num_envs)]

    def close(self):
        pass
