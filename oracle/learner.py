"""Oracle for the whole reference learner step on CPU.  TEST INFRASTRUCTURE ONLY.

``RefAgent`` restates the reference network (``model.py:56-137``) and its
``get_action`` (``model.py:165-216``) including the per-component Python loop;
``ref_learn_step`` restates ``PPO_learn`` (``libs/utils.py:223-342``).  Parameter
names equal the reference's so a reference ``state_dict`` loads unchanged.
This is what ``bench.py``'s ``cpu_baseline`` / ``--impl reference`` legs time
(the reference itself is Python and cannot travel to the GPU box).
See ``oracle/__init__.py`` for the rules on who may import this.
"""
from __future__ import annotations

import torch
import torch.nn as nn
import torch.nn.functional as F

from . import NVEC_CELL, LOGITS_PER_CELL
from .head import MaskedCategorical
from .vtrace import vtrace_losses


def _ortho(layer: nn.Linear, std: float) -> nn.Linear:
    nn.init.orthogonal_(layer.weight, std)          # model.py:24-27
    nn.init.constant_(layer.bias, 0.0)
    return layer


class _Res(nn.Module):                                # model.py:56-73
    def __init__(self, ch):
        super().__init__()
        self.conv0 = nn.Conv2d(ch, ch, 3, padding=1)
        self.conv1 = nn.Conv2d(ch, ch, 3, padding=1)

    def forward(self, x):
        return self.conv1(F.relu(self.conv0(F.relu(x)))) + x


class _Stage(nn.Module):                              # model.py:77-109
    def __init__(self, cin, cout):
        super().__init__()
        self.conv = nn.Conv2d(cin, cout, 3, padding=1)
        self.res_block0 = _Res(cout)
        self.res_block1 = _Res(cout)

    def forward(self, x):
        x = F.max_pool2d(self.conv(x), kernel_size=3, stride=2, padding=1)
        return self.res_block1(self.res_block0(x))


class RefAgent(nn.Module):
    def __init__(self, obs_space_shape, nvec, mapsize=64):
        super().__init__()
        h, w, c = obs_space_shape                     # model.py:115
        self.nvec = list(nvec)
        self.mapsize = mapsize
        stages, cin = [], c
        for cout in (16, 32, 32):                     # model.py:121-124
            stages.append(_Stage(cin, cout))
            cin, h, w = cout, (h + 1) // 2, (w + 1) // 2
        self.network = nn.Sequential(*stages, nn.Flatten(), nn.ReLU(),
                                     nn.Linear(cin * h * w, 256), nn.ReLU())   # model.py:126-133
        self.actor = _ortho(nn.Linear(256, sum(self.nvec)), 0.0)              # model.py:136
        self.critic = _ortho(nn.Linear(256, 1), 1.0)                          # model.py:137

    def forward(self, input_dict, learning):          # model.py:152-160
        obs = input_dict["obs"] if learning else input_dict["obs"][0, 0, ...]
        return self.network(obs.permute((0, 3, 1, 2)))

    def get_value(self, input_dict, learning):        # model.py:219-220
        return self.critic(self.forward(input_dict, learning))

    def get_action(self, input_dict, learning=False, inds=(), agent_state=()):
        logits = self.actor(self.forward(input_dict, learning))               # model.py:167
        split_logits = torch.split(logits, self.nvec, dim=1)                  # model.py:168
        if not learning:
            raise NotImplementedError("oracle actor branch: use oracle.head.head_sample")
        action = input_dict["action"].transpose(1, 0)                         # model.py:183-186
        action_mask = input_dict["action_mask"]                               # model.py:188
        split_mask = torch.split(action_mask, self.nvec, dim=1)               # model.py:189
        cats = [MaskedCategorical(l, m) for l, m in zip(split_logits, split_mask)]  # model.py:190
        logprob = torch.stack([c.log_prob(a) for a, c in zip(action, cats)])  # model.py:194
        entropy = torch.stack([c.entropy() for c in cats])                    # model.py:196
        k = len(self.nvec)
        logprob = logprob.T.view(-1, k)                                       # model.py:199
        entropy = entropy.T.view(-1, k)                                       # model.py:200
        action = action.T.view(-1, k)                                         # model.py:201
        baseline = self.get_value(input_dict, learning).view(1, -1)           # model.py:205 (2nd pass, Q6)
        return dict(action=action, policy_logits=logits, logprobs=logprob.sum(1),
                    baseline=baseline, entropy=entropy.sum(1)), ()            # model.py:213-214


def ref_learn_step(model: RefAgent, batch: dict, optimizer, B: int, T: int, n_envs: int,
                   return_intermediates: bool = False):
    """``PPO_learn`` (``libs/utils.py:223-342``) without printing / weight self-copy.

    ``batch`` holds flat ``[N, ...]`` tensors, ``N = B*n_envs*(T+1)`` (Appendix B).
    Returns ``[pg, value, entropy, total]`` python floats (``utils.py:342``).
    """
    rows = B * n_envs
    out, _ = model.get_action(batch, learning=True)                           # utils.py:234
    out["baseline"] = out["baseline"].flatten()                               # utils.py:237
    out = {k: v.reshape((rows, T + 1) + v.shape[1:]) for k, v in out.items()}         # utils.py:243-246
    bt = {k: v.reshape((rows, T + 1) + v.shape[1:]) for k, v in batch.items()}        # utils.py:251-254
    bootstrap_value = out["baseline"][:, -1]                                  # utils.py:258
    bt = {k: v[:, 1:] for k, v in bt.items()}                                 # utils.py:263
    out = {k: v[:, :-1] for k, v in out.items()}                              # utils.py:265
    res = vtrace_losses(bt["logprobs"], out["logprobs"], out["baseline"], bootstrap_value,
                        bt["reward"], bt["done"], out["entropy"])             # utils.py:277-329
    optimizer.zero_grad()                                                     # utils.py:333
    res["total_loss"].backward()                                              # utils.py:334
    optimizer.step()                                                          # utils.py:335
    four = [res["pg_loss"].item(), res["value_loss"].item(),
            res["entropy_loss"].item(), res["total_loss"].item()]             # utils.py:342
    if return_intermediates:
        return four, res, out
    return four


def nvec_for(cells: int):
    return list(NVEC_CELL) * cells


assert sum(NVEC_CELL) == LOGITS_PER_CELL
