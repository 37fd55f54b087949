"""Measured parity of one learner step: GPU fp32 and CPU fp32 against an fp64 evaluation.  TEST INFRASTRUCTURE ONLY.

The reference runs in fp32 on the CPU (``libs/utils.py:223-342``).  The product runs in fp32 on the GPU.  Two fp32
evaluations of the same 15-layer encoder differ by their summation orders, so "GPU equals CPU to 1e-5" is only
meaningful next to how far EACH of them is from the exact value.  This module computes that exact value: the same
function the reference evaluates, in float64 --

* encoder / actor / critic: the ``RefAgent`` weights promoted to float64 (``oracle/learner.py``);
* policy head: masked log-softmax, log-prob gather and masked entropy per component (``model.py:33-52,165-216``) in
  float64, with the one property of the fp32 reference that is NOT a rounding error kept on purpose: a component
  with no valid entry contributes exactly 0 log-prob, 0 entropy and 0 gradient (quirk Q1 -- in fp32 the ``-1e8``
  sentinel absorbs ``ln n``; evaluated naively in float64 it would contribute ``-ln n`` instead);
* V-trace and the three losses: ``oracle.vtrace.vtrace_losses`` on float64 tensors (its ``gamma`` stays the fp32
  constant ``0.99f``, as in the reference and in the kernel).

``step_report`` returns, for every quantity north_star names (log-prob, entropy, V-trace targets, advantages, the
four losses) and for the parameter gradients, the error of the GPU step and of the CPU fp32 oracle against that
float64 value and against each other, all in the norm of SURVEY 7.3-2: ``max |x - ref| / max(|ref|, rms(ref))``.
Users: ``tests/test_gpu_parity.py`` and the parity gate of ``bench.py``.
"""
from __future__ import annotations

import copy
from typing import Dict

import torch

from . import COMPONENTS_PER_CELL, LOGITS_PER_CELL, NVEC_CELL
from .learner import RefAgent, ref_learn_step
from .vtrace import vtrace_losses

_SEG_START = (0, 6, 10, 14, 18, 22, 29)


def head_eval_exact64(logits: torch.Tensor, mask: torch.Tensor, action: torch.Tensor):
    """``(logprobs [N], entropy [N])`` in float64, differentiable wrt ``logits`` (float64)."""
    n, width = logits.shape
    cells = width // LOGITS_PER_CELL
    lg = logits.view(n, cells, LOGITS_PER_CELL)
    mk = mask.view(n, cells, LOGITS_PER_CELL) != 0
    ac = action.view(n, cells, COMPONENTS_PER_CELL)
    logp = logits.new_zeros(n)
    ent = logits.new_zeros(n)
    zero = logits.new_zeros(())
    for k, (s, w) in enumerate(zip(_SEG_START, NVEC_CELL)):
        m = mk[:, :, s:s + w]
        z = torch.where(m, lg[:, :, s:s + w], logits.new_tensor(-1e8))          # model.py:43
        l = z - torch.logsumexp(z, dim=-1, keepdim=True)                        # Categorical.__init__
        p = l.exp()                                                             # Categorical.probs
        lp = l.gather(2, ac[:, :, k:k + 1]).squeeze(2)                          # log_prob, model.py:194
        h = -torch.where(m, l * p, zero).sum(-1)                                # model.py:46-52
        alive = m.any(-1)                                                       # quirk Q1 (see module docstring)
        logp = logp + torch.where(alive, lp, zero).sum(1)                       # model.py:214
        ent = ent + torch.where(alive, h, zero).sum(1)
    return logp, ent


def _shift(x, rows, T):
    return x.reshape((rows, T + 1) + tuple(x.shape[1:]))


def truth64(ref_agent: RefAgent, batch: Dict[str, torch.Tensor], rows: int, T: int) -> Dict[str, object]:
    """One learner step (``utils.py:234-334``, no optimiser) in float64; returns values and parameter gradients."""
    agent = copy.deepcopy(ref_agent).double()
    for p in agent.parameters():
        p.grad = None
    obs = batch["obs"].double()
    hidden = agent.network(obs.permute((0, 3, 1, 2)))
    logits = agent.actor(hidden)
    baseline = agent.critic(hidden).flatten()
    logp, ent = head_eval_exact64(logits, batch["action_mask"], batch["action"])
    lp, base, en = _shift(logp, rows, T), _shift(baseline, rows, T), _shift(ent, rows, T)       # utils.py:237-253
    res = vtrace_losses(_shift(batch["logprobs"].double(), rows, T)[:, 1:], lp[:, :-1], base[:, :-1], base[:, -1],
                        _shift(batch["reward"].double(), rows, T)[:, 1:], _shift(batch["done"], rows, T)[:, 1:],
                        en[:, :-1])                                                             # utils.py:258-329
    res["total_loss"].backward()
    return dict(logprobs=logp.detach(), entropy=ent.detach(), baseline=baseline.detach(), vs=res["vs"].detach(),
                pg_advantages=res["pg_advantages"].detach(),
                loss4=torch.stack([res[k].detach() for k in ("pg_loss", "value_loss", "entropy_loss", "total_loss")]),
                grads={n: p.grad.detach() for n, p in agent.named_parameters()})


def oracle32(ref_agent: RefAgent, batch: Dict[str, torch.Tensor], rows: int, T: int) -> Dict[str, object]:
    """The same step through the fp32 restatement of the reference (``oracle.learner.ref_learn_step``, lr = 0)."""
    agent = copy.deepcopy(ref_agent)
    opt = torch.optim.SGD(agent.parameters(), lr=0.0)
    with torch.no_grad():
        full, _ = agent.get_action(batch, learning=True)
    four, res, _out = ref_learn_step(agent, {k: v.clone() for k, v in batch.items()}, opt, rows, T, 1,
                                     return_intermediates=True)
    return dict(logprobs=full["logprobs"].detach(), entropy=full["entropy"].detach(),
                baseline=full["baseline"].flatten().detach(), vs=res["vs"].detach(),
                pg_advantages=res["pg_advantages"].detach(), loss4=torch.tensor(four, dtype=torch.float64),
                grads={n: p.grad.detach() for n, p in agent.named_parameters()})


def err_norm(x, ref) -> float:
    """``max |x - ref| / max(|ref|, rms(ref))``; 0 for empty tensors."""
    x = torch.as_tensor(x).detach().double().cpu().flatten()
    ref = torch.as_tensor(ref).detach().double().cpu().flatten()
    if ref.numel() == 0:
        return 0.0
    rms = ref.pow(2).mean().sqrt()
    scale = torch.maximum(ref.abs(), rms.expand_as(ref)).clamp_min(1e-300)
    return float(((x - ref).abs() / scale).max())


def loss_rel(x, ref, floor) -> list:
    """Relative error of the four losses; pg and total are sums of signed terms, so their denominator is at least
    ``floor`` = mean |logp * adv| (the same rule as ``tests/test_gpu_vtrace.py``)."""
    x = torch.as_tensor(x).double().cpu()
    ref = torch.as_tensor(ref).double().cpu()
    fl = torch.tensor([floor, 0.0, 0.0, floor], dtype=torch.float64)
    return [float(v) for v in ((x - ref).abs() / torch.maximum(ref.abs(), fl).clamp_min(1e-300))]


def step_report(ours: Dict[str, object], ref_agent: RefAgent, batch: Dict[str, torch.Tensor], rows: int, T: int):
    """``ours``: the GPU step's ``logprobs, entropy, baseline [N]``, ``vs, pg_advantages [R,T]``, ``loss4 [4]`` and
    ``grads {name: tensor}``.  Returns the parity dictionary described in the module docstring."""
    t64 = truth64(ref_agent, batch, rows, T)
    c32 = oracle32(ref_agent, batch, rows, T)
    lp_adv = float((t64["pg_advantages"] * _shift(t64["logprobs"], rows, T)[:, :-1]).abs().mean())
    rep = {"rows": rows, "T": T, "norm": "max |x-ref| / max(|ref|, rms(ref)); losses: relative",
           "gpu_vs_fp64": {}, "cpu_fp32_vs_fp64": {}, "gpu_vs_cpu_fp32": {}}
    for key in ("logprobs", "entropy", "baseline", "vs", "pg_advantages"):
        rep["gpu_vs_fp64"][key] = err_norm(ours[key], t64[key])
        rep["cpu_fp32_vs_fp64"][key] = err_norm(c32[key], t64[key])
        rep["gpu_vs_cpu_fp32"][key] = err_norm(ours[key], c32[key])
    rep["gpu_vs_fp64"]["loss4_rel"] = loss_rel(ours["loss4"], t64["loss4"], lp_adv)
    rep["cpu_fp32_vs_fp64"]["loss4_rel"] = loss_rel(c32["loss4"], t64["loss4"], lp_adv)
    rep["gpu_vs_cpu_fp32"]["loss4_rel"] = loss_rel(ours["loss4"], c32["loss4"], lp_adv)
    for side, a, b in (("gpu_vs_fp64", ours, t64), ("cpu_fp32_vs_fp64", c32, t64), ("gpu_vs_cpu_fp32", ours, c32)):
        worst, name_worst = 0.0, ""
        for name, g in b["grads"].items():
            e = err_norm(a["grads"][name], g)
            if e >= worst:
                worst, name_worst = e, name
        rep[side]["grad_max"] = worst
        rep[side]["grad_max_param"] = name_worst
    return rep
