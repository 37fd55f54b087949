"""Oracle for the masked multi-discrete policy head.  TEST INFRASTRUCTURE ONLY.

Restates reference ``model.py:33-52`` (``CategoricalMasked``) and the learner /
actor branches of ``model.py:165-216`` (``Agent.get_action``) on CPU torch.
See ``oracle/__init__.py`` for the rules on who may import this.
"""
from __future__ import annotations

import numpy as np
import torch
from torch.distributions.categorical import Categorical

from . import NVEC_CELL, LOGITS_PER_CELL, COMPONENTS_PER_CELL

_SEG_START = np.concatenate([[0], np.cumsum(NVEC_CELL)[:-1]]).tolist()


class MaskedCategorical(Categorical):
    """One masked categorical component.

    Follows reference ``model.py:33-52``: invalid entries get the fp32 sentinel
    ``-1e8`` (``model.py:43``) before ``Categorical.__init__`` normalises the
    logits; ``entropy`` sums ``logits * probs`` over valid entries only
    (``model.py:46-52``).  The mask is converted with ``.bool()`` -- on CPU this
    is the same conversion as the reference's ``.type(torch.BoolTensor)``
    (``model.py:40``): non-zero means valid.
    """

    def __init__(self, logits: torch.Tensor, masks: torch.Tensor):
        self.masks = masks.bool()                                           # model.py:40
        fill = torch.tensor(-1e8)                                           # model.py:43 (fp32 0-dim)
        super().__init__(logits=torch.where(self.masks, logits, fill))      # model.py:43-44

    def entropy(self) -> torch.Tensor:
        plogp = self.logits * self.probs                                    # model.py:50
        plogp = torch.where(self.masks, plogp, torch.tensor(0.0))           # model.py:51
        return -plogp.sum(-1)                                               # model.py:52


def _nvec(cells: int):
    return list(NVEC_CELL) * cells


def head_eval_loop(logits: torch.Tensor, mask: torch.Tensor, action: torch.Tensor):
    """Learner branch of ``Agent.get_action`` after the actor Linear.

    ``logits [N,78C]`` f32 (may require grad), ``mask [N,78C]`` u8,
    ``action [N,7C]`` i64  ->  ``(logprobs [N], entropy [N])``.
    One distribution object per component, exactly like the reference
    (``model.py:168`` split, ``:186`` transpose, ``:188-190`` objects,
    ``:194`` log_prob, ``:196`` entropy, ``:199-201`` ``.T.view``, ``:214`` sums).
    This is also what the CPU baseline times.
    """
    n, width = logits.shape
    cells = width // LOGITS_PER_CELL
    nvec = _nvec(cells)
    split_logits = torch.split(logits, nvec, dim=1)                         # model.py:168
    action_t = action.transpose(1, 0)                                       # model.py:186
    split_mask = torch.split(mask, nvec, dim=1)                             # model.py:189
    cats = [MaskedCategorical(l, m) for l, m in zip(split_logits, split_mask)]   # model.py:190
    logprob = torch.stack([c.log_prob(a) for a, c in zip(action_t, cats)])  # model.py:194
    entropy = torch.stack([c.entropy() for c in cats])                      # model.py:196
    k = len(nvec)
    logprob = logprob.T.view(-1, k)                                         # model.py:199
    entropy = entropy.T.view(-1, k)                                         # model.py:200
    return logprob.sum(1), entropy.sum(1)                                   # model.py:214


def head_eval_components(logits: torch.Tensor, mask: torch.Tensor, action: torch.Tensor):
    """Same arithmetic as :func:`head_eval_loop`, batched per component *type*.

    The seven component types of all cells are evaluated as ``[N, C, n_k]``
    batches through the same torch ops, so per-component values are the ones
    the loop produces; returns ``(logp [N,7C], entropy [N,7C])`` in the
    reference's cell-major component order.  Used for larger test shapes where
    ``7C`` Python objects per call would take minutes.
    """
    n, width = logits.shape
    cells = width // LOGITS_PER_CELL
    lg = logits.view(n, cells, LOGITS_PER_CELL)
    mk = mask.view(n, cells, LOGITS_PER_CELL)
    ac = action.view(n, cells, COMPONENTS_PER_CELL)
    lps, ents = [], []
    for k, (s, w) in enumerate(zip(_SEG_START, NVEC_CELL)):
        cat = MaskedCategorical(lg[:, :, s:s + w], mk[:, :, s:s + w])
        lps.append(cat.log_prob(ac[:, :, k]))
        ents.append(cat.entropy())
    logp = torch.stack(lps, dim=2).reshape(n, cells * COMPONENTS_PER_CELL)
    ent = torch.stack(ents, dim=2).reshape(n, cells * COMPONENTS_PER_CELL)
    return logp, ent


def head_eval_fast(logits, mask, action):
    """``(logprobs [N], entropy [N])`` via :func:`head_eval_components`."""
    logp, ent = head_eval_components(logits, mask, action)
    return logp.sum(1), ent.sum(1)


def head_probs64(logits: torch.Tensor, mask: torch.Tensor):
    """fp64 probabilities of every component, ``[N, C, 78]`` float64 numpy.

    ``CategoricalMasked(logits=z.double(), masks=m).probs`` per component type
    (reference class semantics, fp64 arithmetic).
    """
    n, width = logits.shape
    cells = width // LOGITS_PER_CELL
    lg = logits.detach().double().view(n, cells, LOGITS_PER_CELL)
    mk = mask.view(n, cells, LOGITS_PER_CELL)
    out = np.empty((n, cells, LOGITS_PER_CELL), dtype=np.float64)
    for s, w in zip(_SEG_START, NVEC_CELL):
        cat = MaskedCategorical(lg[:, :, s:s + w], mk[:, :, s:s + w])
        out[:, :, s:s + w] = cat.probs.numpy()
    return out


def sample_invcdf(logits: torch.Tensor, mask: torch.Tensor, uniforms: torch.Tensor) -> torch.Tensor:
    """DEFINITION of the inverse-CDF sampler (the reference has none, see D4).

    For every component with ``n`` entries, fp32 uniform ``u`` in ``[0,1)``:

    1. ``p = CategoricalMasked(logits.double(), mask).probs``  (fp64);
    2. ``cdf_k = p_0 + ... + p_k`` accumulated left to right in fp64;
    3. ``a = #{k < n-1 : double(u) >= cdf_k}``;
    4. if the component has at least one valid entry and ``mask[a] == 0`` (only
       reachable through the ``n-1`` cap when rounding leaves ``cdf`` short of
       ``u``), ``a`` becomes the last valid index.

    A fully masked component has ``p = 1/n`` (quirk Q1) and so samples
    uniformly over all ``n`` entries, like the reference's ``.sample()``.
    ``uniforms [N,7C]`` f32 -> ``action [N,7C]`` i64.
    """
    n, width = logits.shape
    cells = width // LOGITS_PER_CELL
    p = head_probs64(logits, mask)
    mk = mask.view(n, cells, LOGITS_PER_CELL).numpy() != 0
    u = uniforms.detach().double().view(n, cells, COMPONENTS_PER_CELL).numpy()
    act = np.empty((n, cells, COMPONENTS_PER_CELL), dtype=np.int64)
    for k, (s, w) in enumerate(zip(_SEG_START, NVEC_CELL)):
        cdf = np.cumsum(p[:, :, s:s + w], axis=2)            # sequential fp64 prefix sum
        a = (u[:, :, k:k + 1] >= cdf[:, :, :w - 1]).sum(axis=2)
        m = mk[:, :, s:s + w]
        any_valid = m.any(axis=2)
        last_valid = (w - 1) - np.argmax(m[:, :, ::-1], axis=2)
        landed_valid = np.take_along_axis(m, a[:, :, None], axis=2)[:, :, 0]
        a = np.where(any_valid & ~landed_valid, last_valid, a)
        act[:, :, k] = a
    return torch.from_numpy(act.reshape(n, cells * COMPONENTS_PER_CELL))


def sample_expdiv(logits: torch.Tensor, mask: torch.Tensor, exponentials: torch.Tensor) -> torch.Tensor:
    """The reference's sampling rule with the randomness made explicit: ``Categorical.sample()`` (``model.py:177``)
    is ``torch.multinomial(probs, 1, True)``, which draws ``q ~ Exp(1)`` per category and returns
    ``argmax(probs / q)`` (``tests/test_oracle_golden.py`` checks that this reproduces
    ``torch.multinomial`` bit for bit on the CPU with the same generator state).
    ``exponentials [N, 78C]`` f32 -> ``action [N, 7C]`` i64; fp32 arithmetic on the reference class's ``probs``."""
    n, width = logits.shape
    cells = width // LOGITS_PER_CELL
    lg = logits.view(n, cells, LOGITS_PER_CELL)
    mk = mask.view(n, cells, LOGITS_PER_CELL)
    q = exponentials.view(n, cells, LOGITS_PER_CELL)
    acts = []
    for s, w in zip(_SEG_START, NVEC_CELL):
        p = MaskedCategorical(lg[:, :, s:s + w], mk[:, :, s:s + w]).probs
        acts.append((p / q[:, :, s:s + w]).argmax(dim=2))
    return torch.stack(acts, dim=2).reshape(n, cells * COMPONENTS_PER_CELL)


def head_sample(logits, mask, uniforms):
    """Actor branch (``model.py:171-177,194,199-210``) with the defined sampler.

    Returns ``(action [N,7C] i64, logprobs [N] f32, entropy [N] f32)``;
    log-prob / entropy are the fp32 reference arithmetic at the sampled action.
    """
    action = sample_invcdf(logits, mask, uniforms)
    logp, ent = head_eval_fast(logits, mask, action)
    return action, logp, ent


def head_eval_grad(logits, mask, action, g_logp, g_ent, loop=False):
    """dlogits via torch autograd through the reference arithmetic."""
    z = logits.detach().clone().requires_grad_(True)
    fn = head_eval_loop if loop else head_eval_fast
    logp, ent = fn(z, mask, action)
    (logp * g_logp).sum().add((ent * g_ent).sum()).backward()
    return z.grad
