"""Oracle for V-trace targets/advantages and the loss reduction.  TEST INFRASTRUCTURE ONLY.

Line-by-line restatement of reference ``libs/utils.py:277-329`` (the inline block
of ``PPO_learn``), returning the intermediates the reference throws away.
See ``oracle/__init__.py`` for the rules on who may import this.
"""
from __future__ import annotations

import torch


def vtrace_losses(old_logprobs, new_logprobs, values, bootstrap_value, rewards, done, entropy,
                  gamma: float = 0.99, rho_bar: float = 1.0, c_bar: float = 1.0,
                  pg_rho_bar: float = 1.0, entropy_coef: float = 0.01):
    """All tensors ``[R, T]`` (time contiguous, D2) except ``bootstrap_value [R]``.

    ``new_logprobs``, ``values`` and ``entropy`` may require grad; the V-trace
    block itself runs under ``no_grad`` exactly like the reference.  The four
    constants default to the reference's literals (Q10).  Returns a dict with
    ``vs``, ``pg_advantages`` (no grad) and the four loss tensors.
    """
    rows = values.shape[0]
    discounts = (~done).float() * gamma                                      # utils.py:277

    ratio = new_logprobs - old_logprobs                                      # utils.py:285
    with torch.no_grad():                                                    # utils.py:289
        ratio_exp = ratio.exp()                                              # utils.py:290
        rhos = torch.clamp(ratio_exp, max=rho_bar)                           # utils.py:293
        cs = torch.clamp(ratio_exp, max=c_bar)                               # utils.py:294
        values_t_plus_one = torch.cat(
            [values[:, 1:], bootstrap_value.view((rows, 1))], dim=1)         # utils.py:296
        deltas = rhos * (rewards + discounts * values_t_plus_one - values)   # utils.py:298

        acc = torch.zeros((rows,))                                           # utils.py:300
        result = []
        for t in range(discounts.shape[1] - 1, -1, -1):                      # utils.py:302
            acc = deltas[:, t] + discounts[:, t] * cs[:, t] * acc            # utils.py:303
            result.append(acc)                                               # utils.py:304
        result.reverse()                                                     # utils.py:306
        vs_minus_v_xs = torch.stack(result).transpose(1, 0)                  # utils.py:307
        vs = vs_minus_v_xs + values                                          # utils.py:310

        broadcasted = torch.ones_like(vs[:, 0]) * bootstrap_value            # utils.py:313
        vs_t_plus_1 = torch.cat([vs[:, 1:], broadcasted.view((rows, 1))], dim=1)   # utils.py:314
        clipped_pg_rhos = torch.clamp(rhos, max=pg_rho_bar)                  # utils.py:316
        pg_advantages = clipped_pg_rhos * (rewards + discounts * vs_t_plus_1 - values)  # utils.py:318

    pg_loss = (new_logprobs * pg_advantages).mean()                          # utils.py:320
    value_loss = 0.5 * ((vs - values) ** 2).mean()                           # utils.py:323
    entropy_loss = entropy.mean()                                            # utils.py:326
    total_loss = pg_loss + value_loss - entropy_coef * entropy_loss          # utils.py:329
    return dict(vs=vs, pg_advantages=pg_advantages, pg_loss=pg_loss, value_loss=value_loss,
                entropy_loss=entropy_loss, total_loss=total_loss)


def vtrace_losses_and_grads(old_logprobs, new_logprobs, values, bootstrap_value, rewards, done,
                            entropy, **kw):
    """Runs :func:`vtrace_losses` and autograd; adds ``d_logp``, ``d_values``, ``d_entropy``."""
    lp = new_logprobs.detach().clone().requires_grad_(True)
    v = values.detach().clone().requires_grad_(True)
    e = entropy.detach().clone().requires_grad_(True)
    out = vtrace_losses(old_logprobs, lp, v, bootstrap_value.detach(), rewards, done, e, **kw)
    out["total_loss"].backward()
    out["d_logp"], out["d_values"], out["d_entropy"] = lp.grad, v.grad, e.grad
    return {k: (t.detach() if torch.is_tensor(t) else t) for k, t in out.items()}
