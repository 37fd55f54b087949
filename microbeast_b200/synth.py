"""Synthetic microRTS-shaped rollouts.

gym-microRTS (the Java engine behind reference ``libs/utils.py:59-76``) is not
installable here, so every config of BASELINE.json runs on synthetic tensors
with the shapes and dtypes of the reference rollout buffers
(``libs/utils.py:34-46``).  Generation uses only torch ops, so it runs on the
device for throughput runs and on the CPU for parity fixtures.
"""
from __future__ import annotations

import torch

NVEC_CELL = (6, 4, 4, 4, 4, 7, 49)          # reference microbeast.py:145
SEG_START = (0, 6, 10, 14, 18, 22, 29)
LOGITS_PER_CELL = 78
COMPONENTS_PER_CELL = 7
OBS_PLANES = (5, 5, 3, 8, 6)                # gym-microRTS one-hot groups, 27 planes

# probability that an entry of a component is valid inside an active cell
_VALID_P = (0.5, 0.5, 0.5, 0.5, 0.5, 0.3, 0.05)


def _gen(seed: int, device) -> torch.Generator:
    g = torch.Generator(device=device)
    g.manual_seed(int(seed))
    return g


def make_masks(n: int, cells: int, preset: str = "dense", *, gen: torch.Generator,
               device="cpu") -> torch.Tensor:
    """``action_mask [n, 78*cells]`` u8.

    presets: ``dense`` (6 % of cells active), ``realistic`` (2 % active, most
    components single-valid), ``all_valid``, ``all_masked``, ``one_valid``.
    """
    shape = (n, cells, LOGITS_PER_CELL)
    if preset == "all_valid":
        return torch.ones((n, cells * LOGITS_PER_CELL), dtype=torch.uint8, device=device)
    if preset == "all_masked":
        return torch.zeros((n, cells * LOGITS_PER_CELL), dtype=torch.uint8, device=device)
    r = torch.rand(shape, generator=gen, device=device)
    if preset == "one_valid":
        m = torch.zeros(shape, dtype=torch.uint8, device=device)
        for s, w in zip(SEG_START, NVEC_CELL):
            idx = r[:, :, s:s + w].argmax(dim=2, keepdim=True)
            m[:, :, s:s + w].scatter_(2, idx, 1)
        return m.view(n, -1)
    if preset not in ("dense", "realistic"):
        raise ValueError(f"unknown mask preset {preset!r}")
    active_p = 0.06 if preset == "dense" else 0.02
    thresh = torch.empty(LOGITS_PER_CELL, device=device)
    for k, (s, w) in enumerate(zip(SEG_START, NVEC_CELL)):
        thresh[s:s + w] = _VALID_P[k]
    m = r < thresh
    m[:, :, 0] = True                                   # NOOP always valid in an active cell
    if preset == "realistic":
        # 5 of 7 components keep exactly one valid entry
        r2 = torch.rand(shape, generator=gen, device=device)
        single = torch.rand((n, cells, COMPONENTS_PER_CELL), generator=gen, device=device) < 5.0 / 7.0
        for k, (s, w) in enumerate(zip(SEG_START, NVEC_CELL)):
            one = torch.zeros((n, cells, w), dtype=torch.bool, device=device)
            one.scatter_(2, r2[:, :, s:s + w].argmax(dim=2, keepdim=True), True)
            sel = single[:, :, k:k + 1]
            m[:, :, s:s + w] = torch.where(sel, one, m[:, :, s:s + w])
    active = torch.rand((n, cells, 1), generator=gen, device=device) < active_p
    return (m & active).to(torch.uint8).view(n, -1)


def make_obs(n: int, h: int, w: int, *, gen: torch.Generator, device="cpu",
             dtype=torch.float32) -> torch.Tensor:
    """``obs [n,h,w,27]`` in {0,1}: five one-hot groups 5/5/3/8/6, index 0 w.p. 0.9."""
    obs = torch.zeros((n, h, w, sum(OBS_PLANES)), dtype=dtype, device=device)
    base = 0
    for width in OBS_PLANES:
        zero = torch.rand((n, h, w), generator=gen, device=device) < 0.9
        other = torch.randint(1, width, (n, h, w), generator=gen, device=device)
        idx = torch.where(zero, torch.zeros_like(other), other) + base
        obs.scatter_(3, idx.unsqueeze(-1), 1)
        base += width
    return obs


def make_valid_actions(mask: torch.Tensor, *, gen: torch.Generator) -> torch.Tensor:
    """``action [n,7C]`` i64: a uniformly random valid index per component
    (any index when the component is fully masked)."""
    n, width = mask.shape
    cells = width // LOGITS_PER_CELL
    mk = mask.view(n, cells, LOGITS_PER_CELL)
    act = torch.empty((n, cells, COMPONENTS_PER_CELL), dtype=torch.int64, device=mask.device)
    for k, (s, w) in enumerate(zip(SEG_START, NVEC_CELL)):
        r = torch.rand((n, cells, w), generator=gen, device=mask.device)
        r = r + (mk[:, :, s:s + w] != 0).to(r.dtype)
        act[:, :, k] = r.argmax(dim=2)
    return act.view(n, cells * COMPONENTS_PER_CELL)


def make_logits(n: int, cells: int, *, gen: torch.Generator, device="cpu", std: float = 2.0):
    """Head-only runs: ``logits [n,78C]`` ~ N(0, std^2) fp32."""
    return torch.randn((n, cells * LOGITS_PER_CELL), generator=gen, device=device) * std


def make_uniforms(n: int, cells: int, *, gen: torch.Generator, device="cpu"):
    """``uniforms [n,7C]`` fp32 in [0,1) (multiples of 2^-24)."""
    return torch.rand((n, cells * COMPONENTS_PER_CELL), generator=gen, device=device)


def make_vtrace_inputs(rows: int, T: int, *, gen: torch.Generator, device="cpu"):
    """C5 inputs, all ``[rows,T]``: behaviour/learner log-probs with rho around 1
    (about half clipped), values N(0,1), sparse rewards, done ~ Bern(1/250)."""
    lp_new = -torch.rand((rows, T), generator=gen, device=device) * 20.0
    lp_old = lp_new + 0.3 * torch.randn((rows, T), generator=gen, device=device)
    values = torch.randn((rows, T), generator=gen, device=device)
    bootstrap = torch.randn((rows,), generator=gen, device=device)
    hit = torch.rand((rows, T), generator=gen, device=device) < 0.1
    rewards = torch.where(hit, torch.randn((rows, T), generator=gen, device=device),
                          torch.zeros((), device=device))
    done = torch.rand((rows, T), generator=gen, device=device) < (1.0 / 250.0)
    entropy = torch.rand((rows, T), generator=gen, device=device) * 30.0
    return dict(lp_old=lp_old, lp_new=lp_new, values=values, bootstrap=bootstrap,
                rewards=rewards, done=done, entropy=entropy)


def make_rollout_batch(h: int, w: int, T: int, rows: int, *, seed: int = 1234, device="cpu",
                       preset: str = "dense", obs_dtype=torch.float32, full_keys: bool = True):
    """Flat learner batch, ``N = rows*(T+1)`` frames, keys/dtypes of the reference
    buffers (``libs/utils.py:34-46``; every key is reshaped at ``:251-253``).

    ``logprobs`` (behaviour policy) is filled with zeros here; callers that need
    rho around 1 overwrite it with ``learner_logp.detach() + N(0, 0.3^2)`` once a
    model exists (random values give rho == 0 and all-zero losses).
    ``full_keys=False`` drops the keys the learner never reads
    (``policy_logits``, ``last_action``, ``ep_return``, ``ep_step``).
    """
    g = _gen(seed, device)
    cells = h * w
    n = rows * (T + 1)
    batch = dict(
        obs=make_obs(n, h, w, gen=g, device=device, dtype=obs_dtype),
        action_mask=make_masks(n, cells, preset, gen=g, device=device),
    )
    batch["action"] = make_valid_actions(batch["action_mask"], gen=g)
    hit = torch.rand((n,), generator=g, device=device) < 0.1
    batch["reward"] = torch.where(hit, torch.randn((n,), generator=g, device=device),
                                  torch.zeros((), device=device))
    batch["done"] = torch.rand((n,), generator=g, device=device) < (1.0 / 250.0)
    batch["logprobs"] = torch.zeros((n,), device=device)
    batch["baseline"] = torch.randn((n,), generator=g, device=device)
    if full_keys:
        batch["ep_return"] = torch.zeros((n,), device=device)
        batch["ep_step"] = torch.zeros((n,), dtype=torch.int32, device=device)
        batch["policy_logits"] = torch.zeros((n, cells * LOGITS_PER_CELL), device=device)
        batch["last_action"] = torch.zeros((n, cells * COMPONENTS_PER_CELL), dtype=torch.int64,
                                           device=device)
    return batch
