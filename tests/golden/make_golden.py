"""Generate the golden fixtures from the UNMODIFIED reference (authoring container only).

    python tests/golden/make_golden.py

Imports ``/root/reference`` (``oracle/refload.py``: stub gym_microrts/matplotlib),
runs the real ``CategoricalMasked``, ``Agent.get_action(learning=True)`` and
``PPO_learn`` on small seeded inputs and writes their outputs next to this file:

* ``head_edge.npz``   real ``CategoricalMasked`` on hand-built edge cases
* ``agent_4x4.npz``   real ``Agent.get_action`` + real ``PPO_learn`` on a 4x4 map
* ``vtrace_small.npz`` restated V-trace block; the script asserts that the
  restatement reproduces the real ``PPO_learn`` floats bit for bit first
* ``sampler.npz``     the *defined* inverse-CDF sampler (the reference has no
  uniform-driven sampler); the script chi-square-checks it against the real
  ``.sample()`` before writing

torch version used is recorded in every file.  Inputs are stored, so the
fixtures do not depend on RNG stability; only the 4x4 Agent's weights are
re-created from ``numpy.random.RandomState`` (a frozen stream).
"""
import contextlib
import io
import os
import sys

import numpy as np
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)

from oracle import NVEC_CELL, refload                     # noqa: E402
from oracle import head as ohead                          # noqa: E402
from oracle import vtrace as ovtrace                      # noqa: E402
from oracle.learner import RefAgent, ref_learn_step       # noqa: E402

ref_model, ref_utils = refload.load()
SEG = np.concatenate([[0], np.cumsum(NVEC_CELL)]).tolist()


def agent_weights(state_dict, seed):
    """Deterministic weights for a model with the reference's parameter names."""
    rs = np.random.RandomState(seed)
    out = {}
    for name, t in state_dict.items():
        scale = 0.05 if name.startswith("actor") else 0.15
        out[name] = torch.from_numpy(rs.standard_normal(tuple(t.shape)).astype(np.float32) * scale)
    return out


def real_head(logits, mask, action, cells):
    """The learner branch of the reference get_action on given logits, real classes
    (``model.py:168,186-214`` executed with the reference's own ``CategoricalMasked``)."""
    nvec = list(NVEC_CELL) * cells
    split_logits = torch.split(logits, nvec, dim=1)
    split_mask = torch.split(mask, nvec, dim=1)
    act_t = action.transpose(1, 0)
    cats = [ref_model.CategoricalMasked(logits=l, masks=m) for l, m in zip(split_logits, split_mask)]
    logprob = torch.stack([c.log_prob(a) for a, c in zip(act_t, cats)])
    entropy = torch.stack([c.entropy() for c in cats])
    k = len(nvec)
    return logprob.T.reshape(-1, k), entropy.T.reshape(-1, k)


def build_edge_cases():
    """Frames of 2 cells (156 logits) covering the edge cases of SURVEY section 4."""
    rs = np.random.RandomState(7)
    cells = 2
    frames = []

    def blank():
        return (rs.standard_normal(78 * cells).astype(np.float32) * 2.0,
                np.zeros(78 * cells, np.uint8), np.zeros(7 * cells, np.int64))

    # 0: everything masked (Q1), arbitrary actions
    z, m, a = blank(); a[:] = [rs.randint(n) for n in NVEC_CELL] * cells; frames.append((z, m, a))
    # 1: everything valid, random in-range actions
    z, m, a = blank(); m[:] = 1; a[:] = [rs.randint(n) for n in NVEC_CELL] * cells; frames.append((z, m, a))
    # 2: exactly one valid entry per component, action = that entry
    z, m, a = blank()
    for c in range(cells):
        for k, n in enumerate(NVEC_CELL):
            j = rs.randint(n); m[78 * c + SEG[k] + j] = 1; a[7 * c + k] = j
    frames.append((z, m, a))
    # 3: partly valid, action on a MASKED entry (Q2: log_prob == -1e8)
    z, m, a = blank()
    for c in range(cells):
        for k, n in enumerate(NVEC_CELL):
            m[78 * c + SEG[k]] = 1; a[7 * c + k] = n - 1
    frames.append((z, m, a))
    # 4: large positive / negative logits, random masks with valid actions
    z, m, a = blank(); z[::3] = 50.0; z[1::3] = -60.0
    m[:] = rs.rand(78 * cells) < 0.5
    frames.append((z, m, a))
    # 5: very large magnitudes (exp underflow inside a component)
    z, m, a = blank(); z[:] = rs.standard_normal(78 * cells) * 300.0; m[:] = rs.rand(78 * cells) < 0.7
    frames.append((z, m, a))
    # 6: all-equal logits (actor init std=0: model.py:136), mixed masks
    z, m, a = blank(); z[:] = 0.0; m[:] = rs.rand(78 * cells) < 0.4
    frames.append((z, m, a))
    # 7: mask bytes other than 1 count as valid (non-zero -> True, model.py:40)
    z, m, a = blank(); m[:] = (rs.rand(78 * cells) < 0.5) * rs.randint(1, 255, 78 * cells)
    frames.append((z, m, a))
    # 8-11: synthetic "dense" style frames
    for _ in range(4):
        z, m, a = blank(); act = rs.rand(cells) < 0.7
        mm = (rs.rand(cells, 78) < 0.3) & act[:, None]; m[:] = mm.reshape(-1)
        frames.append((z, m, a))
    zs, ms, as_ = map(np.stack, zip(*frames))
    # frames 4..: make actions valid where possible
    for f in range(4, len(frames)):
        for c in range(cells):
            for k, n in enumerate(NVEC_CELL):
                seg = ms[f, 78 * c + SEG[k]: 78 * c + SEG[k] + n]
                valid = np.flatnonzero(seg)
                as_[f, 7 * c + k] = valid[rs.randint(len(valid))] if len(valid) else rs.randint(n)
    return torch.from_numpy(zs), torch.from_numpy(ms), torch.from_numpy(as_), cells


def gen_head_edge():
    logits, mask, action, cells = build_edge_cases()
    z = logits.clone().requires_grad_(True)
    lp_c, ent_c = real_head(z, mask, action, cells)
    logp, ent = lp_c.sum(1), ent_c.sum(1)
    rs = np.random.RandomState(11)
    g_lp = torch.from_numpy(rs.standard_normal(logp.shape).astype(np.float32))
    g_ent = torch.from_numpy(rs.standard_normal(ent.shape).astype(np.float32))
    ((logp * g_lp).sum() + (ent * g_ent).sum()).backward()
    # restatement must agree bit for bit with the real classes
    o_lp, o_ent = ohead.head_eval_loop(logits, mask, action)
    assert torch.equal(o_lp, logp.detach()) and torch.equal(o_ent, ent.detach())
    f_lp, f_ent = ohead.head_eval_components(logits, mask, action)
    assert torch.equal(f_lp, lp_c.detach()) and torch.equal(f_ent, ent_c.detach())
    assert torch.equal(ohead.head_eval_grad(logits, mask, action, g_lp, g_ent), z.grad)
    assert torch.equal(ohead.head_eval_grad(logits, mask, action, g_lp, g_ent, loop=True), z.grad)
    np.savez_compressed(os.path.join(HERE, "head_edge.npz"), torch_version=torch.__version__,
                        cells=cells, logits=logits.numpy(), mask=mask.numpy(), action=action.numpy(),
                        logp_comp=lp_c.detach().numpy(), ent_comp=ent_c.detach().numpy(),
                        logp=logp.detach().numpy(), ent=ent.detach().numpy(),
                        g_logp=g_lp.numpy(), g_ent=g_ent.numpy(), dlogits=z.grad.numpy())
    print("head_edge: frames", logits.shape[0], "logp", logp.detach().numpy()[:4])


def gen_agent_case():
    h = w = 4
    cells = h * w
    nvec = list(NVEC_CELL) * cells
    B, T, n_envs = 2, 5, 3
    rows = B * n_envs
    n = rows * (T + 1)
    real = ref_model.Agent((h, w, 27), nvec, cells)
    weights = agent_weights(real.state_dict(), seed=2024)
    real.load_state_dict(weights)
    mine = RefAgent((h, w, 27), nvec, cells)
    mine.load_state_dict(weights)                         # identical parameter names

    rs = np.random.RandomState(5)
    obs = np.zeros((n, h, w, 27), np.float32)
    base = 0
    for width in (5, 5, 3, 8, 6):
        idx = np.where(rs.rand(n, h, w) < 0.8, 0, rs.randint(1, width, (n, h, w))) + base
        np.put_along_axis(obs, idx[..., None], 1.0, axis=3); base += width
    active = rs.rand(n, cells, 1) < 0.4
    mask = ((rs.rand(n, cells, 78) < 0.35) & active).astype(np.uint8).reshape(n, -1)
    action = np.zeros((n, cells, 7), np.int64)
    for k, nk in enumerate(NVEC_CELL):
        seg = mask.reshape(n, cells, 78)[:, :, SEG[k]:SEG[k] + nk]
        action[:, :, k] = (rs.rand(n, cells, nk) + seg).argmax(2)
    action = action.reshape(n, -1)
    batch = dict(
        obs=torch.from_numpy(obs), action_mask=torch.from_numpy(mask), action=torch.from_numpy(action),
        reward=torch.from_numpy(np.where(rs.rand(n) < 0.3, rs.standard_normal(n), 0).astype(np.float32)),
        done=torch.from_numpy(rs.rand(n) < 0.1),
        baseline=torch.from_numpy(rs.standard_normal(n).astype(np.float32)),
        ep_return=torch.zeros(n), ep_step=torch.zeros(n, dtype=torch.int32),
        policy_logits=torch.zeros(n, 78 * cells), last_action=torch.zeros(n, 7 * cells, dtype=torch.int64),
    )
    with torch.no_grad():
        out, _ = real.get_action(batch, learning=True)
    # behaviour log-probs: learner log-probs + noise so that rho straddles 1
    batch["logprobs"] = out["logprobs"] + torch.from_numpy(rs.standard_normal(n).astype(np.float32) * 0.3)

    def fresh():
        return {k: v.clone() for k, v in batch.items()}

    opt = torch.optim.SGD(real.parameters(), lr=0.0)
    with contextlib.redirect_stdout(io.StringIO()):
        four_real = ref_utils.PPO_learn(real, real, "/tmp/golden_", fresh(), opt, B, T, n_envs)
    opt2 = torch.optim.SGD(mine.parameters(), lr=0.0)
    four_mine, res, lo = ref_learn_step(mine, fresh(), opt2, B, T, n_envs, return_intermediates=True)
    assert four_real == four_mine, (four_real, four_mine)      # restated == real, bit for bit
    for (na, pa), (nb, pb) in zip(real.named_parameters(), mine.named_parameters()):
        assert na == nb and torch.equal(pa.grad, pb.grad), na

    grads = {n_: p.grad.numpy() for n_, p in real.named_parameters()}
    np.savez_compressed(
        os.path.join(HERE, "agent_4x4.npz"), torch_version=torch.__version__,
        h=h, w=w, B=B, T=T, n_envs=n_envs, weight_seed=2024,
        obs=obs.astype(np.uint8), action_mask=mask, action=action.astype(np.int16),
        reward=batch["reward"].numpy(), done=batch["done"].numpy(),
        baseline_in=batch["baseline"].numpy(), logprobs_in=batch["logprobs"].numpy(),
        out_logprobs=out["logprobs"].numpy(), out_entropy=out["entropy"].numpy(),
        out_baseline=out["baseline"].numpy(), out_logits_sample=out["policy_logits"].numpy()[:, :156],
        four=np.array(four_real, np.float64),
        vs=res["vs"].numpy(), pg_advantages=res["pg_advantages"].numpy(),
        grad_actor_bias=grads["actor.bias"], grad_critic_weight=grads["critic.weight"],
        grad_critic_bias=grads["critic.bias"], grad_conv0_bias=grads["network.0.conv.bias"],
        grad_fc_bias=grads["network.5.bias"],
        grad_actor_weight_rowsum=grads["actor.weight"].sum(1),
        grad_norms=np.array([float(np.linalg.norm(g)) for g in grads.values()]),
        grad_names=np.array(list(grads.keys())))
    print("agent_4x4: four", four_real)
    return batch, out


def gen_vtrace_small():
    rs = np.random.RandomState(3)
    R, T = 5, 7
    lp_new = torch.from_numpy((-rs.rand(R, T) * 10).astype(np.float32))
    lp_old = lp_new + torch.from_numpy((rs.standard_normal((R, T)) * 0.5).astype(np.float32))
    values = torch.from_numpy(rs.standard_normal((R, T)).astype(np.float32))
    boot = torch.from_numpy(rs.standard_normal(R).astype(np.float32))
    rew = torch.from_numpy(np.where(rs.rand(R, T) < 0.4, rs.standard_normal((R, T)), 0).astype(np.float32))
    done = torch.from_numpy(rs.rand(R, T) < 0.2)
    ent = torch.from_numpy((rs.rand(R, T) * 5).astype(np.float32))
    out = ovtrace.vtrace_losses_and_grads(lp_old, lp_new, values, boot, rew, done, ent)
    np.savez_compressed(os.path.join(HERE, "vtrace_small.npz"), torch_version=torch.__version__,
                        lp_old=lp_old.numpy(), lp_new=lp_new.numpy(), values=values.numpy(),
                        bootstrap=boot.numpy(), rewards=rew.numpy(), done=done.numpy(),
                        entropy=ent.numpy(),
                        **{k: (v.numpy() if torch.is_tensor(v) else v) for k, v in out.items()})
    print("vtrace_small: total", float(out["total_loss"]))


def gen_sampler():
    logits, mask, _, cells = build_edge_cases()
    n = logits.shape[0]
    rs = np.random.RandomState(17)
    reps = 24
    lg = logits.repeat(reps, 1)
    mk = mask.repeat(reps, 1)
    u = torch.from_numpy((rs.randint(0, 1 << 24, (n * reps, 7 * cells)) / float(1 << 24)).astype(np.float32))
    # adversarial uniforms: 0, largest fp32 below 1, and fp32 roundings of exact CDF edges
    u[0:n] = 0.0
    u[n:2 * n] = float(np.float32(1.0) - np.float32(2.0 ** -24))
    p = ohead.head_probs64(lg, mk)
    for k, (s, wd) in enumerate(zip(SEG[:-1], NVEC_CELL)):
        cdf = np.cumsum(p[:, :, s:s + wd], axis=2)
        for rep, j in ((2, 0), (3, wd // 2), (4, wd - 2)):
            edge = cdf[rep * n:(rep + 1) * n, :, j].astype(np.float32)
            edge = np.minimum(edge, np.float32(1.0) - np.float32(2.0 ** -24))
            u.view(n * reps, cells, 7)[rep * n:(rep + 1) * n, :, k] = torch.from_numpy(edge)
    act = ohead.sample_invcdf(lg, mk, u)
    # every sampled index has non-zero mass unless the component is fully masked
    mv = mk.view(-1, cells, 78).numpy() != 0
    for k, (s, wd) in enumerate(zip(SEG[:-1], NVEC_CELL)):
        a = act.view(-1, cells, 7)[:, :, k].numpy()
        seg = mv[:, :, s:s + wd]
        ok = np.take_along_axis(seg, a[:, :, None], 2)[:, :, 0] | ~seg.any(2)
        assert ok.all()
    # distribution-level check against the reference's own .sample()
    torch.manual_seed(0)
    z = logits[8:9, 29:78]; m = mask[8:9, 29:78]
    if not m.any():
        m = m.clone(); m[0, :5] = 1
    draws = 40000
    cat = ref_model.CategoricalMasked(logits=z.expand(draws, -1), masks=m.expand(draws, -1))
    ref_counts = np.bincount(cat.sample().numpy(), minlength=49)
    zz = torch.zeros(draws, 156); zz[:, 29:78] = z; mm = torch.zeros(draws, 156, dtype=torch.uint8); mm[:, 29:78] = m
    uu = torch.rand(draws, 14)
    mine_counts = np.bincount(ohead.sample_invcdf(zz, mm, uu)[:, 6].numpy(), minlength=49)
    pr = ohead.head_probs64(zz[:1], mm[:1])[0, 0, 29:78]
    sel = pr * draws > 5
    chi_ref = (((ref_counts - pr * draws) ** 2)[sel] / (pr * draws)[sel]).sum()
    chi_mine = (((mine_counts - pr * draws) ** 2)[sel] / (pr * draws)[sel]).sum()
    dof = int(sel.sum()) - 1
    assert ref_counts[~(pr > 0)].sum() == 0 and mine_counts[~(pr > 0)].sum() == 0
    assert chi_ref < 3 * dof + 20 and chi_mine < 3 * dof + 20, (chi_ref, chi_mine, dof)
    np.savez_compressed(os.path.join(HERE, "sampler.npz"), torch_version=torch.__version__, cells=cells,
                        logits=lg.numpy(), mask=mk.numpy(), uniforms=u.numpy(), action=act.numpy().astype(np.int16),
                        chi2=np.array([chi_ref, chi_mine, dof]))
    print("sampler: chi2 ref/mine/dof", chi_ref, chi_mine, dof)


if __name__ == "__main__":
    torch.set_num_threads(1)
    gen_head_edge()
    gen_agent_case()
    gen_vtrace_small()
    gen_sampler()
    for f in sorted(os.listdir(HERE)):
        if f.endswith(".npz"):
            print(f, os.path.getsize(os.path.join(HERE, f)), "bytes")
