"""The oracle against the golden vectors produced by the unmodified reference
(tests/golden/make_golden.py).  CPU only."""
import numpy as np
import pytest
import torch

from oracle import NVEC_CELL, refload
from oracle import head as ohead
from oracle import vtrace as ovtrace
from oracle.learner import RefAgent, ref_learn_step

def _tol(g):
    """bit-exact when run on the torch that wrote the fixture, 2e-6 otherwise."""
    return 0.0 if str(g["torch_version"]) == torch.__version__ else 2e-6


def _check(x, ref, tol):
    x = np.asarray(x); ref = np.asarray(ref)
    if tol == 0.0:
        assert np.array_equal(x, ref)
    else:
        np.testing.assert_allclose(x, ref, rtol=tol, atol=tol * max(1.0, float(np.abs(ref).max())))


def test_head_edge_cases(golden):
    g = golden("head_edge")
    tol = _tol(g)
    z, m, a = (torch.from_numpy(g[k]) for k in ("logits", "mask", "action"))
    lp, ent = ohead.head_eval_loop(z, m, a)
    _check(lp, g["logp"], tol); _check(ent, g["ent"], tol)
    lpc, entc = ohead.head_eval_components(z, m, a)
    _check(lpc, g["logp_comp"], tol); _check(entc, g["ent_comp"], tol)
    dz = ohead.head_eval_grad(z, m, a, torch.from_numpy(g["g_logp"]), torch.from_numpy(g["g_ent"]))
    _check(dz, g["dlogits"], tol)
    # quirks the fixture pins: Q1 fully masked frame, Q2 masked action
    assert g["logp"][0] == 0.0 and g["ent"][0] == 0.0 and not g["dlogits"][0].any()
    assert g["logp"][3] == np.float32(-1e8) * 14


def test_sampler_definition_stable(golden):
    g = golden("sampler")
    act = ohead.sample_invcdf(torch.from_numpy(g["logits"]), torch.from_numpy(g["mask"]),
                              torch.from_numpy(g["uniforms"]))
    assert np.array_equal(act.numpy(), g["action"].astype(np.int64))


def test_sampler_python_loop_matches_vectorised(golden):
    """The definition, written out as plain Python loops on a few frames."""
    g = golden("sampler")
    z = g["logits"][36:60].astype(np.float64); m = g["mask"][36:60]; u = g["uniforms"][36:60].astype(np.float64)
    ref = g["action"][36:60]
    cells = int(g["cells"])
    for f in range(z.shape[0]):
        for c in range(cells):
            s = 78 * c
            for k, n in enumerate(NVEC_CELL):
                zz = np.where(m[f, s:s + n] != 0, z[f, s:s + n], -1e8)
                mx = zz.max(); lse = mx + np.log(np.exp(zz - mx).sum()); l = zz - lse
                e = np.exp(l - l.max()); p = e / e.sum()
                acc, a = 0.0, 0
                for j in range(n - 1):
                    acc += p[j]
                    a += u[f, 7 * c + k] >= acc
                if (m[f, s:s + n] != 0).any() and m[f, s + a] == 0:
                    a = int(np.flatnonzero(m[f, s:s + n])[-1])
                assert a == ref[f, 7 * c + k], (f, c, k)
                s += n


def test_vtrace_small(golden):
    g = golden("vtrace_small")
    tol = _tol(g)
    t = {k: torch.from_numpy(g[k]) for k in ("lp_old", "lp_new", "values", "bootstrap", "rewards", "done", "entropy")}
    out = ovtrace.vtrace_losses_and_grads(t["lp_old"], t["lp_new"], t["values"], t["bootstrap"],
                                          t["rewards"], t["done"], t["entropy"])
    for k in ("vs", "pg_advantages", "pg_loss", "value_loss", "entropy_loss", "total_loss",
              "d_logp", "d_values", "d_entropy"):
        _check(out[k], g[k], tol)
    # hand-derived gradients of the three means
    rt = g["vs"].size
    np.testing.assert_allclose(g["d_logp"], g["pg_advantages"] / rt, rtol=1e-6, atol=1e-8)
    np.testing.assert_allclose(g["d_values"], -(g["vs"] - g["values"]) / rt, rtol=1e-5, atol=1e-8)
    np.testing.assert_allclose(g["d_entropy"], np.full_like(g["entropy"], -0.01 / rt), rtol=1e-6)


def _agent_batch(g):
    n = g["obs"].shape[0]
    cells = int(g["h"]) * int(g["w"])
    return dict(
        obs=torch.from_numpy(g["obs"]).float(), action_mask=torch.from_numpy(g["action_mask"]),
        action=torch.from_numpy(g["action"].astype(np.int64)), reward=torch.from_numpy(g["reward"]),
        done=torch.from_numpy(g["done"]), baseline=torch.from_numpy(g["baseline_in"]),
        logprobs=torch.from_numpy(g["logprobs_in"]), ep_return=torch.zeros(n),
        ep_step=torch.zeros(n, dtype=torch.int32), policy_logits=torch.zeros(n, 78 * cells),
        last_action=torch.zeros(n, 7 * cells, dtype=torch.int64))


def golden_agent_weights(state_dict, seed):
    rs = np.random.RandomState(seed)
    out = {}
    for name, t in state_dict.items():
        scale = 0.05 if name.startswith("actor") else 0.15
        out[name] = torch.from_numpy(rs.standard_normal(tuple(t.shape)).astype(np.float32) * scale)
    return out


def test_learner_step_against_real_ppo_learn(golden):
    g = golden("agent_4x4")
    tol = _tol(g)
    h, w = int(g["h"]), int(g["w"])
    nvec = list(NVEC_CELL) * (h * w)
    agent = RefAgent((h, w, 27), nvec, h * w)
    agent.load_state_dict(golden_agent_weights(agent.state_dict(), int(g["weight_seed"])))
    opt = torch.optim.SGD(agent.parameters(), lr=0.0)
    four, res, out = ref_learn_step(agent, _agent_batch(g), opt, int(g["B"]), int(g["T"]), int(g["n_envs"]),
                                    return_intermediates=True)
    _check(np.array(four), g["four"], tol)
    _check(res["vs"].detach(), g["vs"], tol)
    _check(res["pg_advantages"].detach(), g["pg_advantages"], tol)
    grads = dict(agent.named_parameters())
    gtol = max(tol, 2e-6)        # batch reductions depend on the host thread count
    _check(grads["actor.bias"].grad, g["grad_actor_bias"], gtol)
    _check(grads["critic.weight"].grad, g["grad_critic_weight"], gtol)
    _check(grads["network.0.conv.bias"].grad, g["grad_conv0_bias"], gtol)
    with torch.no_grad():
        o, _ = agent.get_action(_agent_batch(g), learning=True)
    _check(o["logprobs"], g["out_logprobs"], tol)
    _check(o["entropy"], g["out_entropy"], tol)
    _check(o["baseline"], g["out_baseline"], tol)


@pytest.mark.skipif(not refload.available(), reason="reference checkout not present")
def test_restatement_equals_imported_reference_live():
    """Where /root/reference exists, run the real class beside the restatement."""
    ref_model, _ = refload.load()
    torch.manual_seed(1)
    for n in NVEC_CELL:
        z = torch.randn(64, n) * 3
        m = (torch.rand(64, n) < 0.5).to(torch.uint8)
        a = torch.randint(0, n, (64,))
        real = ref_model.CategoricalMasked(logits=z, masks=m)
        mine = ohead.MaskedCategorical(z, m)
        assert torch.equal(real.log_prob(a), mine.log_prob(a))
        assert torch.equal(real.entropy(), mine.entropy())
        assert torch.equal(real.probs, mine.probs)


def test_expdiv_sampler_reproduces_the_reference_sample_call():
    """``oracle.head.sample_expdiv`` fed with exponentials drawn in the order the reference draws them equals the real
    ``CategoricalMasked(...).sample()`` of every component (``model.py:176-177``), bit for bit."""
    if not refload.available():
        pytest.skip("reference checkout not present")
    ref_model, _ = refload.load()
    from oracle.head import sample_expdiv
    n, cells = 40, 6
    g = torch.Generator().manual_seed(21)
    logits = torch.randn(n, 78 * cells, generator=g) * 2
    mask = (torch.rand(n, 78 * cells, generator=g) < 0.4).to(torch.uint8)
    mask[:, :78] = 0                                        # a fully masked cell: uniform over all entries (Q1)
    nvec = list(NVEC_CELL) * cells
    torch.manual_seed(77)
    real = torch.stack([ref_model.CategoricalMasked(logits=l, masks=m).sample()
                        for l, m in zip(torch.split(logits, nvec, dim=1), torch.split(mask, nvec, dim=1))]).T
    torch.manual_seed(77)
    expo = torch.cat([torch.empty(n, w).exponential_() for w in nvec], dim=1)      # same draws, same order
    assert torch.equal(sample_expdiv(logits, mask, expo), real)
