"""V-trace + loss kernel against the oracle, through the C ABI.  Needs a B200."""
import numpy as np
import pytest
import torch

from conftest import close_rms

pytestmark = pytest.mark.gpu

RTOL = 1e-5      # north_star: V-trace targets, advantages and loss within 1e-5 relative (fp32)


def _mods():
    import microbeast_b200
    from microbeast_b200 import ops, synth
    from oracle import vtrace as ovt
    return microbeast_b200, ops, synth, ovt


def _assert_close(x, ref, what, rtol=RTOL):
    ok, worst = close_rms(x, ref, rtol)
    assert ok, f"{what}: worst error {worst:.2f}x the {rtol:g} bound"


def _run(ops, t, **kw):
    c = {k: v.cuda() for k, v in t.items()}
    return ops.vtrace_loss(c["lp_old"], c["lp_new"], c["values"], c["bootstrap"], c["rewards"], c["done"],
                           c["entropy"], **kw)


def _compare(out, ref, scalars_rtol=RTOL):
    _assert_close(out["vs"], ref["vs"], "vs")
    _assert_close(out["pg_advantages"], ref["pg_advantages"], "pg_advantages")
    _assert_close(out["d_logp"], ref["d_logp"], "d_logp")
    _assert_close(out["d_values"], ref["d_values"], "d_values")
    four = out["loss4"].cpu().double().numpy()
    want = np.array([float(ref[k]) for k in ("pg_loss", "value_loss", "entropy_loss", "total_loss")])
    # pg and total are sums of signed terms: compare against the magnitude of the mean |term|
    lp_adv = (torch.as_tensor(ref["pg_advantages"]) * torch.as_tensor(ref["lp_new_in"])).abs().mean().item()
    floor = np.array([lp_adv, 0.0, 0.0, lp_adv])
    err = np.abs(four - want)
    bound = scalars_rtol * np.maximum(np.abs(want), floor)
    assert (err <= bound).all(), f"loss4 {four} vs {want}: err/bound {err / bound}"
    assert abs(out["d_entropy"] - float(torch.as_tensor(ref["d_entropy"]).flatten()[0])) <= 1e-6 * abs(out["d_entropy"])


def test_golden_small(golden):
    _, ops, _, _ = _mods()
    g = golden("vtrace_small")
    t = {k: torch.from_numpy(g[k]) for k in ("lp_old", "lp_new", "values", "bootstrap", "rewards", "done", "entropy")}
    out = _run(ops, t)
    ref = {k: g[k] for k in g.files}
    ref["lp_new_in"] = g["lp_new"]
    _compare(out, ref)


@pytest.mark.parametrize("rows,T", [(256, 80), (32, 64), (33, 80), (7, 16), (100, 20), (64, 33), (5, 1),
                                    (1024, 80), (96, 256), (40, 512), (3, 48), (1000, 44), (300, 12), (70, 100)])
def test_vs_oracle_shapes(rows, T):
    """TMA path (T % 16 == 0) and generic path (other T) against the reference arithmetic."""
    mb, ops, synth, ovt = _mods()
    g = torch.Generator().manual_seed(rows * 1000 + T)
    t = synth.make_vtrace_inputs(rows, T, gen=g)
    t["done"] = torch.rand((rows, T), generator=g) < 0.05
    ref = ovt.vtrace_losses_and_grads(t["lp_old"], t["lp_new"], t["values"], t["bootstrap"], t["rewards"],
                                      t["done"], t["entropy"])
    ref["lp_new_in"] = t["lp_new"]
    out = _run(ops, t)
    _compare(out, ref)
    # the generic kernel gives the same bits as the TMA kernel
    mb.set_tuning("vtrace_force_generic", 1)
    try:
        out2 = _run(ops, t)
    finally:
        mb.set_tuning("vtrace_force_generic", 0)
    for k in ("vs", "pg_advantages", "d_logp", "d_values"):
        assert torch.equal(out[k], out2[k]), k
    assert torch.allclose(out["loss4"], out2["loss4"], rtol=1e-6, atol=0)
    # two calls in a row share the cached workspace: the ticket word must have been restored
    out3 = _run(ops, t)
    assert torch.equal(out["loss4"], out3["loss4"]) and torch.equal(out["vs"], out3["vs"])


@pytest.mark.parametrize("variant", list(range(16)))
def test_pipeline_shapes_agree(variant):
    """Every stages x steps-per-tile shape of the TMA kernel gives the same targets, bit for bit."""
    mb, ops, synth, _ = _mods()
    g = torch.Generator().manual_seed(5)
    t = synth.make_vtrace_inputs(300, 160, gen=g)
    base = _run(ops, t)
    mb.set_tuning("vtrace_variant", variant)
    try:
        out = _run(ops, t)
    finally:
        mb.set_tuning("vtrace_variant", -1)
    for k in ("vs", "pg_advantages", "d_logp", "d_values"):
        assert torch.equal(base[k], out[k]), k
    # loss sums: the persistent grid (CTAs per SM) depends on the shape, so the fp64 fold order does too
    assert torch.allclose(base["loss4"], out["loss4"], rtol=1e-6, atol=0)
    mb.set_tuning("vtrace_l2_promo", 128)
    try:
        out = _run(ops, t)
    finally:
        mb.set_tuning("vtrace_l2_promo", 256)
    for k in ("vs", "pg_advantages", "d_logp", "d_values", "loss4"):
        assert torch.equal(base[k], out[k]), k


@pytest.mark.parametrize("variant", range(0, 16))
@pytest.mark.parametrize("rows,T", [(300, 160), (33, 80), (70, 112), (1000, 48), (17, 512), (5000, 96), (40, 16)])
def test_request_order_of_every_pipeline_shape(variant, rows, T):
    """Tiles are requested in ascending blocks of 256 B per row while they are consumed backwards in time; every
    stages x steps shape (block lengths 1, 2, 4; rows shorter than a block; tile counts that are not a multiple of the
    block; more groups than resident CTAs) must still give the generic kernel's bits and the oracle's values."""
    mb, ops, synth, ovt = _mods()
    g = torch.Generator().manual_seed(rows * 977 + T)
    t = synth.make_vtrace_inputs(rows, T, gen=g)
    t["done"] = torch.rand((rows, T), generator=g) < 0.05
    mb.set_tuning("vtrace_force_generic", 1)
    try:
        base = _run(ops, t)
    finally:
        mb.set_tuning("vtrace_force_generic", 0)
    mb.set_tuning("vtrace_variant", variant)
    try:
        out = _run(ops, t)
        again = _run(ops, t)
    finally:
        mb.set_tuning("vtrace_variant", -1)
    for k in ("vs", "pg_advantages", "d_logp", "d_values"):
        assert torch.equal(base[k], out[k]), k
        assert torch.equal(again[k], out[k]), k
    assert torch.equal(again["loss4"], out["loss4"])
    ref = ovt.vtrace_losses_and_grads(t["lp_old"], t["lp_new"], t["values"], t["bootstrap"], t["rewards"],
                                      t["done"], t["entropy"])
    ref["lp_new_in"] = t["lp_new"]
    _compare(out, ref)


def test_persistent_groups_more_rows_than_resident_ctas():
    """More 32-row groups than resident CTAs: every CTA walks several groups through one pipeline."""
    _, ops, synth, ovt = _mods()
    g = torch.Generator().manual_seed(9)
    rows, T = 32 * 1500 + 5, 48
    t = synth.make_vtrace_inputs(rows, T, gen=g)
    t["done"] = torch.rand((rows, T), generator=g) < 0.05
    ref = ovt.vtrace_losses_and_grads(t["lp_old"], t["lp_new"], t["values"], t["bootstrap"], t["rewards"],
                                      t["done"], t["entropy"])
    ref["lp_new_in"] = t["lp_new"]
    _compare(_run(ops, t), ref)


def test_non_default_constants_and_global_normalisation():
    _, ops, synth, ovt = _mods()
    g = torch.Generator().manual_seed(2)
    t = synth.make_vtrace_inputs(64, 48, gen=g)
    kw = dict(gamma=0.9, rho_bar=1.5, c_bar=0.8, pg_rho_bar=1.2, entropy_coef=0.05)
    ref = ovt.vtrace_losses_and_grads(t["lp_old"], t["lp_new"], t["values"], t["bootstrap"], t["rewards"],
                                      t["done"], t["entropy"], **kw)
    ref["lp_new_in"] = t["lp_new"]
    _compare(_run(ops, t, **kw), ref)
    # two half shards normalised by the global count add up to the full batch (data-parallel contract)
    full = _run(ops, t)
    inv_n = 1.0 / (64 * 48)
    a = _run(ops, {k: v[:32] for k, v in t.items()}, inv_n=inv_n)
    b = _run(ops, {k: v[32:] for k, v in t.items()}, inv_n=inv_n)
    assert torch.allclose(a["loss4"] + b["loss4"], full["loss4"], rtol=2e-6, atol=1e-7)
    assert torch.equal(torch.cat([a["d_logp"], b["d_logp"]]), full["d_logp"])
    assert torch.equal(torch.cat([a["vs"], b["vs"]]), full["vs"])


def test_bandwidth_sweep_size_properties():
    """Largest C5 shape (T=512, R=65536): done everywhere => vs = V + rho*(r - V), adv equal; and
    scaling the rewards by a power of two scales nothing non-linearly (exact in fp32)."""
    _, ops, synth, _ = _mods()
    dev = torch.device("cuda")
    g = torch.Generator(device=dev).manual_seed(11)
    rows, T = 65536, 512
    t = synth.make_vtrace_inputs(rows, T, gen=g, device=dev)
    t["done"] = torch.ones((rows, T), dtype=torch.bool, device=dev)
    out = ops.vtrace_loss(t["lp_old"], t["lp_new"], t["values"], t["bootstrap"], t["rewards"], t["done"],
                          t["entropy"])
    rho = (t["lp_new"] - t["lp_old"]).exp().clamp(max=1.0)
    delta = rho * (t["rewards"] - t["values"])
    assert torch.allclose(out["vs"], delta + t["values"], rtol=2e-6, atol=1e-6)
    assert torch.allclose(out["pg_advantages"], delta, rtol=2e-6, atol=1e-6)
    ent_mean = t["entropy"].double().mean().item()
    assert abs(out["loss4"][2].item() - ent_mean) <= 1e-5 * ent_mean
    v_mean = 0.5 * (delta.double() ** 2).mean().item()
    assert abs(out["loss4"][1].item() - v_mean) <= 1e-5 * v_mean


def test_autograd_function():
    _, ops, synth, ovt = _mods()
    g = torch.Generator().manual_seed(4)
    t = synth.make_vtrace_inputs(48, 32, gen=g)
    ref = ovt.vtrace_losses_and_grads(t["lp_old"], t["lp_new"], t["values"], t["bootstrap"], t["rewards"],
                                      t["done"], t["entropy"])
    c = {k: v.cuda() for k, v in t.items()}
    lp = c["lp_new"].clone().requires_grad_(True)
    v = c["values"].clone().requires_grad_(True)
    e = c["entropy"].clone().requires_grad_(True)
    loss4, vs, adv = ops.VTraceLoss.apply(lp, v, e, c["lp_old"], c["bootstrap"], c["rewards"], c["done"],
                                          0.99, 1.0, 1.0, 1.0, 0.01, None)
    loss4[3].backward()
    _assert_close(lp.grad, ref["d_logp"], "autograd d_logp")
    _assert_close(v.grad, ref["d_values"], "autograd d_values")
    _assert_close(e.grad, ref["d_entropy"], "autograd d_entropy")


@pytest.mark.parametrize("rows,T", [(256, 80), (33, 7), (1000, 44), (5, 1), (4096, 129), (130, 512)])
def test_time_major_kernel(rows, T):
    """mb_vtrace_loss_tmajor on [T, R] arrays: the oracle on the transposed inputs, and the row-major
    kernel bit for bit."""
    _, ops, synth, ovt = _mods()
    g = torch.Generator().manual_seed(rows + 31 * T)
    t = synth.make_vtrace_inputs(rows, T, gen=g)
    t["done"] = torch.rand((rows, T), generator=g) < 0.05
    ref = ovt.vtrace_losses_and_grads(t["lp_old"], t["lp_new"], t["values"], t["bootstrap"], t["rewards"],
                                      t["done"], t["entropy"])
    ref["lp_new_in"] = t["lp_new"]
    c = {k: (v.t().contiguous() if v.dim() == 2 else v).cuda() for k, v in t.items()}
    out = ops.vtrace_loss(c["lp_old"], c["lp_new"], c["values"], c["bootstrap"], c["rewards"], c["done"],
                          c["entropy"], time_major=True)
    assert out["vs"].shape == (T, rows)
    back = {k: (v.t() if torch.is_tensor(v) and v.dim() == 2 else v) for k, v in out.items()}
    _compare(back, ref)
    rm = _run(ops, t)
    for k in ("vs", "pg_advantages", "d_logp", "d_values"):
        assert torch.equal(back[k], rm[k]), k
    assert torch.allclose(out["loss4"], rm["loss4"], rtol=1e-6, atol=0)
