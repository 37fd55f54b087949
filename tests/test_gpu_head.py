"""Policy-head kernels against the oracle, through the C ABI (ctypes).  Needs a B200."""
import numpy as np
import pytest
import torch

from conftest import close_rms

pytestmark = pytest.mark.gpu

RTOL = 1e-5      # north_star: logp / entropy within 1e-5 relative in fp32 (norm: SURVEY 7.3-2)


def _mods():
    from microbeast_b200 import ops, synth
    from oracle import head as ohead
    return ops, synth, ohead


def _case(cells, n, preset, seed, std=2.0):
    ops, synth, ohead = _mods()
    g = torch.Generator().manual_seed(seed)
    logits = synth.make_logits(n, cells, gen=g, std=std)
    mask = synth.make_masks(n, cells, preset, gen=g)
    action = synth.make_valid_actions(mask, gen=g)
    return logits, mask, action


def _assert_close(x, ref, what, rtol=RTOL, support=None):
    ok, worst = close_rms(x, ref, rtol, support)
    assert ok, f"{what}: worst error {worst:.2f}x the {rtol:g} bound"


def test_golden_edge_cases_fwd_bwd(golden):
    ops, _, _ = _mods()
    g = golden("head_edge")
    z, m, a = (torch.from_numpy(g[k]).cuda() for k in ("logits", "mask", "action"))
    logp, ent = ops.head_eval_fwd(z, m, a)
    _assert_close(logp, g["logp"], "logp")
    _assert_close(ent, g["ent"], "entropy")
    # quirks: fully masked frame is exactly zero (Q1); masked action gives exactly -1e8 per component (Q2)
    assert float(logp[0]) == 0.0 and float(ent[0]) == 0.0
    assert float(logp[3]) == float(np.float32(-1e8) * 14)
    dz = ops.head_eval_bwd(z, m, a, torch.from_numpy(g["g_logp"]).cuda(), torch.from_numpy(g["g_ent"]).cuda())
    _assert_close(dz, g["dlogits"], "dlogits", support=g["mask"] != 0)
    assert not dz[0].any()                                  # Q1: zero gradient
    assert not dz.cpu()[torch.from_numpy(g["mask"]) == 0].any()   # masked entries get exactly 0
    ops.check_status(z.device)


@pytest.mark.parametrize("cells,n,preset", [
    (64, 37, "dense"), (64, 300, "realistic"), (256, 9, "dense"), (256, 5, "all_valid"),
    (64, 7, "all_masked"), (64, 11, "one_valid"), (100, 6, "dense"), (4, 50, "all_valid"),
    (16, 33, "dense"), (1024, 3, "dense"), (72, 5, "all_valid"),
])
def test_eval_fwd_bwd_vs_oracle(cells, n, preset):
    ops, _, ohead = _mods()
    logits, mask, action = _case(cells, n, preset, seed=cells * 131 + n)
    if preset == "dense":
        # densify so the comparison is not dominated by exact zeros
        g = torch.Generator().manual_seed(5)
        from microbeast_b200 import synth
        mask = mask | synth.make_masks(n, cells, "one_valid", gen=g) * (torch.rand(n, 1, generator=g) < 0.5)
        action = synth.make_valid_actions(mask, gen=g)
    ref_lp, ref_ent = ohead.head_eval_fast(logits, mask, action)
    lp, ent = ops.head_eval_fwd(logits.cuda(), mask.cuda(), action.cuda())
    _assert_close(lp, ref_lp, "logp")
    _assert_close(ent, ref_ent, "entropy")
    g = torch.Generator().manual_seed(9)
    g_lp = torch.randn(n, generator=g)
    g_en = torch.randn(n, generator=g)
    ref_dz = ohead.head_eval_grad(logits, mask, action, g_lp, g_en)
    dz = ops.head_eval_bwd(logits.cuda(), mask.cuda(), action.cuda(), g_lp.cuda(), g_en.cuda())
    _assert_close(dz, ref_dz, "dlogits", support=mask != 0)
    assert not dz.cpu()[mask == 0].any() and not ref_dz[mask == 0].any()
    ops.check_status(torch.device("cuda"))


def test_eval_is_deterministic_and_unaligned_views_work():
    ops, _, ohead = _mods()
    logits, mask, action = _case(64, 40, "all_valid", seed=3)
    zc, mc, ac = logits.cuda(), mask.cuda(), action.cuda()
    a1 = ops.head_eval_fwd(zc, mc, ac)
    a2 = ops.head_eval_fwd(zc, mc, ac)
    assert torch.equal(a1[0], a2[0]) and torch.equal(a1[1], a2[1])
    # a slice starting at frame 1 of a buffer whose mask base is then not 16-byte aligned for C=9
    logits, mask, action = _case(9, 21, "all_valid", seed=4)
    ref = ohead.head_eval_fast(logits[1:], mask[1:], action[1:])
    got = ops.head_eval_fwd(logits.cuda()[1:], mask.cuda()[1:], action.cuda()[1:])
    _assert_close(got[0], ref[0], "logp (unaligned)")
    _assert_close(got[1], ref[1], "entropy (unaligned)")


def test_autograd_function_matches_reference_autograd():
    ops, _, ohead = _mods()
    logits, mask, action = _case(64, 12, "all_valid", seed=21)
    w = torch.randn(12, generator=torch.Generator().manual_seed(1))
    z = logits.clone().requires_grad_(True)
    lp, ent = ohead.head_eval_fast(z, mask, action)
    ((lp * w).sum() - 0.01 * ent.sum()).backward()
    zc = logits.cuda().requires_grad_(True)
    lpc, entc = ops.HeadEval.apply(zc, mask.cuda(), action.cuda())
    ((lpc * w.cuda()).sum() - 0.01 * entc.sum()).backward()
    _assert_close(zc.grad, z.grad, "autograd dlogits", support=mask != 0)


@pytest.mark.parametrize("variant", [0, 1, 2, 3, 4, 5, 6, 7])
def test_tile_shapes_agree(variant):
    """Every tile / pipeline shape of the head kernel gives the same bits (the per-frame reduction
    order depends on the tile size, so log-prob / entropy may differ in the last ulp only)."""
    import microbeast_b200
    ops, synth, ohead = _mods()
    logits, mask, action = _case(100, 23, "all_valid", seed=31)
    g = torch.Generator().manual_seed(1)
    u = synth.make_uniforms(23, 100, gen=g)
    zc, mc, ac, uc = logits.cuda(), mask.cuda(), action.cuda(), u.cuda()
    gl = torch.randn(23, generator=g).cuda()
    base = (ops.head_eval_fwd(zc, mc, ac), ops.head_eval_bwd(zc, mc, ac, gl, gl), ops.head_sample(zc, mc, uc))
    microbeast_b200.set_tuning("head_variant", variant)
    try:
        got = (ops.head_eval_fwd(zc, mc, ac), ops.head_eval_bwd(zc, mc, ac, gl, gl), ops.head_sample(zc, mc, uc))
    finally:
        microbeast_b200.set_tuning("head_variant", -1)
    assert torch.allclose(got[0][0], base[0][0], rtol=2e-6, atol=1e-4)
    assert torch.allclose(got[0][1], base[0][1], rtol=2e-6, atol=1e-4)
    assert torch.equal(got[1], base[1])
    assert torch.equal(got[2][0], base[2][0])
    assert torch.equal(got[2][0].cpu(), ohead.sample_invcdf(logits, mask, u))


def test_bad_action_sets_status_like_validate_args():
    ops, _, _ = _mods()
    logits, mask, action = _case(64, 4, "all_valid", seed=8)
    action[2, 5] = 7           # component 5 of cell 0 has 7 entries: index 7 is out of support
    ops.head_eval_fwd(logits.cuda(), mask.cuda(), action.cuda())
    with pytest.raises(ValueError):
        ops.check_status(torch.device("cuda"))
    ops.check_status(torch.device("cuda"))      # cleared


def test_sampler_golden_bit_exact(golden):
    ops, _, _ = _mods()
    g = golden("sampler")
    z, m, u = (torch.from_numpy(g[k]).cuda() for k in ("logits", "mask", "uniforms"))
    act, logp, ent = ops.head_sample(z, m, u)
    assert torch.equal(act.cpu(), torch.from_numpy(g["action"].astype(np.int64)))


@pytest.mark.parametrize("cells,n,preset,std", [
    (64, 64, "dense", 2.0), (64, 33, "all_valid", 2.0), (64, 40, "all_masked", 2.0),
    (64, 25, "one_valid", 2.0), (256, 6, "realistic", 2.0), (100, 5, "all_valid", 6.0),
    (64, 30, "all_valid", 0.0), (16, 9, "dense", 20.0),
])
def test_sampler_bit_exact_vs_oracle(cells, n, preset, std):
    """north_star: sampled actions and mask handling bit-exact on identical inputs and uniforms."""
    ops, synth, ohead = _mods()
    logits, mask, _ = _case(cells, n, preset, seed=cells + 7 * n, std=std)
    g = torch.Generator().manual_seed(77)
    u = synth.make_uniforms(n, cells, gen=g)
    u[0] = 0.0
    u[-1] = float(np.float32(1.0) - np.float32(2.0 ** -24))
    if std == 0.0:
        # all-equal logits (actor init, model.py:136): CDF edges are j/m, put uniforms ON them
        u[1] = 0.25; u[2] = 0.5; u[3] = 0.75; u[4] = float(np.float32(1 / 3)); u[5] = float(np.float32(1 / 7))
    ref_act = ohead.sample_invcdf(logits, mask, u)
    act, logp, ent = ops.head_sample(logits.cuda(), mask.cuda(), u.cuda())
    assert torch.equal(act.cpu(), ref_act), f"{int((act.cpu() != ref_act).sum())} sampled actions differ"
    # the sampled index is valid wherever the component has a valid entry
    mk = mask.view(n, cells, 78)
    seg = [0, 6, 10, 14, 18, 22, 29, 78]
    a3 = act.cpu().view(n, cells, 7)
    for k in range(7):
        m = mk[:, :, seg[k]:seg[k + 1]]
        picked = torch.gather(m, 2, a3[:, :, k:k + 1])[:, :, 0] != 0
        assert bool((picked | ~m.bool().any(2)).all())
    # log-prob / entropy returned with the sample equal the eval kernel at that action
    lp2, ent2 = ops.head_eval_fwd(logits.cuda(), mask.cuda(), act)
    assert torch.equal(logp, lp2) and torch.equal(ent, ent2)
    ref_lp, ref_ent = ohead.head_eval_fast(logits, mask, ref_act)
    _assert_close(logp, ref_lp, "sample logp")
    _assert_close(ent, ref_ent, "sample entropy")


def test_full_size_properties_c2():
    """BASELINE config 2 size (16x16, T=80, R=256: N=20736 frames): size-independent checks."""
    ops, synth, ohead = _mods()
    cells, n = 256, 20736
    dev = torch.device("cuda")
    g = torch.Generator(device=dev).manual_seed(1234)
    logits = synth.make_logits(n, cells, gen=g, device=dev)
    mask = synth.make_masks(n, cells, "dense", gen=g, device=dev)
    u = synth.make_uniforms(n, cells, gen=g, device=dev)
    act, logp, ent = ops.head_sample(logits, mask, u)
    lp2, ent2 = ops.head_eval_fwd(logits, mask, act)
    assert torch.equal(logp, lp2) and torch.equal(ent, ent2)
    # sampled entries are valid wherever anything is valid
    mk = mask.view(n, cells, 78)
    a3 = act.view(n, cells, 7)
    seg = [0, 6, 10, 14, 18, 22, 29, 78]
    for k in range(7):
        m = mk[:, :, seg[k]:seg[k + 1]]
        picked = torch.gather(m, 2, a3[:, :, k:k + 1])[:, :, 0] != 0
        assert bool((picked | ~(m != 0).any(2)).all())
        assert int(a3[:, :, k].min()) >= 0 and int(a3[:, :, k].max()) < seg[k + 1] - seg[k]
    # a random subset of frames against the oracle, bit-exact actions
    idx = torch.randperm(n, generator=torch.Generator().manual_seed(3))[:48]
    ref_act = ohead.sample_invcdf(logits[idx.cuda()].cpu(), mask[idx.cuda()].cpu(), u[idx.cuda()].cpu())
    assert torch.equal(act[idx.cuda()].cpu(), ref_act)
    ref_lp, ref_ent = ohead.head_eval_fast(logits[idx.cuda()].cpu(), mask[idx.cuda()].cpu(), ref_act)
    _assert_close(logp[idx.cuda()], ref_lp, "logp subset")
    _assert_close(ent[idx.cuda()], ref_ent, "entropy subset")
    # backward: linear in the upstream gradients, zero at masked entries, and with g_ent = 0 the
    # gradient of every component sums to zero over its entries (softmax Jacobian)
    g1 = torch.randn(n, generator=g, device=dev)
    zero = torch.zeros(n, device=dev)
    dz1 = ops.head_eval_bwd(logits, mask, act, g1, zero)
    assert not dz1[mask == 0].any()
    cell_sum = dz1.view(n, cells, 78)[:, :, 29:].sum(2)
    scale = g1.abs().view(n, 1)
    assert float((cell_sum.abs() / (scale + 1e-6)).max()) < 1e-4
    dz2 = ops.head_eval_bwd(logits, mask, act, 2 * g1, zero)
    assert torch.allclose(dz2, 2 * dz1, rtol=1e-6, atol=1e-7)
    ops.check_status(dev)


@pytest.mark.parametrize("n,cells,offset", [(1, 1, 0), (3, 16, 0), (7, 37, 0), (129, 64, 0), (5, 16, 6), (2000, 256, 0)])
def test_active_cells_matches_any(n, cells, offset):
    """mb_active_cells == (mask != 0).any over each cell's 78 entries; ragged tails and a base that
    is not 16-byte aligned take the scalar path."""
    ops, _, _ = _mods()
    g = torch.Generator().manual_seed(n * 31 + cells)
    flat = torch.zeros(offset + n * cells * 78, dtype=torch.uint8)
    body = (torch.rand(n, cells, 78, generator=g) < 0.02) & (torch.rand(n, cells, 1, generator=g) < 0.3)
    body[0, 0, 77] = True                      # last byte of a cell alone
    if n * cells > 1:
        body[-1, -1] = False
        body[-1, -1, 0] = True                 # first byte of the last cell alone
    flat[offset:] = body.reshape(-1).to(torch.uint8) * 255      # any non-zero byte counts
    mask = flat.cuda()[offset:].view(n, cells * 78)
    got = ops.active_cells(mask, cells)
    assert got.dtype == torch.bool and got.shape == (n, cells)
    assert torch.equal(got.cpu(), body.any(dim=2))


def test_multinomial_rule_sampler_matches_oracle_and_torch_distribution():
    """mb_head_sample_exp: the reference's own rule, argmax(probs / Exp(1)) (model.py:177 -> torch.multinomial).
    (1) with the same exponentials the actions equal the oracle's fp32 evaluation (which reproduces the real
    ``.sample()`` bit for bit, tests/test_oracle_golden.py) except where the two largest ratios are a rounding apart;
    (2) over many draws the action frequencies follow the masked soft-max (chi-square), as do torch.multinomial's."""
    import numpy as np
    from microbeast_b200 import ops, synth
    from oracle.head import MaskedCategorical, sample_expdiv
    n, cells = 4096, 16
    g = torch.Generator().manual_seed(5)
    logits = synth.make_logits(n, cells, gen=g)
    mask = synth.make_masks(n, cells, "dense", gen=g) | synth.make_masks(n, cells, "one_valid", gen=g)
    mask[:, :78] = 0                                         # one fully masked cell per frame (uniform, quirk Q1)
    expo = torch.empty(n, 78 * cells).exponential_(generator=g)
    act, logp, ent = ops.head_sample_exp(logits.cuda(), mask.cuda(), expo.cuda(), want_entropy=True)
    want = sample_expdiv(logits, mask, expo)
    same = (act.cpu() == want)
    assert float(same.float().mean()) >= 0.9999, float(same.float().mean())
    lp_ref, en_ref = ohead_eval(logits, mask, act.cpu())
    assert torch.allclose(logp.cpu(), lp_ref, rtol=1e-5, atol=1e-4) and torch.allclose(ent.cpu(), en_ref, rtol=1e-5, atol=1e-4)
    # masked entries of partly valid components are never drawn
    seg_start = (0, 6, 10, 14, 18, 22, 29)
    mk = mask.view(n, cells, 78)
    a = act.cpu().view(n, cells, 7)
    for k, s in enumerate(seg_start):
        w = (6, 4, 4, 4, 4, 7, 49)[k]
        seg = mk[:, :, s:s + w]
        landed = seg.gather(2, a[:, :, k:k + 1]).squeeze(2) != 0
        assert bool((landed | (seg.sum(2) == 0)).all())
    # distribution: one (logits, mask) row repeated, fresh exponentials from torch's CUDA generator
    reps = 20000
    row_l, row_m = logits[:1].repeat(reps, 1).cuda(), mask[:1].repeat(reps, 1).cuda()
    gen = torch.Generator(device="cuda").manual_seed(1)
    drawn, _, _ = ops.head_sample_exp(row_l, row_m, generator=gen)
    drawn = drawn.cpu().view(reps, cells, 7)
    worst = 0.0
    for cell in (0, 1, 5):
        for k, s in enumerate(seg_start):
            w = (6, 4, 4, 4, 4, 7, 49)[k]
            p = MaskedCategorical(logits[0, cell * 78 + s: cell * 78 + s + w].double()[None],
                                  mask[0, cell * 78 + s: cell * 78 + s + w][None]).probs[0].numpy()
            counts = np.bincount(drawn[:, cell, k].numpy(), minlength=w).astype(np.float64)
            assert counts[p < 1e-12].sum() == 0
            keep = p * reps >= 5
            if keep.sum() < 2:
                continue
            chi2 = float((((counts - p * reps) ** 2)[keep] / (p * reps)[keep]).sum())
            dof = int(keep.sum()) - 1
            worst = max(worst, (chi2 - dof) / np.sqrt(2.0 * dof))
    assert worst < 6.0, f"chi-square {worst:.1f} standard deviations above its mean"


def ohead_eval(logits, mask, action):
    from oracle.head import head_eval_fast
    return head_eval_fast(logits, mask, action)
