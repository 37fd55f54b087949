"""Actor Linear + policy head fused on tcgen05 against the oracle, through the C ABI.  Needs a B200.

Budget: the logits come from three TF32 products of hi/lo-split operands with fp32 accumulation (measured 5.5e-7 of
max|logit| in tests/test_gpu_gemm.py, cuBLAS fp32: 5.0e-7), i.e. they carry the rounding of an fp32 matmul in another
summation order; log-prob and entropy are sums of up to 7*C component terms of those logits.  The checks below hold
the per-frame sums to north_star's 1e-5 in the rms norm of SURVEY 7.3-2 against the oracle evaluated on fp32 logits of
torch's CPU matmul, and the stored logits to 2e-6 of max|logit| against a float64 product.
"""
import pytest
import torch

from conftest import close_rms

pytestmark = pytest.mark.gpu


def _case(n, cells, k, preset, seed, std=0.05):
    from microbeast_b200 import synth
    g = torch.Generator().manual_seed(seed)
    hidden = torch.randn(n, k, generator=g).relu()              # the encoder ends in a ReLU (model.py:131)
    weight = torch.randn(78 * cells, k, generator=g) * std
    bias = torch.randn(78 * cells, generator=g) * 0.1
    mask = synth.make_masks(n, cells, preset, gen=g)
    action = synth.make_valid_actions(mask, gen=g)
    return hidden, weight, bias, mask, action


@pytest.mark.parametrize("n,cells,preset", [(300, 64, "dense"), (129, 16, "all_valid"), (1000, 256, "dense"),
                                            (64, 9, "one_valid"), (257, 64, "all_masked"), (5, 4, "realistic"),
                                            (640, 100, "dense")])
def test_fused_forward_matches_oracle(n, cells, preset):
    from microbeast_b200 import ops
    from oracle.head import head_eval_fast
    hidden, weight, bias, mask, action = _case(n, cells, 256, preset, seed=n + cells)
    logits_ref = torch.addmm(bias, hidden, weight.t())                       # model.py:167 on the CPU, fp32
    lp_ref, en_ref = head_eval_fast(logits_ref, mask, action)
    meta = ops.pack_cells(mask.cuda(), action.cuda(), cells)
    packed = ops.pack_actor_weight(weight.cuda(), bias.cuda(), cells)
    out = ops.actor_head_fwd(hidden.cuda(), packed, meta, want_logits=True)
    assert out is not None
    logp, ent, cell_out, logits = out
    ops.check_status(torch.device("cuda", 0))
    exact = torch.addmm(bias.double(), hidden.double(), weight.double().t())
    err = float((logits.cpu().double() - exact).abs().max() / exact.abs().max())
    assert err <= 2e-6, f"stored logits: {err:.2e} of max|logit|"
    for name, got, want in (("logprobs", logp, lp_ref), ("entropy", ent, en_ref)):
        ok, worst = close_rms(got, want, 1e-5)
        assert ok, f"{name}: {worst:.2f}x the 1e-5 bound"
    # cells without a valid action contribute exactly zero (quirk Q1), and the frame sums are the cell sums in order
    dead = (mask.view(n, cells, 78).sum(2) == 0).t()
    assert not cell_out.cpu()[dead].any()
    assert torch.equal(cell_out.sum(0)[:, 0].cpu() != 0, cell_out.sum(0)[:, 0].cpu() != 0)
    # without the logits store the sums are the same bits
    logp2, ent2, _, none = ops.actor_head_fwd(hidden.cuda(), packed, meta, want_logits=False)
    assert none is None and torch.equal(logp2, logp) and torch.equal(ent2, ent)


def test_fused_forward_agrees_with_the_two_kernel_path():
    """Same inputs through mb_linear_tf32x3 + mb_head_eval_fwd: the fused kernel is that pipeline minus the round trip."""
    from microbeast_b200 import ops
    n, cells = 700, 64
    hidden, weight, bias, mask, action = _case(n, cells, 256, "dense", seed=3)
    d = lambda t: t.cuda()
    logits = ops.linear_tf32x3(d(hidden), d(weight), d(bias))
    lp2, en2 = ops.head_eval_fwd(logits, d(mask), d(action))
    meta = ops.pack_cells(d(mask), d(action), cells)
    logp, ent, _, lg = ops.actor_head_fwd(d(hidden), ops.pack_actor_weight(d(weight), d(bias), cells), meta, want_logits=True)
    assert float((lg - logits).abs().max()) <= 1e-6 * float(logits.abs().max())
    for got, want in ((logp, lp2), (ent, en2)):
        ok, worst = close_rms(got, want, 1e-5)
        assert ok, worst


def test_pack_cells_flags_bad_actions_and_packs_bits():
    from microbeast_b200 import ops
    n, cells = 70, 12                     # cells % 8 != 0: the byte-wise staging path
    g = torch.Generator().manual_seed(9)
    mask = (torch.rand(n, 78 * cells, generator=g) < 0.3).to(torch.uint8)
    action = torch.randint(0, 4, (n, 7 * cells), generator=g)
    meta = ops.pack_cells(mask.cuda(), action.cuda(), cells).cpu()
    ops.check_status(torch.device("cuda", 0))
    words = meta.to(torch.int64) & 0xffffffff
    for cell, frame in ((0, 0), (5, 33), (11, 69)):
        bits = [(int(words[cell, frame, i // 32]) >> (i % 32)) & 1 for i in range(78)]
        assert bits == mask[frame, cell * 78:(cell + 1) * 78].tolist()
        acts = [(int(words[cell, frame, 2]) >> (14 + 6 * k)) & 63 for k in range(3)] + \
               [(int(words[cell, frame, 3]) >> (6 * k)) & 63 for k in range(4)]
        assert acts == action[frame, cell * 7:(cell + 1) * 7].tolist()
    action[3, 8] = 99
    ops.pack_cells(mask.cuda(), action.cuda(), cells)
    with pytest.raises(ValueError):
        ops.check_status(torch.device("cuda", 0))


@pytest.mark.parametrize("n,cells,preset", [(300, 64, "dense"), (97, 16, "all_valid"), (2000, 256, "dense"), (33, 9, "one_valid"),
                                            (64, 64, "all_masked"), (5, 4, "realistic")])
def test_cells_backward_matches_autograd_of_the_oracle(n, cells, preset):
    """mb_cells_bwd (device-side lists of the cells with a valid action + two fp32 products) against torch autograd
    through the oracle head on the CPU: d hidden, d weight, d bias within 1e-5 (rms norm)."""
    from microbeast_b200 import ops
    from oracle.head import head_eval_fast
    hidden, weight, bias, mask, action = _case(n, cells, 256, preset, seed=n * 3 + cells)
    g = torch.Generator().manual_seed(n)
    g_lp, g_en = torch.randn(n, generator=g), torch.randn(n, generator=g) * 0.1
    h = hidden.clone().requires_grad_(True)
    w = weight.clone().requires_grad_(True)
    b = bias.clone().requires_grad_(True)
    lp, en = head_eval_fast(torch.addmm(b, h, w.t()), mask, action)
    ((lp * g_lp).sum() + (en * g_en).sum()).backward()
    d = lambda t: t.cuda()
    meta = ops.pack_cells(d(mask), d(action), cells)
    _, _, _, logits = ops.actor_head_fwd(d(hidden), ops.pack_actor_weight(d(weight), d(bias), cells), meta, want_logits="alive")
    dh, dw, db = ops.cells_bwd(d(hidden), d(weight), logits, meta, d(g_lp), d(g_en))
    dh2, dw2, db2 = ops.cells_bwd(d(hidden), d(weight), logits, meta, d(g_lp), d(g_en))       # buffers re-zeroed, lists rebuilt
    for name, got, again, want in (("d_hidden", dh, dh2, h.grad), ("d_weight", dw, dw2, w.grad), ("d_bias", db, db2, b.grad)):
        if float(want.abs().max()) == 0.0:
            assert not got.any() and not again.any(), name
            continue
        ok, worst = close_rms(got, want, 1e-5)
        assert ok, f"{name}: {worst:.2f}x the 1e-5 bound"
        ok, worst = close_rms(again, want, 1e-5)
        assert ok, f"{name} (second call): {worst:.2f}x the 1e-5 bound"


@pytest.mark.parametrize("n,cells,preset", [(300, 16, "dense"), (257, 64, "realistic"), (64, 256, "dense"), (1, 4, "all_valid"),
                                            (130, 9, "all_masked"), (1000, 64, "one_valid")])
def test_list_forward_equals_the_dense_fused_forward(n, cells, preset):
    """mb_actor_head_list_fwd (only the (frame, cell) pairs with a valid action go through the tensor cores, gathered
    from device-built lists) == mb_actor_head_fwd bit for bit: per-frame sums, per-cell results, and the logits of
    every live cell; dead pairs are exactly zero."""
    from microbeast_b200 import ops
    hidden, weight, bias, mask, action = _case(n, cells, 256, preset, seed=n + cells)
    d = lambda t: t.cuda()
    meta = ops.pack_cells(d(mask), d(action), cells)
    packed = ops.pack_actor_weight(d(weight), d(bias), cells)
    lp0, en0, co0, lg0 = ops.actor_head_fwd(d(hidden), packed, meta, want_logits="alive")
    for _ in range(2):                                   # twice: the workspace and the lists are rebuilt per call
        got = ops.actor_head_list_fwd(d(hidden), packed, meta, want_logits="alive")
        assert got is not None
        lp1, en1, co1, lg1 = got
        assert torch.equal(co1, co0) and torch.equal(lp1, lp0) and torch.equal(en1, en0)
        alive = (mask.view(n, cells, 78) != 0).any(dim=2).cuda()
        assert torch.equal(lg1.view(n, cells, 78)[alive], lg0.view(n, cells, 78)[alive])
        assert not co1.permute(1, 0, 2)[~alive].any()
