"""mb_conv3x3_wgrad (csrc/convgrad.cu) against autograd's convolution weight gradient.  Needs a B200.

The encoder stays on PyTorch (reference model.py:56-137); this kernel replaces cuDNN's fp32 wgrad
kernels inside its backward pass.  Bar: same mathematics in fp32 -- compared with an fp64 evaluation
of the same sum, the error must be of the size of fp32 rounding of that sum and no worse than
cuDNN's own."""
import pytest
import torch
import torch.nn.functional as F

from conftest import close_rms

pytestmark = pytest.mark.gpu

SHAPES = [(27, 16, 16), (27, 16, 8), (16, 16, 8), (16, 16, 4), (16, 32, 16), (16, 32, 8), (16, 32, 4),
          (32, 32, 8), (32, 32, 4), (32, 32, 2)]


def _ref_wgrad(x, dy, dtype):
    x = x.to(dtype).requires_grad_(False)
    w = torch.zeros(dy.shape[1], x.shape[1], 3, 3, dtype=dtype, device=x.device, requires_grad=True)
    y = F.conv2d(x, w, None, padding=1)
    (gw,) = torch.autograd.grad(y, w, dy.to(dtype))
    return gw


@pytest.mark.parametrize("cin,cout,h", SHAPES)
@pytest.mark.parametrize("n", [1, 3, 37, 1000])
def test_wgrad_matches_autograd(cin, cout, h, n):
    from microbeast_b200 import ops
    g = torch.Generator(device="cuda").manual_seed(cin * 1000 + cout * 10 + h + n)
    x = torch.randn(n, cin, h, h, device="cuda", generator=g)
    x = torch.where(torch.rand(x.shape, device="cuda", generator=g) < 0.5, x, torch.zeros_like(x))   # relu-like
    dy = torch.randn(n, cout, h, h, device="cuda", generator=g)
    got, gb = ops.conv3x3_wgrad(x, dy)
    assert got.shape == (cout, cin, 3, 3) and gb.shape == (cout,)
    ref64 = _ref_wgrad(x, dy, torch.float64)
    ok, worst = close_rms(got.double(), ref64, 2e-5)
    assert ok, worst
    ok, worst = close_rms(gb.double(), dy.double().sum(dim=(0, 2, 3)), 2e-5)
    assert ok, ("bias gradient", worst)
    # default: CTAs add into the result with atomics (order varies from run to run); DETERMINISTIC_WGRAD: slots + a
    # fixed-order sum, bit-reproducible
    ops.DETERMINISTIC_WGRAD = True
    try:
        det, gbd = ops.conv3x3_wgrad(x, dy)
        again, gb2 = ops.conv3x3_wgrad(x, dy)
    finally:
        ops.DETERMINISTIC_WGRAD = False
    assert torch.equal(det, again) and torch.equal(gbd, gb2), "the kernel and its fixed-order partial sum are deterministic"
    ok, worst = close_rms(det.double(), ref64, 2e-5)
    assert ok, worst


def test_wgrad_one_hot_taps():
    """Every tap lands where the definition says: x has a single 1, dy a single 1."""
    from microbeast_b200 import ops
    cin, cout, h = 16, 16, 8
    for (yi, xi, yo, xo) in [(0, 0, 0, 0), (0, 0, 1, 1), (3, 4, 2, 5), (7, 7, 6, 6), (7, 0, 7, 1), (2, 2, 5, 5)]:
        x = torch.zeros(5, cin, h, h, device="cuda")
        dy = torch.zeros(5, cout, h, h, device="cuda")
        x[3, 5, yi, xi] = 1.0
        dy[3, 9, yo, xo] = 2.0
        got, _ = ops.conv3x3_wgrad(x, dy)
        want = torch.zeros(cout, cin, 3, 3, device="cuda")
        ky, kx = yi - yo + 1, xi - xo + 1
        if 0 <= ky < 3 and 0 <= kx < 3:
            want[9, 5, ky, kx] = 2.0
        assert torch.equal(got, want), (yi, xi, yo, xo)


def test_wgrad_uncovered_shape_returns_none():
    from microbeast_b200 import ops
    x = torch.randn(4, 8, 8, 8, device="cuda")
    dy = torch.randn(4, 8, 8, 8, device="cuda")
    assert ops.conv3x3_wgrad(x, dy) is None
    assert ops.conv3x3_wgrad(torch.randn(4, 32, 1, 1, device="cuda"), torch.randn(4, 32, 1, 1, device="cuda")) is None


def test_wgrad_at_learner_size_no_worse_than_cudnn():
    """Config-2 sizes: error against fp64 compared with cuDNN's own fp32 error."""
    from microbeast_b200 import ops
    torch.backends.cudnn.allow_tf32 = False
    n = 20736
    for cin, cout, h in [(16, 16, 8), (32, 32, 4), (32, 32, 2), (27, 16, 16)]:
        g = torch.Generator(device="cuda").manual_seed(h)
        x = torch.relu(torch.randn(n, cin, h, h, device="cuda", generator=g))
        dy = torch.randn(n, cout, h, h, device="cuda", generator=g) * 1e-3
        ref64 = _ref_wgrad(x, dy, torch.float64)
        got, _ = ops.conv3x3_wgrad(x, dy)
        cud = _ref_wgrad(x, dy, torch.float32)
        scale = ref64.pow(2).mean().sqrt()
        err_mine = float((got.double() - ref64).abs().max() / scale)
        err_cudnn = float((cud.double() - ref64).abs().max() / scale)
        assert err_mine < 2e-5, (cin, cout, h, err_mine)
        assert err_mine < 4 * err_cudnn + 1e-6, (cin, cout, h, err_mine, err_cudnn)


def test_encoder_backward_with_and_without_fast_wgrad():
    """Agent gradients with FAST_CONV_WGRAD on == off (cuDNN wgrad)."""
    from microbeast_b200 import model as mmodel, synth
    from microbeast_b200.learner import PPO_learn
    h = 8
    torch.manual_seed(3)
    agent = mmodel.Agent((h, h, 27), [6, 4, 4, 4, 4, 7, 49] * (h * h), h * h).cuda()
    with torch.no_grad():
        agent.actor.weight.normal_(0, 0.01)
    T, rows = 6, 9
    batch = synth.make_rollout_batch(h, h, T, rows, seed=4, preset="dense", full_keys=False)
    batch["logprobs"] = -torch.rand(batch["reward"].shape) * 20
    grads = {}
    for flag in (True, False):
        mmodel.FAST_CONV_WGRAD = flag
        try:
            PPO_learn(agent, agent, "w", batch, torch.optim.SGD(agent.parameters(), lr=0.0), rows, T, 1)
        finally:
            mmodel.FAST_CONV_WGRAD = True
        grads[flag] = {k: p.grad.clone() for k, p in agent.named_parameters()}
    for k in grads[True]:
        ok, worst = close_rms(grads[True][k], grads[False][k], 1e-4)
        assert ok, (k, worst)


@pytest.mark.parametrize("n,c,h", [(1, 1, 2), (3, 5, 4), (7, 16, 8), (5, 16, 16), (2, 3, 7), (4, 2, 1), (1000, 32, 4)])
def test_maxpool_backward_matches_autograd(n, c, h):
    """mb_maxpool3s2_bwd == autograd's max_pool2d backward, bit for bit (ties included: both follow the
    forward's argmax indices; a pixel that wins several windows gets their gradients added)."""
    from microbeast_b200 import ops
    g = torch.Generator(device="cuda").manual_seed(n + c + h)
    x = torch.randint(0, 4, (n, c, h, h), device="cuda", generator=g).float()       # many ties
    x.requires_grad_(True)
    y, idx = F.max_pool2d(x, 3, 2, 1, return_indices=True)
    dy = torch.randn(y.shape, device="cuda", generator=g)
    (want,) = torch.autograd.grad(y, x, dy)
    got = ops.maxpool3s2_bwd(dy, idx, h, h)
    # a pixel that is the argmax of several windows sums 2-4 gradients; ATen adds them with atomics in
    # arbitrary order, so allow one rounding of that small sum
    assert torch.allclose(got, want, rtol=0, atol=1e-6), float((got - want).abs().max())


FWD_SHAPES = [(27, 16, 16), (27, 16, 8), (16, 16, 16), (16, 16, 8), (16, 16, 4), (16, 32, 16), (16, 32, 8), (16, 32, 4),
              (32, 16, 16), (32, 16, 8), (32, 16, 4), (32, 32, 8), (32, 32, 4), (32, 32, 2)]


@pytest.mark.parametrize("cin,cout,h", FWD_SHAPES)
@pytest.mark.parametrize("n", [1, 5, 67, 300])
def test_conv_forward_matches_conv2d(cin, cout, h, n):
    """mb_conv3x3_fwd == F.conv2d(padding=1) (fp64 reference), with and without bias; frames past the last
    chunk are not written."""
    from microbeast_b200 import ops
    g = torch.Generator(device="cuda").manual_seed(cin * 100 + cout + h + n)
    x = torch.randn(n, cin, h, h, device="cuda", generator=g)
    w = torch.randn(cout, cin, 3, 3, device="cuda", generator=g) * 0.2
    b = torch.randn(cout, device="cuda", generator=g)
    wk = w.permute(1, 0, 2, 3).reshape(cin, cout, 9).contiguous()
    for bias in (b, None):
        got = ops.conv3x3_fwd(x, wk, bias)
        assert got is not None and got.shape == (n, cout, h, h)
        ref = F.conv2d(x.double(), w.double(), None if bias is None else bias.double(), padding=1)
        ok, worst = close_rms(got.double(), ref, 1e-5)
        assert ok, worst
        assert torch.equal(ops.conv3x3_fwd(x, w, bias, wmode=1), got), "torch weight layout read in the kernel"


def test_conv_forward_as_data_gradient():
    """The data gradient through mb_conv3x3_fwd (flipped kernel, channel axes swapped) == autograd's."""
    from microbeast_b200 import ops
    for cin, cout, h in [(16, 16, 8), (16, 32, 8), (32, 32, 4), (32, 32, 2)]:
        g = torch.Generator(device="cuda").manual_seed(h)
        x = torch.randn(33, cin, h, h, device="cuda", generator=g, dtype=torch.float64, requires_grad=True)
        w = torch.randn(cout, cin, 3, 3, device="cuda", generator=g, dtype=torch.float64)
        dy = torch.randn(33, cout, h, h, device="cuda", generator=g, dtype=torch.float64)
        (want,) = torch.autograd.grad(F.conv2d(x, w, None, padding=1), x, dy)
        got = ops.conv3x3_fwd(dy.float(), w.float().flip(2, 3).reshape(cout, cin, 9).contiguous(), None)
        ok, worst = close_rms(got.double(), want, 1e-5)
        assert ok, (cin, cout, h, worst)
        assert torch.equal(ops.conv3x3_fwd(dy.float(), w.float(), None, wmode=2), got), "tap flip done in the kernel"


def test_conv_forward_uncovered_shape_returns_none():
    from microbeast_b200 import ops
    assert ops.conv3x3_fwd(torch.randn(2, 8, 8, 8, device="cuda"), torch.randn(8, 8, 9, device="cuda"), None) is None


def test_encoder_with_and_without_fast_forward():
    """Agent outputs and gradients with FAST_CONV_FORWARD on == off (cuDNN forward and data gradient)."""
    from microbeast_b200 import model as mmodel, synth
    from microbeast_b200.learner import PPO_learn
    for h in (8, 16):
        torch.manual_seed(5)
        agent = mmodel.Agent((h, h, 27), [6, 4, 4, 4, 4, 7, 49] * (h * h), h * h).cuda()
        with torch.no_grad():
            agent.actor.weight.normal_(0, 0.01)
        T, rows = 5, 7
        batch = synth.make_rollout_batch(h, h, T, rows, seed=6, preset="dense", full_keys=False)
        batch["logprobs"] = -torch.rand(batch["reward"].shape) * 20
        res = {}
        for flag in (True, False):
            mmodel.FAST_CONV_FORWARD = flag
            try:
                four = PPO_learn(agent, agent, "w", batch, torch.optim.SGD(agent.parameters(), lr=0.0), rows, T, 1)
            finally:
                mmodel.FAST_CONV_FORWARD = True
            res[flag] = (four, {k: p.grad.clone() for k, p in agent.named_parameters()})
        import numpy as np
        np.testing.assert_allclose(res[True][0], res[False][0], rtol=2e-5, atol=1e-7)
        # two fp32 implementations of 15 convolution layers and their data gradients: rounding differences of
        # the activations (~1e-6) reach the gradients amplified by the network; 3e-4 of the gradient's rms
        for k in res[True][1]:
            ok, worst = close_rms(res[True][1][k], res[False][1][k], 3e-4)
            assert ok, (h, k, worst)


@pytest.mark.parametrize("c,h,n", [(16, 8, 67), (32, 4, 300), (32, 2, 129), (16, 4, 5)])
def test_fused_residual_block_matches_the_unfused_composition(c, h, n):
    """model._ResBlock (ReLU, skip add and ReLU backward folded into the convolution kernels) against float64 autograd
    of relu-conv-relu-conv + x (reference model.py:67-73): output, input gradient and all four parameter gradients."""
    from microbeast_b200 import model as mmodel
    g = torch.Generator(device="cuda").manual_seed(c + h + n)
    blk = mmodel.Residual_Block(c).cuda()
    x = torch.randn(n, c, h, h, device="cuda", generator=g, requires_grad=True)
    gout = torch.randn(n, c, h, h, device="cuda", generator=g)
    assert mmodel._res_block_fusable(x, c)
    out = blk(x)
    out.backward(gout)
    got = [out.detach(), x.grad] + [p.grad for p in blk.parameters()]
    ref = mmodel.Residual_Block(c).cuda().double()
    ref.load_state_dict(blk.state_dict())
    x64 = x.detach().double().requires_grad_(True)
    mmodel.FUSED_RES_BLOCK = False
    try:
        y = F.conv2d(F.relu(F.conv2d(F.relu(x64), ref.conv0.weight, ref.conv0.bias, padding=1)), ref.conv1.weight, ref.conv1.bias,
                     padding=1) + x64
    finally:
        mmodel.FUSED_RES_BLOCK = True
    y.backward(gout.double())
    want = [y.detach(), x64.grad] + [p.grad for p in ref.parameters()]
    for name, a, b in zip(("out", "dx", "dW0", "db0", "dW1", "db1"), got, want):
        ok, worst = close_rms(a.double(), b, 1e-5)
        assert ok, f"{name}: {worst:.2f}x the 1e-5 bound"


@pytest.mark.parametrize("n,c,h", [(1, 1, 2), (3, 5, 4), (7, 16, 8), (5, 16, 16), (2, 3, 7), (4, 2, 1), (1000, 32, 4)])
def test_maxpool_forward_with_byte_codes_matches_aten(n, c, h):
    """mb_maxpool3s2_fwd / mb_maxpool3s2_bwd_code == ATen's max_pool2d and its backward, bit for bit, ties included
    (first maximum in row-major scan order)."""
    from microbeast_b200 import ops
    g = torch.Generator(device="cuda").manual_seed(n + c + h)
    for ties in (True, False):
        x = torch.randint(0, 4, (n, c, h, h), device="cuda", generator=g).float() if ties else \
            torch.randn(n, c, h, h, device="cuda", generator=g)
        x.requires_grad_(True)
        want, idx = F.max_pool2d(x, 3, 2, 1, return_indices=True)
        y, code = ops.maxpool3s2_fwd(x.detach())
        assert torch.equal(y, want)
        oh = (h + 1) // 2
        oy = torch.arange(oh, device="cuda").view(1, 1, oh, 1)
        ox = torch.arange(oh, device="cuda").view(1, 1, 1, oh)
        pix = (2 * oy - 1 + code.long() // 3) * h + (2 * ox - 1 + code.long() % 3)
        assert torch.equal(pix, idx)
        dy = torch.randn(want.shape, device="cuda", generator=g)
        (gx,) = torch.autograd.grad(want, x, dy)
        assert torch.allclose(ops.maxpool3s2_bwd_code(dy, code, h, h), gx, rtol=0, atol=1e-6)


@pytest.mark.parametrize("h,n", [(16, 37), (8, 130), (8, 3), (16, 1)])
def test_first_convolution_from_bits(h, n):
    """mb_conv1_bits_fwd on bit-packed observations == conv2d on the expanded planes (float64), for one-hot groups and
    for arbitrary bit patterns; its autograd node gives the weight / bias gradients of the dense convolution."""
    from microbeast_b200 import model as mmodel, ops, synth
    g = torch.Generator().manual_seed(h + n)
    conv = torch.nn.Conv2d(27, 16, 3, padding=1).cuda()
    for pattern in ("one_hot", "random", "full", "empty"):
        if pattern == "one_hot":
            planes = synth.make_obs(n, h, h, gen=g).permute(0, 3, 1, 2).contiguous()
        elif pattern == "random":
            planes = (torch.rand(n, 27, h, h, generator=g) < 0.3).float()
        else:
            planes = torch.full((n, 27, h, h), 1.0 if pattern == "full" else 0.0)
        shifts = torch.arange(27, dtype=torch.int64).view(1, 27, 1, 1)
        bits = (planes.long() << shifts).sum(1).to(torch.int32).cuda()
        want = F.conv2d(planes.double().cuda(), conv.weight.double(), conv.bias.double(), padding=1)
        got = ops.conv1_bits_fwd(bits, conv.weight, conv.bias)
        ok, worst = close_rms(got.double(), want, 1e-5)
        assert ok, (pattern, worst)
        # weight / bias gradients: from the bits (set cells, or unset cells against the rectangle sums for planes that
        # are mostly set -- "one_hot" and "full" take that branch) and, with the switch off, from the expanded planes
        gout = torch.randn(want.shape, generator=torch.Generator(device="cuda").manual_seed(1), device="cuda")
        ref = torch.nn.Conv2d(27, 16, 3, padding=1).cuda().double()
        ref.load_state_dict(conv.state_dict())
        ref(planes.double().cuda()).backward(gout.double())
        for from_bits in (True, False):
            conv.zero_grad()
            old = mmodel.FAST_CONV1_BITS_WGRAD
            mmodel.FAST_CONV1_BITS_WGRAD = from_bits
            before = ops.launch_counts.get(f"conv1_bits_wgrad_{h}", 0)
            try:
                mmodel.conv1_from_bits(conv, bits).backward(gout)
            finally:
                mmodel.FAST_CONV1_BITS_WGRAD = old
            assert (ops.launch_counts.get(f"conv1_bits_wgrad_{h}", 0) - before == 1) == from_bits
            for name, a, b in (("dW", conv.weight.grad, ref.weight.grad), ("db", conv.bias.grad, ref.bias.grad)):
                err = (a.double() - b).abs().max() / b.abs().max().clamp_min(1.0)
                assert err < 2e-6, (pattern, from_bits, name, float(err))


@pytest.mark.parametrize("cin,cout,h", [(16, 16, 8), (16, 16, 4), (16, 32, 8), (16, 32, 4), (32, 16, 8), (32, 16, 4),
                                         (32, 32, 8), (32, 32, 4), (32, 32, 2)])
def test_tensor_core_convolution_against_float64(cin, cout, h):
    """csrc/convtc.cu (tcgen05, one M128 x N144 product per pixel tile against all nine taps + shift-and-add epilogue)
    for every shape it covers: forward with a torch weight, the data-gradient form with ReLU on the input, the ReLU
    mask and the residual add folded in, frame counts that leave the last 128-pixel tile ragged -- against float64
    conv2d / conv_transpose2d.  Measured on B200: 2.3e-7 of max |out| (the fp32 FMA kernel: 6.8e-7); bound 2e-6."""
    from microbeast_b200 import ops, set_tuning
    g = torch.Generator(device="cuda").manual_seed(cin * 100 + cout + h)
    rnd = lambda *shape: torch.randn(shape, generator=g, device="cuda")
    frames_per_tile = 128 // (h * h)
    for n in (1, frames_per_tile + 1, 5 * frames_per_tile - 1, 301):
        x, wt, b = rnd(n, cin, h, h), rnd(cout, cin, 3, 3) * 0.2, rnd(cout)
        want = F.conv2d(x.double(), wt.double(), b.double(), padding=1)
        before = dict(ops.launch_counts)
        got = ops.conv3x3_fwd(x, wt, b, wmode=1)
        assert float((got.double() - want).abs().max() / want.abs().max()) < 2e-6, ("forward", n)
        # [cin][cout][9] weight layout (wmode 0)
        got0 = ops.conv3x3_fwd(x, wt.permute(1, 0, 2, 3).reshape(cin, cout, 9).contiguous(), b, wmode=0)
        assert torch.equal(got0, got), ("wmode 0", n)
        wt2 = rnd(cin, cout, 3, 3) * 0.2                  # weight of a convolution cout -> cin whose input gradient this is
        msk, add = rnd(n, cout, h, h), rnd(n, cout, h, h)
        want2 = F.conv_transpose2d(torch.relu(x).double(), wt2.double(), padding=1)
        want2 = torch.where(msk > 0, want2, torch.zeros_like(want2)) + add.double()
        got2 = ops.conv3x3_fwd(x, wt2, None, wmode=2, relu_in=True, mask_src=msk, add_src=add)
        assert float((got2.double() - want2).abs().max() / want2.abs().max()) < 2e-6, ("data gradient + tails", n)
        # the fp32 FMA kernel behind the same entry point (mb_set_tuning("conv_tc", 0)) agrees to fp32 rounding
        set_tuning("conv_tc", 0)
        try:
            old = ops.conv3x3_fwd(x, wt, b, wmode=1)
        finally:
            set_tuning("conv_tc", 1)
        assert float((old - got).abs().max() / want.abs().max()) < 2e-6, ("FMA kernel", n)
