"""Canary checks: no kernel writes outside the output ranges it was given (compute-sanitizer is
closed on the GPU pool, so out-of-range writes are looked for with guard bands around every output
and ragged shapes that exercise the clipped / fallback paths)."""
import ctypes

import pytest
import torch

pytestmark = pytest.mark.gpu
PAD = 4096


def _guarded(shape, dtype, dev, fill):
    n = 1
    for s in shape:
        n *= s
    flat = torch.full((n + 2 * PAD,), fill, dtype=dtype, device=dev)
    return flat, flat[PAD:PAD + n].view(shape)


def _intact(flat, n, fill):
    return bool((flat[:PAD] == fill).all()) and bool((flat[PAD + n:] == fill).all())


def _p(t):
    return ctypes.c_void_p(t.data_ptr())


@pytest.mark.parametrize("cells,n", [(64, 7), (100, 5), (9, 6), (72, 3), (256, 2)])
@pytest.mark.parametrize("variant", [-1, 0, 3, 5])
def test_head_outputs_stay_inside_their_buffers(cells, n, variant):
    import microbeast_b200
    from microbeast_b200 import _lib, synth
    lib = _lib.lib
    dev = torch.device("cuda", 0)
    g = torch.Generator().manual_seed(cells + n)
    logits = synth.make_logits(n, cells, gen=g).to(dev)
    mask = synth.make_masks(n, cells, "all_valid", gen=g).to(dev)
    act = synth.make_valid_actions(mask.cpu(), gen=g).to(dev)
    u = synth.make_uniforms(n, cells, gen=g).to(dev)
    gl = torch.randn(n, generator=g).to(dev)
    stream = ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)
    f_lp, lp = _guarded((n,), torch.float32, dev, 7.5)
    f_en, en = _guarded((n,), torch.float32, dev, 7.5)
    f_dz, dz = _guarded((n, 78 * cells), torch.float32, dev, 7.5)
    f_ac, ac = _guarded((n, 7 * cells), torch.int64, dev, -77)
    microbeast_b200.set_tuning("head_variant", variant)
    try:
        assert lib.mb_head_eval_fwd(_p(logits), _p(mask), _p(act), _p(lp), _p(en), n, cells, None, stream) == 0
        assert lib.mb_head_eval_bwd(_p(logits), _p(mask), _p(act), _p(gl), _p(gl), _p(dz), n, cells, None, stream) == 0
        assert lib.mb_head_sample(_p(logits), _p(mask), _p(u), _p(ac), _p(lp), _p(en), n, cells, None, stream) == 0
    finally:
        microbeast_b200.set_tuning("head_variant", -1)
    torch.cuda.synchronize()
    assert _intact(f_lp, n, 7.5) and _intact(f_en, n, 7.5)
    assert _intact(f_dz, n * 78 * cells, 7.5) and _intact(f_ac, n * 7 * cells, -77)
    assert bool((dz != 7.5).all()) and bool((ac >= 0).all()) and bool((lp != 7.5).all())


@pytest.mark.parametrize("rows,T", [(70, 80), (33, 32), (5, 48), (40, 20), (3, 7), (129, 16)])
@pytest.mark.parametrize("variant", [-1, 0, 3, 7])
def test_vtrace_outputs_stay_inside_their_buffers(rows, T, variant):
    import microbeast_b200
    from microbeast_b200 import _lib, ops, synth
    lib = _lib.lib
    dev = torch.device("cuda", 0)
    g = torch.Generator().manual_seed(rows * 7 + T)
    t = {k: v.to(dev) for k, v in synth.make_vtrace_inputs(rows, T, gen=g).items()}
    done = t["done"].view(torch.uint8)
    outs = {k: _guarded((rows, T), torch.float32, dev, 7.5) for k in ("vs", "adv", "dlp", "dv")}
    f_l4, l4 = _guarded((4,), torch.float32, dev, 7.5)
    ws = ops.vtrace_workspace(dev, rows)
    stream = ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)
    microbeast_b200.set_tuning("vtrace_variant", variant)
    try:
        rc = lib.mb_vtrace_loss(_p(t["lp_old"]), _p(t["lp_new"]), _p(t["values"]), _p(t["bootstrap"]),
                                _p(t["rewards"]), _p(done), _p(t["entropy"]), rows, T, T, T,
                                0.99, 1.0, 1.0, 1.0, 0.01, 0.0, _p(outs["vs"][1]), _p(outs["adv"][1]), _p(l4),
                                _p(outs["dlp"][1]), _p(outs["dv"][1]), _p(ws), ws.numel(), stream)
    finally:
        microbeast_b200.set_tuning("vtrace_variant", -1)
    assert rc == 0, _lib.last_error()
    torch.cuda.synchronize()
    for k, (flat, view) in outs.items():
        assert _intact(flat, rows * T, 7.5), k
        assert bool((view != 7.5).all()), k
    assert _intact(f_l4, 4, 7.5) and bool(torch.isfinite(l4).all())
    assert int(ws[:4].view(torch.int32)[0]) == 0          # ticket word restored


def test_empty_and_noncontiguous_inputs():
    from microbeast_b200 import ops, synth
    from oracle import head as ohead
    dev = torch.device("cuda", 0)
    z = torch.empty((0, 78 * 64), device=dev)
    m = torch.empty((0, 78 * 64), dtype=torch.uint8, device=dev)
    a = torch.empty((0, 7 * 64), dtype=torch.int64, device=dev)
    lp, en = ops.head_eval_fwd(z, m, a)
    assert lp.shape == (0,) and en.shape == (0,)
    assert ops.head_eval_bwd(z, m, a, lp, en).shape == (0, 78 * 64)
    act, lp, en = ops.head_sample(z, m, torch.empty((0, 7 * 64), device=dev))
    assert act.shape == (0, 7 * 64)
    e = torch.empty((0, 16), device=dev)
    out = ops.vtrace_loss(e, e, e, torch.empty(0, device=dev), e, e.bool(), e)
    assert out["vs"].shape == (0, 16) and not out["loss4"].any()
    # bool masks and strided views are accepted (made contiguous by the wrapper)
    g = torch.Generator().manual_seed(1)
    logits = synth.make_logits(12, 64, gen=g)
    mask = synth.make_masks(12, 64, "all_valid", gen=g) * (torch.rand(12, 78 * 64, generator=g) < 0.5)
    action = synth.make_valid_actions(mask, gen=g)
    ref = ohead.head_eval_fast(logits[::2], mask[::2], action[::2])
    got = ops.head_eval_fwd(logits.to(dev)[::2], (mask != 0).to(dev)[::2], action.to(dev)[::2])
    assert torch.allclose(got[0].cpu(), ref[0], rtol=1e-5, atol=1e-4)
    assert torch.allclose(got[1].cpu(), ref[1], rtol=1e-5, atol=1e-4)
    with pytest.raises(TypeError):
        ops.head_eval_fwd(logits.to(dev).double(), mask.to(dev), action.to(dev))
    with pytest.raises(ValueError):
        ops.head_eval_fwd(logits.to(dev), mask.to(dev)[:, :-78], action.to(dev))
