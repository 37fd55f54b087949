"""mb_vtrace_loss_static: V-trace + losses for callers that cannot allocate device memory.

Who needs it: a host that binds the C ABI without an allocator of its own, a step loop that has to
avoid cudaMalloc's device-wide synchronisation, or a thread that runs while another stream of the
process is being captured (cudaMalloc is refused then).  The entry point takes no workspace; its
scratch is a fixed array inside the library, shared by every call on the GPU, so the test below
calls it twice in a row (the second call must not see leftovers of the first) and compares each
result with what the ordinary entry point, mb_vtrace_loss with a caller-owned workspace, returns
for the same inputs.  Everything is pre-allocated by the test; the call itself goes straight
through ctypes.

This file sorts last on purpose: a run that stops at the first failure has then already executed
every other test.
"""
import ctypes

import pytest
import torch


def _ptr(t):
    return ctypes.c_void_p(0 if t is None else t.data_ptr())


def _call_static(lib, t, out, T, rows):
    stream = ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)
    return lib.mb_vtrace_loss_static(
        _ptr(t["lp_old"]), _ptr(t["lp_new"]), _ptr(t["values"]), _ptr(t["bootstrap"]), _ptr(t["rewards"]),
        _ptr(t["done_u8"]), _ptr(t["entropy"]), rows, T, T, T, 0.99, 1.0, 1.0, 1.0, 0.01, 0.0,
        _ptr(out["vs"]), _ptr(out["adv"]), _ptr(out["loss4"]), _ptr(out["dlp"]), _ptr(out["dv"]), stream)


@pytest.mark.gpu
def test_static_scratch_entry_point_matches_workspace_entry_point_twice():
    from microbeast_b200 import _lib, ops, synth
    dev = torch.device("cuda", 0)
    for call, (rows, T, seed) in enumerate([(300, 160, 1), (96, 80, 2)]):
        g = torch.Generator().manual_seed(seed)
        t = {k: v.to(dev) for k, v in synth.make_vtrace_inputs(rows, T, gen=g).items()}
        t["done_u8"] = t["done"].view(torch.uint8)
        out = dict(vs=torch.full((rows, T), 7.0, device=dev), adv=torch.full((rows, T), 7.0, device=dev),
                   dlp=torch.full((rows, T), 7.0, device=dev), dv=torch.full((rows, T), 7.0, device=dev),
                   loss4=torch.full((4,), 7.0, device=dev))
        rc = _call_static(_lib.lib, t, out, T, rows)
        assert rc == 0, _lib.last_error()
        ref = ops.vtrace_loss(t["lp_old"], t["lp_new"], t["values"], t["bootstrap"], t["rewards"], t["done"],
                              t["entropy"])
        torch.cuda.synchronize()
        # same kernel, same launch shape: the existing tests hold such pairs to bit equality
        assert torch.equal(out["vs"], ref["vs"]), f"call {call}: vs"
        assert torch.equal(out["adv"], ref["pg_advantages"]), f"call {call}: advantages"
        assert torch.equal(out["dlp"], ref["d_logp"]), f"call {call}: d_logp"
        assert torch.equal(out["dv"], ref["d_values"]), f"call {call}: d_values"
        assert torch.equal(out["loss4"], ref["loss4"]), f"call {call}: losses"
        assert bool(torch.isfinite(out["loss4"]).all())


def _outputs(rows, T, dev, fill=7.0):
    return dict(vs=torch.full((rows, T), fill, device=dev), adv=torch.full((rows, T), fill, device=dev),
                dlp=torch.full((rows, T), fill, device=dev), dv=torch.full((rows, T), fill, device=dev),
                loss4=torch.full((4,), fill, device=dev))


def _inputs(rows, T, seed, dev):
    from microbeast_b200 import synth
    g = torch.Generator().manual_seed(seed)
    host = synth.make_vtrace_inputs(rows, T, gen=g)
    t = {k: v.to(dev) for k, v in host.items()}
    t["done_u8"] = t["done"].view(torch.uint8)
    return host, t


@pytest.mark.gpu
def test_static_scratch_entry_point_agrees_with_the_oracle_on_a_dirty_scratch():
    """The array keeps whatever the last user left in it.  The test hook "vtrace_static_poison" fills it with
    0xA5 bytes on the stream before the call's own clearing, which is as dirty as it can get (a ticket word that
    is not zero would make no CTA the "last" one and the losses would never be written).  Results are held to
    the CPU restatement of the reference (libs/utils.py:277-329) at the 1e-5 norm the other V-trace tests use."""
    from conftest import close_rms
    from microbeast_b200 import _lib, set_tuning
    from oracle import vtrace as ovt
    dev = torch.device("cuda", 0)
    for rows, T, seed in [(200, 96, 11), (33, 20, 12)]:          # a TMA layout and one that takes the generic kernel
        host, t = _inputs(rows, T, seed, dev)
        out = _outputs(rows, T, dev, fill=float("nan"))
        set_tuning("vtrace_static_poison", 1)
        rc = _call_static(_lib.lib, t, out, T, rows)
        assert rc == 0, _lib.last_error()
        torch.cuda.synchronize()
        ref = ovt.vtrace_losses_and_grads(host["lp_old"], host["lp_new"], host["values"], host["bootstrap"],
                                          host["rewards"], host["done"], host["entropy"])
        for mine, key in (("vs", "vs"), ("adv", "pg_advantages"), ("dlp", "d_logp"), ("dv", "d_values")):
            ok, worst = close_rms(out[mine], ref[key], rtol=1e-5)
            assert ok, f"{key} rows={rows} T={T}: {worst:.2f}x the 1e-5 bound"
        want = torch.stack([ref[k] for k in ("pg_loss", "value_loss", "entropy_loss", "total_loss")])
        assert torch.allclose(out["loss4"].cpu(), want, rtol=1e-5, atol=1e-6), (out["loss4"], want)


@pytest.mark.gpu
def test_static_scratch_entry_point_at_its_row_limit():
    """65 536 rows is the most the built-in array holds partial sums for (one triple per 32-row group)."""
    from microbeast_b200 import _lib, ops
    dev = torch.device("cuda", 0)
    rows, T = 65536, 32
    _, t = _inputs(rows, T, 21, dev)
    out = _outputs(rows, T, dev)
    assert _call_static(_lib.lib, t, out, T, rows) == 0, _lib.last_error()
    ref = ops.vtrace_loss(t["lp_old"], t["lp_new"], t["values"], t["bootstrap"], t["rewards"], t["done"], t["entropy"])
    torch.cuda.synchronize()
    assert torch.equal(out["vs"], ref["vs"]) and torch.equal(out["adv"], ref["pg_advantages"])
    assert torch.equal(out["loss4"], ref["loss4"]) and bool(torch.isfinite(out["loss4"]).all())


@pytest.mark.gpu
def test_static_scratch_entry_point_on_a_second_device():
    """Every GPU holds its own copy of the array; the address is looked up per device on first use there."""
    if torch.cuda.device_count() < 2:
        pytest.skip("one GPU on this box")
    from microbeast_b200 import _lib, ops
    dev = torch.device("cuda", 1)
    with torch.cuda.device(dev):
        _, t = _inputs(130, 64, 31, dev)
        out = _outputs(130, 64, dev)
        assert _call_static(_lib.lib, t, out, 64, 130) == 0, _lib.last_error()
        ref = ops.vtrace_loss(t["lp_old"], t["lp_new"], t["values"], t["bootstrap"], t["rewards"], t["done"],
                              t["entropy"])
        torch.cuda.synchronize(dev)
        assert torch.equal(out["vs"], ref["vs"]) and torch.equal(out["loss4"], ref["loss4"])


def test_static_scratch_entry_point_rejects_sizes_beyond_its_limit():
    from microbeast_b200 import _lib
    null = ctypes.c_void_p(0)
    rc = _lib.lib.mb_vtrace_loss_static(null, null, null, null, null, null, null, 65537, 80, 80, 80,
                                        0.99, 1.0, 1.0, 1.0, 0.01, 0.0, null, null, null, null, null, null)
    assert rc == 1
    msg = _lib.last_error()
    assert "65536" in msg and "mb_vtrace_loss" in msg
