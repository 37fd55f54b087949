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


def test_static_scratch_entry_point_rejects_sizes_beyond_its_limit():
    from microbeast_b200 import _lib
    null = ctypes.c_void_p(0)
    rc = _lib.lib.mb_vtrace_loss_static(null, null, null, null, null, null, null, 65537, 80, 80, 80,
                                        0.99, 1.0, 1.0, 1.0, 0.01, 0.0, null, null, null, null, null, null)
    assert rc == 1
    msg = _lib.last_error()
    assert "65536" in msg and "mb_vtrace_loss" in msg
