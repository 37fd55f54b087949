"""Rollout store / batcher host logic (no GPU): layout, reference-compatible keys, queue protocol."""
import queue

import pytest
import torch

from microbeast_b200 import buffers as mbuf
from oracle import refload


def _fill(bufs, n_buffers, T, n_envs):
    g = torch.Generator().manual_seed(0)
    for key in bufs:
        for i in range(n_buffers):
            slot = bufs[key][i]
            if not isinstance(slot, torch.Tensor):
                continue
            if slot.dtype == torch.bool:
                slot.copy_(torch.rand(slot.shape, generator=g) < 0.3)
            elif slot.dtype.is_floating_point:
                slot.copy_(torch.randn(slot.shape, generator=g))
            else:
                slot.copy_(torch.randint(0, 4, slot.shape, generator=g).to(slot.dtype))


def test_reference_layout_when_not_lean():
    n_buffers, n_envs, B, T, obs = 5, 3, 2, 4, (8, 8, 27)
    b = mbuf.create_buffers(n_buffers, n_envs, B, T, obs, lean=False, obs_dtype=torch.float32)
    assert sorted(b) == sorted(["obs", "reward", "done", "ep_return", "ep_step", "policy_logits", "baseline",
                                "last_action", "action", "action_mask", "logprobs"])
    assert all(len(v) == n_buffers for v in b.values())
    assert b["obs"][0].shape == (T + 1, n_envs, 8, 8, 27) and b["obs"][0].dtype == torch.float32
    assert b["action_mask"][0].shape == (T + 1, n_envs, 78 * 64) and b["action_mask"][0].dtype == torch.uint8
    assert b["action"][0].shape == (T + 1, n_envs, 7 * 64) and b["action"][0].dtype == torch.int64
    assert b["done"][0].dtype == torch.bool and b["ep_step"][0].dtype == torch.int32
    assert b["policy_logits"][0].shape == (T + 1, n_envs, 78 * 64)
    assert all(t.is_shared() for t in b["obs"])
    # the list entries are views of one slab: a slot write lands in the slab
    b["reward"][3][2, 1] = 7.0
    assert float(b.slabs["reward"][3, 2, 1]) == 7.0


def test_lean_store_discards_dead_keys_but_accepts_their_writes():
    b = mbuf.create_buffers(4, 2, 2, 3, (8, 8, 27))
    assert b["obs"][0].dtype == torch.uint8
    for key in mbuf.DEAD_KEYS:
        assert isinstance(b[key][0], mbuf.DiscardSlot)
        b[key][1][0, ...] = torch.zeros(1)            # microbeast.py:73-100 style slot write: no-op
        assert key not in b.slabs
    # float observations written by the actor are stored exactly as 0/1 bytes
    b["obs"][0][1, ...] = torch.ones(2, 8, 8, 27)
    assert int(b["obs"][0][1].sum()) == 2 * 8 * 8 * 27


def test_get_batch_order_and_queue_protocol():
    n_buffers, n_envs, B, T = 6, 3, 2, 4
    b = mbuf.create_buffers(n_buffers, n_envs, B, T, (8, 8, 27), lean=False, obs_dtype=torch.float32, share=False)
    _fill(b, n_buffers, T, n_envs)
    free_q, full_q = queue.Queue(), queue.Queue()
    for m in (4, 1):
        full_q.put(m)
    batch = mbuf.get_batch(B, {}, 0.99, "cpu", free_q, full_q, b)
    assert sorted([free_q.get_nowait(), free_q.get_nowait()]) == [1, 4] and full_q.empty()
    for key in b:
        want = torch.stack([b[key][4], b[key][1]], dim=1)              # libs/utils.py:197-199
        want = want.reshape((-1,) + tuple(want.shape[3:]))             # libs/utils.py:203-205
        assert torch.equal(batch[key], want), key
    n = (T + 1) * B * n_envs
    assert batch["obs"].shape == (n, 8, 8, 27) and batch["action"].shape == (n, 7 * 64)
    # flat index (t, b, e): entry t*B*n_envs + j*n_envs + e comes from buffer indices[j]
    assert torch.equal(batch["reward"][1 * B * n_envs + 1 * n_envs + 2], b["reward"][1][1, 2])


@pytest.mark.skipif(not refload.available(), reason="reference checkout not present")
def test_get_batch_equals_the_reference_implementation():
    _, ref_utils = refload.load()
    n_buffers, n_envs, B, T = 4, 2, 2, 3
    b = mbuf.create_buffers(n_buffers, n_envs, B, T, (8, 8, 27), lean=False, obs_dtype=torch.float32, share=False)
    _fill(b, n_buffers, T, n_envs)
    import contextlib, io
    with contextlib.redirect_stdout(io.StringIO()):
        ref_b = ref_utils.create_buffers(n_buffers, n_envs, B, T, (8, 8, 27))
    assert sorted(ref_b) == sorted(b)
    for key in b:
        assert ref_b[key][0].shape == b[key][0].shape and ref_b[key][0].dtype == b[key][0].dtype, key
        for i in range(n_buffers):
            ref_b[key][i].copy_(b[key][i])
    fq1, uq1, fq2, uq2 = queue.Queue(), queue.Queue(), queue.Queue(), queue.Queue()
    for m in (2, 0):
        uq1.put(m); uq2.put(m)
    got = mbuf.get_batch(B, {}, 0.99, "cpu", fq1, uq1, b)
    want = ref_utils.get_batch(B, {}, 0.99, "cpu", fq2, uq2, ref_b)
    assert sorted(got) == sorted(want)
    for key in want:
        assert torch.equal(got[key], want[key]), key


def test_compact_store_round_trips_masks_and_actions():
    """compact=True keeps masks as bits and actions as uint8 in the slabs; slot writes pack on the way
    in and the host path of get_batch returns exactly what a plain store returns."""
    from microbeast_b200 import synth
    n_buffers, n_envs, B, T = 3, 2, 2, 4
    plain = mbuf.create_buffers(n_buffers, n_envs, B, T, (8, 8, 27), share=False)
    comp = mbuf.create_buffers(n_buffers, n_envs, B, T, (8, 8, 27), share=False, compact=True)
    assert comp.slabs["action_mask"].shape[-1] == (78 * 64 + 7) // 8 and comp.slabs["action"].dtype == torch.uint8
    src = synth.make_rollout_batch(8, 8, T, n_buffers * n_envs, seed=3, obs_dtype=torch.uint8, full_keys=False)
    for key in ("obs", "action_mask", "action", "reward", "done", "logprobs", "baseline"):
        x = src[key].view((n_buffers, T + 1, n_envs) + tuple(src[key].shape[1:]))
        for i in range(n_buffers):
            for t in range(T + 1):
                plain[key][i][t, ...] = x[i, t]               # per-step slot writes, microbeast.py:95-100
                comp[key][i][t, ...] = x[i, t]
    q = [queue.Queue() for _ in range(4)]
    for m in (2, 0):
        q[1].put(m); q[3].put(m)
    a = mbuf.get_batch(B, {}, 0.99, "cpu", q[0], q[1], plain)
    b = mbuf.get_batch(B, {}, 0.99, "cpu", q[2], q[3], comp)
    assert sorted(a) == sorted(b)
    for key in a:
        assert torch.equal(a[key], b[key]) and a[key].dtype == b[key].dtype, key
    packed = mbuf.pack_batch(src)
    assert packed["action_mask_bits"].shape == (src["action"].shape[0], 624) and packed["action"].dtype == torch.uint8
    assert torch.equal(mbuf.PackedMaskSlot(packed["action_mask_bits"], 78 * 64).unpacked(), src["action_mask"])
