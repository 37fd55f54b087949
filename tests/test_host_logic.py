"""Host-side logic that needs no GPU: sharding, synthetic data, Agent wiring."""
import numpy as np
import pytest
import torch

from microbeast_b200 import model, parallel, synth
from oracle import NVEC_CELL
from oracle.learner import RefAgent


def test_shard_rows_partitions_exactly():
    for total, world in [(256, 1), (256, 2), (1024, 8), (10, 4), (3, 8)]:
        spans = [parallel.shard_rows(total, r, world) for r in range(world)]
        assert spans[0][0] == 0 and spans[-1][1] == total
        for (a, b), (c, d) in zip(spans, spans[1:]):
            assert b == c and b >= a and d >= c
    with pytest.raises(ValueError):
        parallel.shard_rows(8, 2, 2)


def test_shard_batch_keeps_whole_rows():
    rows, T = 6, 3
    flat = torch.arange(rows * (T + 1))
    b = {"x": flat, "y": flat.view(-1, 1).repeat(1, 2)}
    parts = [parallel.shard_batch(b, rows, T, r, 3) for r in range(3)]
    assert torch.equal(torch.cat([p["x"] for p in parts]), flat)
    assert parts[1]["x"].tolist() == list(range(8, 16)) and parts[1]["y"].shape == (8, 2)
    # time-major flat order (t, row): rank 1 of 3 owns rows 2..3 at every t
    tm = parallel.shard_batch(b, rows, T, 1, 3, time_major=True)["x"].view(T + 1, 2)
    assert tm[0].tolist() == [2, 3] and tm[1].tolist() == [8, 9]


def test_synth_shapes_and_mask_presets():
    g = torch.Generator().manual_seed(0)
    b = synth.make_rollout_batch(8, 8, 4, 3, seed=7)
    n = 3 * 5
    assert b["obs"].shape == (n, 8, 8, 27) and b["obs"].dtype == torch.float32
    assert torch.equal(b["obs"].sum(-1), torch.full((n, 8, 8), 5.0))        # five one-hot groups
    assert b["action_mask"].shape == (n, 78 * 64) and b["action_mask"].dtype == torch.uint8
    assert b["action"].shape == (n, 7 * 64) and b["action"].dtype == torch.int64
    assert b["done"].dtype == torch.bool and b["ep_step"].dtype == torch.int32
    # actions are valid wherever the component has a valid entry
    mk = b["action_mask"].view(n, 64, 78)
    a3 = b["action"].view(n, 64, 7)
    seg = np.concatenate([[0], np.cumsum(NVEC_CELL)])
    for k in range(7):
        m = mk[:, :, seg[k]:seg[k + 1]]
        picked = torch.gather(m, 2, a3[:, :, k:k + 1])[:, :, 0] != 0
        assert bool((picked | ~(m != 0).any(2)).all())
        assert int(a3[:, :, k].max()) < NVEC_CELL[k]
    one = synth.make_masks(5, 16, "one_valid", gen=g).view(5, 16, 78)
    for k in range(7):
        assert torch.equal(one[:, :, seg[k]:seg[k + 1]].sum(2), torch.ones(5, 16, dtype=torch.uint8))
    dense = synth.make_masks(400, 64, "dense", gen=g).view(400, 64, 78)
    frac = (dense.sum(2) > 0).float().mean().item()
    assert 0.04 < frac < 0.08
    assert synth.make_masks(2, 4, "all_masked", gen=g).sum() == 0
    with pytest.raises(ValueError):
        synth.make_masks(2, 4, "nope", gen=g)


def test_agent_matches_reference_architecture_and_loads_its_weights():
    nvec = list(NVEC_CELL) * 64
    mine = model.Agent((8, 8, 27), nvec, 64)
    ref = RefAgent((8, 8, 27), nvec, 64)
    assert [k for k, _ in mine.state_dict().items()] == [k for k, _ in ref.state_dict().items()]
    mine.load_state_dict(ref.state_dict())
    assert not mine.actor.weight.any()                        # std=0 init, model.py:136
    b = synth.make_rollout_batch(8, 8, 2, 2, seed=1)
    with torch.no_grad():
        assert torch.equal(mine.forward(b, (), True), ref.forward(b, True))
        assert torch.equal(mine.get_value(b, (), True), ref.get_value(b, True))
    with pytest.raises(ValueError):
        model.Agent((8, 8, 27), [6, 4, 4, 4, 4, 7, 48] * 64, 64)
    # 10x10 / 16x16 / 32x32 flatten sizes of the survey (128 / 128 / 512)
    for side, feat in ((10, 128), (16, 128), (32, 512)):
        a = model.Agent((side, side, 27), list(NVEC_CELL) * side * side, side * side)
        assert a.network[5].in_features == feat
