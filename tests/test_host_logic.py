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


# ---- cell list of the sparse actor-head paths (model.cell_list_from_active / gather_cells / spread_padding) ----
def _active_map(n, cells, p, seed):
    g = torch.Generator().manual_seed(seed)
    return torch.rand(n, cells, generator=g) < p


@pytest.mark.parametrize("n,cells,p", [(300, 16, 0.06), (1, 4, 1.0), (500, 64, 0.02), (129, 8, 0.0), (200, 4, 0.3)])
def test_cell_list_holds_every_active_pair_once(n, cells, p):
    from microbeast_b200.model import cell_list_from_active
    active = _active_map(n, cells, p, seed=n + cells)
    fidx, cidx = cell_list_from_active(active, auto=False)
    lmax = fidx.shape[1]
    assert fidx.shape == (cells, lmax) and cidx.shape == (cells, lmax) and lmax % 64 == 0 and lmax >= 64
    longest = int(active.sum(0).max())
    assert lmax == max(-(-longest // 64) * 64, 64)                    # padded to the longest group, no further
    assert torch.equal(cidx, torch.arange(cells).unsqueeze(1).expand(cells, lmax))
    valid = fidx < n
    assert torch.equal(fidx[~valid], torch.full_like(fidx[~valid], n))     # padding = one past the last frame
    rebuilt = torch.zeros(n + 1, cells, dtype=torch.int64)
    rebuilt.index_put_((fidx.reshape(-1), cidx.reshape(-1)), torch.ones(cells * lmax, dtype=torch.int64), accumulate=True)
    assert torch.equal(rebuilt[:n], active.to(torch.int64))               # each active pair exactly once, nothing else
    for c in range(cells):                                                 # frames ascending inside a cell, padding last
        row = fidx[c]
        assert torch.equal(row, row.sort().values)


def test_cell_list_auto_falls_back_when_dense():
    from microbeast_b200.model import cell_list_from_active
    assert cell_list_from_active(_active_map(400, 8, 0.9, 1), auto=True) is None       # longest group > n/2
    assert cell_list_from_active(_active_map(400, 8, 0.05, 2), auto=True) is not None
    assert cell_list_from_active(_active_map(100, 8, 0.05, 3), auto=True) is None      # 64-entry minimum > n/2


def test_gather_cells_and_list_products_equal_dense_products():
    """The cell-list backward of model._ActorHead in plain torch on the CPU: with d logits zero outside the
    active cells, dW, dbias and dhidden from the list (bmm over gathered rows, index_add over spread rows)
    equal the dense products dlogits^T @ hidden, dlogits.sum(0), dlogits @ W."""
    from microbeast_b200.model import _PAD_ROWS, cell_list_from_active, gather_cells, spread_padding
    n, cells, width = 150, 8, 12
    g = torch.Generator().manual_seed(5)
    active = _active_map(n, cells, 0.2, seed=9)
    mask = (torch.rand(n, cells, 78, generator=g) < 0.5) & active.unsqueeze(2)
    mask[..., 0] |= active                                                  # an active cell has a valid entry
    mask = mask.to(torch.uint8).view(n, cells * 78)
    action = torch.randint(0, 4, (n, cells * 7), generator=g)
    hidden = torch.randn(n, width, generator=g, dtype=torch.float64)
    weight = torch.randn(cells * 78, width, generator=g, dtype=torch.float64)
    dlogits = torch.randn(n, cells, 78, generator=g, dtype=torch.float64) * active.unsqueeze(2)
    fidx, cidx = cell_list_from_active(active, auto=False)
    lmax = fidx.shape[1]
    safe, mask_c, act_c = gather_cells(fidx, cidx, n, mask, action, cells)
    valid = (fidx < n).reshape(-1)
    assert torch.equal(mask_c[~valid], torch.zeros_like(mask_c[~valid]))     # padding = fully masked cells
    assert torch.equal(mask_c[valid], mask.view(n, cells, 78)[fidx.reshape(-1)[valid], cidx.reshape(-1)[valid]])
    assert torch.equal(act_c[valid], action.view(n, cells, 7)[fidx.reshape(-1)[valid], cidx.reshape(-1)[valid]])
    dl = dlogits[safe, cidx] * (fidx < n).unsqueeze(2)                      # what the cell-list kernel returns
    flat = spread_padding(fidx, n)
    assert int(flat.max()) < n + _PAD_ROWS and torch.equal(flat[valid], fidx.reshape(-1)[valid])
    hid = torch.cat([hidden, hidden.new_zeros(_PAD_ROWS, width)]).index_select(0, flat).view(cells, lmax, width)
    g_weight = torch.bmm(dl.transpose(1, 2), hid).reshape(cells * 78, width)
    g_bias = dl.sum(dim=1).reshape(-1)
    g_hid = torch.bmm(dl, weight.view(cells, 78, width))
    g_hidden = hidden.new_zeros((n + _PAD_ROWS, width)).index_add_(0, flat, g_hid.reshape(-1, width))[:n]
    dense = dlogits.view(n, cells * 78)
    torch.testing.assert_close(g_weight, dense.t() @ hidden, rtol=1e-12, atol=1e-12)
    torch.testing.assert_close(g_bias, dense.sum(0), rtol=1e-12, atol=1e-12)
    torch.testing.assert_close(g_hidden, dense @ weight, rtol=1e-12, atol=1e-12)


def test_linear_bias_in_k_equals_addmm():
    """model._linear_bias_in_k (the cuBLAS fallback of the actor Linear: bias as an extra K column)."""
    g = torch.Generator().manual_seed(11)
    x = torch.randn(37, 24, generator=g, dtype=torch.float64)
    w = torch.randn(19, 24, generator=g, dtype=torch.float64)
    b = torch.randn(19, generator=g, dtype=torch.float64)
    torch.testing.assert_close(model._linear_bias_in_k(x, w, b), torch.addmm(b, x, w.t()), rtol=1e-12, atol=1e-12)


def test_graphed_learner_refuses_the_sparse_head():
    """The cell-list forward sizes tensors from the data and synchronises: it cannot be captured."""
    from microbeast_b200.graphed import GraphedLearner
    agent = model.Agent((4, 4, 27), list(NVEC_CELL) * 16, 16)
    agent.sparse_head = True
    with pytest.raises(ValueError):
        GraphedLearner(agent, torch.optim.SGD(agent.parameters(), lr=0.1), 2, 3, 1)
