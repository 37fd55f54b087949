"""The N>1 host path on CPU: two gloo ranks shard the rows, normalise by the global count and
sum one flat gradient bucket; the result equals the single-process full-batch step.  The
compute on each rank is the oracle (the CUDA kernels need a GPU); what is under test is
shard_batch + GradAllReducer + the global-normalisation contract."""
import os
import sys

import numpy as np
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _make(h, rows, T, seed):
    from microbeast_b200 import synth
    from oracle import NVEC_CELL
    from oracle.learner import RefAgent
    torch.manual_seed(seed)
    agent = RefAgent((h, h, 27), list(NVEC_CELL) * (h * h), h * h)
    with torch.no_grad():
        agent.actor.weight.normal_(0, 0.05)
    batch = synth.make_rollout_batch(h, h, T, rows, seed=seed, preset="dense")
    with torch.no_grad():
        o, _ = agent.get_action(batch, learning=True)
    batch["logprobs"] = o["logprobs"] + 0.3 * torch.randn(o["logprobs"].shape)
    return agent, batch


def _loss_and_backward(agent, batch, rows, T, global_rows):
    """Oracle forward with means over the GLOBAL rows*T (what inv_n does in mb_vtrace_loss)."""
    from oracle.vtrace import vtrace_losses
    out, _ = agent.get_action(batch, learning=True)
    rs = lambda x: x.reshape((rows, T + 1) + tuple(x.shape[1:]))
    lp, base, ent = rs(out["logprobs"]), rs(out["baseline"].flatten()), rs(out["entropy"])
    res = vtrace_losses(rs(batch["logprobs"])[:, 1:], lp[:, :-1], base[:, :-1], base[:, -1],
                        rs(batch["reward"])[:, 1:], rs(batch["done"])[:, 1:], ent[:, :-1])
    scale = rows / float(global_rows)
    four = torch.stack([res["pg_loss"], res["value_loss"], res["entropy_loss"], res["total_loss"]]) * scale
    agent.zero_grad()
    four[3].backward()
    return four.detach()


def _worker(rank, world, port, ret):
    sys.path.insert(0, ROOT)
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        from microbeast_b200 import parallel
        torch.set_num_threads(1)
        rows, T = 4, 6
        agent, batch = _make(4, rows, T, seed=11)
        shard = parallel.shard_batch(batch, rows, T, rank, world)
        lo, hi = parallel.shard_rows(rows, rank, world)
        four = _loss_and_backward(agent, shard, hi - lo, T, rows)
        red = parallel.GradAllReducer()
        assert red.world_size == world and red.rank == rank
        four = red.allreduce(agent, four)
        flat = torch.cat([p.grad.reshape(-1) for p in agent.parameters()])
        # the overlapped path (gradients are views of the bucket, head slice reduced from a hook) gives the same sums
        red2 = parallel.GradAllReducer()
        agent2, _ = _make(4, rows, T, seed=11)
        red2.begin(agent2)
        out, _ = agent2.get_action(shard, learning=True)
        from oracle.vtrace import vtrace_losses
        r = hi - lo
        rs = lambda x: x.reshape((r, T + 1) + tuple(x.shape[1:]))
        lp, base, ent = rs(out["logprobs"]), rs(out["baseline"].flatten()), rs(out["entropy"])
        res = vtrace_losses(rs(shard["logprobs"])[:, 1:], lp[:, :-1], base[:, :-1], base[:, -1],
                            rs(shard["reward"])[:, 1:], rs(shard["done"])[:, 1:], ent[:, :-1])
        four2 = torch.stack([res["pg_loss"], res["value_loss"], res["entropy_loss"], res["total_loss"]]) * (r / float(rows))
        four2[3].backward()
        early_fired = red2._early is not None
        four2 = red2.finish(four2.detach())
        flat2 = torch.cat([p.grad.reshape(-1) for p in agent2.parameters()])
        views = all(p.grad.data_ptr() >= red2._grads.data_ptr() and
                    p.grad.data_ptr() < red2._grads.data_ptr() + red2._grads.numel() * 4 for p in agent2.parameters())
        if rank == 0:
            ret["four"] = four.numpy()
            ret["grad"] = flat.numpy()
            ret["calls"] = red.calls
            ret["four2"] = four2.numpy()
            ret["grad2"] = flat2.numpy()
            ret["early"] = early_fired
            ret["views"] = views
    finally:
        dist.destroy_process_group()


def test_two_rank_gradient_sum_equals_full_batch():
    world, port = 2, 29500 + (os.getpid() % 500)
    with mp.Manager() as mgr:
        ret = mgr.dict()
        mp.spawn(_worker, args=(world, port, ret), nprocs=world, join=True)
        four, grad, calls = np.array(ret["four"]), np.array(ret["grad"]), ret["calls"]
        four2, grad2, early, views = np.array(ret["four2"]), np.array(ret["grad2"]), ret["early"], ret["views"]
    torch.set_num_threads(1)
    rows, T = 4, 6
    agent, batch = _make(4, rows, T, seed=11)
    want_four = _loss_and_backward(agent, batch, rows, T, rows).numpy()
    want_grad = torch.cat([p.grad.reshape(-1) for p in agent.parameters()]).numpy()
    assert calls == 1
    np.testing.assert_allclose(four, want_four, rtol=2e-6, atol=1e-7)
    scale = np.abs(want_grad).max()
    assert np.abs(grad - want_grad).max() <= 2e-6 * scale
    assert early and views, "the head slice must go out from the hook, and gradients must stay views of the bucket"
    np.testing.assert_allclose(four2, want_four, rtol=2e-6, atol=1e-7)
    assert np.abs(grad2 - want_grad).max() <= 2e-6 * scale


def test_single_process_reducer_is_a_no_op():
    from microbeast_b200 import parallel
    lin = torch.nn.Linear(3, 2)
    lin(torch.ones(1, 3)).sum().backward()
    before = lin.weight.grad.clone()
    red = parallel.GradAllReducer()
    extra = torch.tensor([1.0, 2.0])
    assert red.world_size == 1 and red.allreduce(lin, extra) is extra
    assert torch.equal(lin.weight.grad, before)


def test_bind_to_gpu_numa_is_harmless_without_nvml():
    """parallel.bind_to_gpu_numa returns the bound core list, or None (and changes nothing) where NVML or
    sched_setaffinity is unavailable -- as in this CPU-only container."""
    import os
    from microbeast_b200.parallel import bind_to_gpu_numa
    before = os.sched_getaffinity(0) if hasattr(os, "sched_getaffinity") else None
    cores = bind_to_gpu_numa(0)
    assert cores is None or (isinstance(cores, list) and set(cores) <= set(before))
    if cores is None and before is not None:
        assert os.sched_getaffinity(0) == before
    if before is not None:
        os.sched_setaffinity(0, before)
