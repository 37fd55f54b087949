"""Where in the step should the next batch be submitted?  (ms per step, losses read every step; run under torchrun for N > 1)"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import torch.distributed as dist
from microbeast_b200 import buffers as mbuf, synth, parallel
from microbeast_b200.learner import BatchPrefetcher, PPO_learn
from microbeast_b200.model import Agent
world = int(os.environ.get("WORLD_SIZE", "1")); rank = int(os.environ.get("RANK", "0")); local = int(os.environ.get("LOCAL_RANK", "0"))
torch.cuda.set_device(local)
dev = torch.device("cuda", local)
red = None
if world > 1:
    dist.init_process_group("nccl", device_id=dev)
    red = parallel.GradAllReducer()
torch.backends.cudnn.allow_tf32 = False; torch.backends.cuda.matmul.allow_tf32 = False
h = w = 16; T = 80; rows = 256; cells = h * w
torch.manual_seed(0)
model = Agent((h, w, 27), [6, 4, 4, 4, 4, 7, 49] * cells, cells).to(dev)
with torch.no_grad():
    model.actor.weight.normal_(0, 0.01)
opt = torch.optim.Adam(model.parameters(), lr=2.5e-4, eps=1e-5, fused=True)
batch = synth.make_rollout_batch(h, w, T, rows, seed=1 + rank, full_keys=False)
batch["logprobs"] = -torch.rand(batch["reward"].shape) * 30
host = {k: v.pin_memory() for k, v in mbuf.pack_batch(batch).items()}
pf = BatchPrefetcher(dev)
state = {"next": pf.submit(host)}


def timed(fn, n=20, warm=4):
    for _ in range(warm):
        fn()
    torch.cuda.synchronize()
    if world > 1:
        dist.barrier()
    s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    s.record()
    for _ in range(n):
        fn()
    e.record(); torch.cuda.synchronize()
    t = torch.tensor([s.elapsed_time(e) / n], device=dev)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return float(t)


def before():
    cur = state["next"]; state["next"] = pf.submit(host)
    return PPO_learn(model, model, "p", cur, opt, rows, T, 1, reducer=red)


def hook():
    def sub():
        state["next"] = pf.submit(host)
        return state["next"]
    return PPO_learn(model, model, "p", state["next"], opt, rows, T, 1, reducer=red, prefetch=sub)


def after():
    out = PPO_learn(model, model, "p", state["next"], opt, rows, T, 1, reducer=red, sync=False)
    state["next"] = pf.submit(host)
    return out.tolist()


res = {"submit before the step": timed(before), "submit from the hook after the forward pass": timed(hook),
       "submit after the step is enqueued": timed(after)}
if rank == 0:
    print(world, {k: round(v, 3) for k, v in res.items()}, flush=True)
if world > 1:
    dist.destroy_process_group()
