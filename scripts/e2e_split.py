"""What separates the e2e step from the resident one?  Same step, four ways (ms per step over 20 steps, CUDA events)."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from microbeast_b200 import buffers as mbuf, synth
from microbeast_b200.learner import BatchPrefetcher, PPO_learn
from microbeast_b200.model import Agent
dev = torch.device("cuda", 0)
torch.backends.cudnn.allow_tf32 = False; torch.backends.cuda.matmul.allow_tf32 = False
h = w = 16; T = 80; rows = 256; cells = h * w
torch.manual_seed(0)
model = Agent((h, w, 27), [6, 4, 4, 4, 4, 7, 49] * cells, cells).to(dev)
with torch.no_grad():
    model.actor.weight.normal_(0, 0.01)
opt = torch.optim.Adam(model.parameters(), lr=2.5e-4, eps=1e-5, fused=True)
batch = synth.make_rollout_batch(h, w, T, rows, seed=1, full_keys=False)
batch["logprobs"] = -torch.rand(batch["reward"].shape) * 30
host = {k: v.pin_memory() for k, v in mbuf.pack_batch(batch).items()}
pf = BatchPrefetcher(dev)
staged = pf.submit(host)
torch.cuda.synchronize()


def timed(fn, n=20, warm=4):
    for _ in range(warm):
        fn()
    torch.cuda.synchronize()
    s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    s.record()
    for _ in range(n):
        fn()
    e.record(); torch.cuda.synchronize()
    return s.elapsed_time(e) / n


a = timed(lambda: PPO_learn(model, model, "p", staged, opt, rows, T, 1, sync=False))
b = timed(lambda: PPO_learn(model, model, "p", staged, opt, rows, T, 1))
state = {"next": pf.submit(host)}


def e2e():
    out = PPO_learn(model, model, "p", state["next"], opt, rows, T, 1, sync=False)
    state["next"] = pf.submit(host)
    return out.tolist()


def e2e_nosync():
    out = PPO_learn(model, model, "p", state["next"], opt, rows, T, 1, sync=False)
    state["next"] = pf.submit(host)
    return out


c = timed(e2e)
d = timed(e2e_nosync)
print(f"staged batch re-used, no loss read {a:.3f} | staged batch re-used, losses read every step {b:.3f} | "
      f"fresh host batch every step, losses read (e2e) {c:.3f} | fresh host batch every step, no loss read {d:.3f}")
