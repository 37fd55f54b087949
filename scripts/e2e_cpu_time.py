"""Host-side cost of the e2e step: wall time of BatchPrefetcher.submit and of PPO_learn's enqueue (sync=False)."""
import os, sys, time
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
nxt = pf.submit(host)
for _ in range(5):
    cur, nxt = nxt, pf.submit(host)
    PPO_learn(model, model, "p", cur, opt, rows, T, 1)
ts, tl, tt = [], [], []
for _ in range(20):
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    cur, nxt = nxt, pf.submit(host)
    t1 = time.perf_counter()
    out = PPO_learn(model, model, "p", cur, opt, rows, T, 1, sync=False)
    t2 = time.perf_counter()
    out.tolist()
    t3 = time.perf_counter()
    ts.append(t1 - t0); tl.append(t2 - t1); tt.append(t3 - t0)
med = lambda v: sorted(v)[len(v) // 2] * 1e3
print(f"submit {med(ts):.3f} ms  PPO_learn enqueue {med(tl):.3f} ms  whole step (to losses on host) {med(tt):.3f} ms")
# step first, submit while it runs, then read
tt2 = []
for _ in range(20):
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    out = PPO_learn(model, model, "p", nxt, opt, rows, T, 1, sync=False)
    nxt = pf.submit(host)
    out.tolist()
    tt2.append(time.perf_counter() - t0)
print(f"step launched first, next batch submitted while it runs: {med(tt2):.3f} ms")
