"""Where does the data-parallel step's extra time go?  (run under torchrun, N ranks)

Config 2 per rank (16x16, T=80, 256 rows), batch resident.  Three runs of the same compute with CUDA events:
  solo     every rank steps on its shard with NO reducer (no collective at all): the compute-only step time;
  dp       the normal data-parallel step (GradAllReducer: head slice all-reduced from the backward hook, tail after),
           with marks at begin / head gradients ready / backward done / gradients summed / optimizer done;
  wire     the two all-reduces alone on an idle GPU (no compute to overlap with).
Prints one JSON line from rank 0 with the max over ranks of every interval (ms)."""
import json, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import torch.distributed as dist
from microbeast_b200 import parallel, synth
from microbeast_b200.learner import LEARNER_KEYS, PPO_learn
from microbeast_b200.model import Agent

rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
torch.cuda.set_device(local)
dev = torch.device("cuda", local)
dist.init_process_group("nccl", device_id=dev)
torch.backends.cudnn.allow_tf32 = False; torch.backends.cuda.matmul.allow_tf32 = False
h = w = 16; T = 80; rows = 256; cells = h * w
torch.manual_seed(0)
model = Agent((h, w, 27), [6, 4, 4, 4, 4, 7, 49] * cells, cells).to(dev)
with torch.no_grad():
    model.actor.weight.normal_(0, 0.01)
opt = torch.optim.Adam(model.parameters(), lr=2.5e-4, eps=1e-5, fused=True)
batch = synth.make_rollout_batch(h, w, T, rows, seed=1 + rank, device=dev, obs_dtype=torch.uint8, full_keys=False)
batch["logprobs"] = -torch.rand(batch["reward"].shape, device=dev) * 30
b = {k: batch[k] for k in LEARNER_KEYS}
shifts = torch.arange(27, device=dev, dtype=torch.int32)
b["obs"] = (batch["obs"].to(torch.int32) << shifts).sum(-1, dtype=torch.int32)
red = parallel.GradAllReducer()
STEPS = 20


def ev():
    e = torch.cuda.Event(enable_timing=True); e.record(); return e


def timed(fn, steps=STEPS, warm=5):
    for _ in range(warm):
        fn()
    dist.barrier(); torch.cuda.synchronize()
    s = ev()
    for _ in range(steps):
        fn()
    e = ev(); torch.cuda.synchronize()
    return s.elapsed_time(e) / steps


solo = timed(lambda: PPO_learn(model, model, "p", b, opt, rows, T, 1, sync=False))
dp = timed(lambda: PPO_learn(model, model, "p", b, opt, rows, T, 1, sync=False, reducer=red))
# marks inside the data-parallel step
acc = {}
for _ in range(STEPS):
    red.marks = {}
    s = ev()
    PPO_learn(model, model, "p", b, opt, rows, T, 1, sync=False, reducer=red)
    e = ev(); torch.cuda.synchronize()
    m = red.marks
    iv = {"forward_to_backward_start": s.elapsed_time(m["begin"]), "backward_start_to_head_grads": m["begin"].elapsed_time(m["head_grads_ready"]),
          "head_grads_to_backward_done (encoder backward, head all-reduce in flight)": m["head_grads_ready"].elapsed_time(m["backward_done"]),
          "backward_done_to_gradients_summed (exposed communication)": m["backward_done"].elapsed_time(m["gradients_summed"]),
          "gradients_summed_to_step_end (Adam)": m["gradients_summed"].elapsed_time(e), "step": s.elapsed_time(e)}
    for k, v in iv.items():
        acc.setdefault(k, []).append(v)
red.marks = None
marks = {k: sorted(v)[len(v) // 2] for k, v in acc.items()}
# the same intervals without any collective: a reducer of world size 1 never touches the wire
# the collectives alone
g = red._grads
head, total = red._head_numel, red._total_numel


def wire():
    a = dist.all_reduce(g[:head], async_op=True); c = dist.all_reduce(g[head: total + 4], async_op=True); a.wait(); c.wait()


wire_ms = timed(wire)
out = {"solo_step": solo, "dp_step": dp, "both_allreduces_alone": wire_ms, **marks}
t = torch.tensor([out[k] for k in sorted(out)], device=dev, dtype=torch.float64)
dist.all_reduce(t, op=dist.ReduceOp.MAX)
if rank == 0:
    res = {k: round(float(v), 3) for k, v in zip(sorted(out), t.tolist())}
    res.update(world_size=world, unit="ms, max over ranks (medians of 20 steps for the marks)", head_bytes=head * 4, tail_bytes=(total - head + 4) * 4,
               workload="config 2 per rank: 16x16, T=80, 256 rows, batch resident")
    print(json.dumps(res), flush=True)
dist.destroy_process_group()
