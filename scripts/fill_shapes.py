"""Which aten::fill_ / zero_ calls does one learner step make, and how large are they?"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from torch.profiler import profile, ProfilerActivity
from microbeast_b200 import synth
from microbeast_b200.learner import LEARNER_KEYS, PPO_learn
from microbeast_b200.model import Agent
dev = torch.device("cuda", 0)
torch.backends.cudnn.allow_tf32 = False; torch.backends.cuda.matmul.allow_tf32 = False
h = w = 16; T = 80; rows = 256; cells = h * w
torch.manual_seed(0)
model = Agent((h, w, 27), [6, 4, 4, 4, 4, 7, 49] * cells, cells).to(dev)
opt = torch.optim.Adam(model.parameters(), lr=2.5e-4, eps=1e-5, fused=True)
batch = synth.make_rollout_batch(h, w, T, rows, seed=1, device=dev, obs_dtype=torch.uint8, full_keys=False)
batch["logprobs"] = -torch.rand(batch["reward"].shape, device=dev) * 30
b = {k: batch[k] for k in LEARNER_KEYS}
shifts = torch.arange(27, device=dev, dtype=torch.int32)
b["obs"] = (batch["obs"].to(torch.int32) << shifts).sum(-1, dtype=torch.int32)
for _ in range(3):
    PPO_learn(model, model, "p", b, opt, rows, T, 1, sync=False)
torch.cuda.synchronize()
with profile(activities=[ProfilerActivity.CUDA, ProfilerActivity.CPU], record_shapes=True, with_stack=False) as prof:
    PPO_learn(model, model, "p", b, opt, rows, T, 1, sync=False)
    torch.cuda.synchronize()
for e in prof.key_averages(group_by_input_shape=True):
    if e.key in ("aten::fill_", "aten::zero_", "aten::zeros", "aten::copy_", "aten::add_", "aten::mul", "aten::sum", "aten::clone", "aten::contiguous", "aten::cat") and e.device_time_total > 0:
        print(f"{e.key:14s} n={e.count:3d} cuda {e.device_time_total:9.1f} us  shapes {e.input_shapes}")
print(prof.key_averages().table(sort_by="cuda_time_total", row_limit=60, max_name_column_width=60))
