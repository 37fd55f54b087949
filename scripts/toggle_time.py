"""Step time at config 2 with individual model switches on / off."""
import os, sys, itertools
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from microbeast_b200 import model as mm, synth, ops
from microbeast_b200.learner import LEARNER_KEYS, PPO_learn
dev = torch.device("cuda", 0)
torch.backends.cudnn.allow_tf32 = False; torch.backends.cuda.matmul.allow_tf32 = False; torch.backends.cudnn.benchmark = True
h = w = 16; T = 80; rows = 256; cells = h * w
torch.manual_seed(0)
model = mm.Agent((h, w, 27), [6, 4, 4, 4, 4, 7, 49] * cells, cells).to(dev)
with torch.no_grad():
    model.actor.weight.normal_(0, 0.01)
opt = torch.optim.Adam(model.parameters(), lr=2.5e-4, eps=1e-5, fused=True)
batch = synth.make_rollout_batch(h, w, T, rows, seed=1, device=dev, obs_dtype=torch.uint8, full_keys=False)
batch["logprobs"] = -torch.rand(batch["reward"].shape, device=dev) * 30
b = {k: batch[k] for k in LEARNER_KEYS}
shifts = torch.arange(27, device=dev, dtype=torch.int32)
b["obs"] = (batch["obs"].to(torch.int32) << shifts).sum(-1, dtype=torch.int32)
names = sys.argv[1:] or ["FUSED_RES_BLOCK", "FAST_MAXPOOL_FORWARD"]
for combo in itertools.product([True, False], repeat=len(names)):
    for n, v in zip(names, combo):
        setattr(mm, n, v)
    for _ in range(4):
        PPO_learn(model, model, "p", b, opt, rows, T, 1, sync=False)
    torch.cuda.synchronize()
    s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    s.record()
    for _ in range(10):
        PPO_learn(model, model, "p", b, opt, rows, T, 1, sync=False)
    e.record(); torch.cuda.synchronize()
    ops.start_kernel_timing()
    for _ in range(3):
        PPO_learn(model, model, "p", b, opt, rows, T, 1, sync=False)
    kt = ops.stop_kernel_timing()
    conv = sum(sum(v) for k, v in kt.items() if k.startswith("conv3x3")) / 3
    pool = sum(sum(v) for k, v in kt.items() if k.startswith("maxpool")) / 3
    bits = {k: round(sum(v) / 3, 3) for k, v in kt.items() if k.startswith("conv1_bits") or k.startswith("unpack")}
    print(bits)
    print(dict(zip(names, combo)), "step %.3f ms" % (s.elapsed_time(e) / 10), "conv kernels %.3f ms" % conv, "pool kernels %.3f" % pool, flush=True)
