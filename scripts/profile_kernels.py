"""Launches each kernel a few times at BASELINE config 2 size (and the largest V-trace sweep
shape) with an L2 flush in between -- the command ncu wraps (see scripts/gpu_profile.sh)."""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from microbeast_b200 import ops, synth

dev = torch.device("cuda", 0)
g = torch.Generator(device=dev).manual_seed(1)
h, w, T, rows = 16, 16, 80, 256
cells, n = h * w, rows * (T + 1)
logits = synth.make_logits(n, cells, gen=g, device=dev)
mask = synth.make_masks(n, cells, "dense", gen=g, device=dev)
act = synth.make_valid_actions(mask, gen=g)
u = synth.make_uniforms(n, cells, gen=g, device=dev)
gl = torch.randn(n, generator=g, device=dev)
out = torch.empty_like(logits)
vt = synth.make_vtrace_inputs(65536, 512, gen=g, device=dev)
flush = torch.empty(512 << 20, dtype=torch.uint8, device=dev)
reps = int(sys.argv[1]) if len(sys.argv) > 1 else 3
for _ in range(reps):
    flush.zero_(); ops.head_eval_fwd(logits, mask, act)
    flush.zero_(); ops.head_eval_bwd(logits, mask, act, gl, gl, out=out)
    flush.zero_(); ops.head_sample(logits, mask, u)
    flush.zero_(); ops.vtrace_loss(vt["lp_old"], vt["lp_new"], vt["values"], vt["bootstrap"], vt["rewards"],
                                   vt["done"], vt["entropy"])
torch.cuda.synchronize()
ops.check_status(dev)
print("profile_kernels ok")
