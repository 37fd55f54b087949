"""V-trace on an L2-resident, one-CTA-per-SM shape (T=512, R=4096): isolates the compute warp's own latency."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from microbeast_b200 import ops, synth
dev = torch.device("cuda", 0)
g = torch.Generator(device=dev).manual_seed(1)
vt = synth.make_vtrace_inputs(4096, 512, gen=g, device=dev)
for _ in range(4):
    ops.vtrace_loss(vt["lp_old"], vt["lp_new"], vt["values"], vt["bootstrap"], vt["rewards"], vt["done"], vt["entropy"])
torch.cuda.synchronize()
print("ok")
