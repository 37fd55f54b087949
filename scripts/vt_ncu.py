import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch, microbeast_b200
from microbeast_b200 import ops, synth
dev = torch.device("cuda", 0)
g = torch.Generator(device=dev).manual_seed(1)
T, R = 512, 65536
t = synth.make_vtrace_inputs(R, T, gen=g, device=dev)
variants = [int(v) for v in sys.argv[1].split(",")] if len(sys.argv) > 1 else [2]
for variant in variants:
    for dbg in (0, 7, 5, 1):
        microbeast_b200.set_tuning("vtrace_variant", variant)
        microbeast_b200.set_tuning("vtrace_debug", dbg)
        for _ in range(2):
            ops.vtrace_loss(t["lp_old"], t["lp_new"], t["values"], t["bootstrap"], t["rewards"], t["done"], t["entropy"])
torch.cuda.synchronize()
print("ok")
