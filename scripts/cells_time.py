"""mb_cells_bwd at config-2 size (for ncu / timing)."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from microbeast_b200 import ops, synth
dev = torch.device("cuda", 0)
g = torch.Generator(device=dev).manual_seed(1)
n, cells, k = 20736, 256, 256
hidden = torch.randn(n, k, device=dev, generator=g).relu()
weight = torch.randn(78 * cells, k, device=dev, generator=g) * 0.01
bias = torch.zeros(78 * cells, device=dev)
mask = synth.make_masks(n, cells, "dense", gen=g, device=dev)
action = synth.make_valid_actions(mask, gen=g)
meta = ops.pack_cells(mask, action, cells)
packed = ops.pack_actor_weight(weight, bias, cells)
_, _, _, logits = ops.actor_head_fwd(hidden, packed, meta, want_logits="alive")
gl = torch.randn(n, device=dev, generator=g); ge = torch.randn(n, device=dev, generator=g)
for _ in range(3):
    ops.cells_bwd(hidden, weight, logits, meta, gl, ge)
torch.cuda.synchronize()
s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
s.record()
for _ in range(10):
    ops.cells_bwd(hidden, weight, logits, meta, gl, ge)
e.record(); torch.cuda.synchronize()
print("cells_bwd %.3f ms per call" % (s.elapsed_time(e) / 10))
