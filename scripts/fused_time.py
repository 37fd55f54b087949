"""Fused actor head vs the two-kernel path at config-2 size (profiles/r02_fused_head.txt)."""
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
flush = torch.empty(512 << 20, dtype=torch.uint8, device=dev)
def t(fn, reps=7):
    ts = []
    for i in range(reps):
        flush.zero_(); flush.zero_()
        s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        s.record(); fn(); e.record(); torch.cuda.synchronize()
        ts.append(s.elapsed_time(e))
    ts.sort(); return ts[len(ts) // 2]
meta = ops.pack_cells(mask, action, cells)
packed = ops.pack_actor_weight(weight, bias, cells)
def two():
    lg = ops.linear_tf32x3(hidden, weight, bias)
    return ops.head_eval_fwd(lg, mask, action)
print("pack_cells            %.3f ms" % t(lambda: ops.pack_cells(mask, action, cells)))
print("pack_actor_weight     %.3f ms" % t(lambda: ops.pack_actor_weight(weight, bias, cells)))
print("linear + head_eval    %.3f ms" % t(two))
print("fused, no logits      %.3f ms" % t(lambda: ops.actor_head_fwd(hidden, packed, meta)))
print("fused, logits stored  %.3f ms" % t(lambda: ops.actor_head_fwd(hidden, packed, meta, want_logits=True)))
ops.start_kernel_timing()
for _ in range(3):
    ops.actor_head_fwd(hidden, packed, meta); two()
for kname, v in ops.stop_kernel_timing().items():
    print("  kernel %-20s %.3f ms" % (kname, sum(v) / len(v)))
