"""Which role bounds mb_actor_head_list_fwd?  Kernel time with one role's work skipped (results wrong by design)."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from microbeast_b200 import ops, synth
from microbeast_b200._lib import lib
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
line = ""
for name, bits in (("all", 0), ("no-epilogue", 1), ("no-gather", 2), ("no-mma", 4), ("nothing", 7)):
    lib.mb_set_tuning(b"head_list_debug", bits)
    for want in ("alive", False):
        for _ in range(3):
            ops.actor_head_list_fwd(hidden, packed, meta, want_logits=want)
        ops.start_kernel_timing(only=("actor_head_list_fwd",))
        for _ in range(10):
            ops.actor_head_list_fwd(hidden, packed, meta, want_logits=want)
        kt = ops.stop_kernel_timing()["actor_head_list_fwd"]
        line += f"  {name}{'' if want else ' (no logits store)'} {sorted(kt)[len(kt) // 2] * 1e3:.0f} us"
lib.mb_set_tuning(b"head_list_debug", 0)
print(line)
