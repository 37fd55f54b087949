"""Round-2 kernels at config-2 size (and the largest V-trace sweep shape), each a few times with an L2 flush in
between -- the command ncu wraps (scripts/gpu_profile_r2.sh)."""
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
logits_dense = synth.make_logits(n, cells, gen=g, device=dev)
u = synth.make_uniforms(n, cells, gen=g, device=dev)
vt = synth.make_vtrace_inputs(65536, 512, gen=g, device=dev)
gl = torch.randn(n, device=dev, generator=g); ge = torch.randn(n, device=dev, generator=g)
flush = torch.empty(512 << 20, dtype=torch.uint8, device=dev)
# encoder kernels of round 2: tensor-core convolutions (the two busiest shapes) and the first convolution from bits
x32, w32, b32 = torch.randn(n, 32, 4, 4, device=dev, generator=g), torch.randn(32, 32, 3, 3, device=dev, generator=g) * 0.2, torch.randn(32, device=dev, generator=g)
x16, w16, b16 = torch.randn(n, 16, 8, 8, device=dev, generator=g), torch.randn(16, 16, 3, 3, device=dev, generator=g) * 0.2, torch.randn(16, device=dev, generator=g)
obs_bits = (synth.make_obs(n, 16, 16, gen=torch.Generator().manual_seed(3)).to(torch.int32) << torch.arange(27, dtype=torch.int32)).sum(-1, dtype=torch.int32).to(dev)
w1, b1 = torch.randn(16, 27, 3, 3, device=dev, generator=g) * 0.2, torch.randn(16, device=dev, generator=g)
dy1 = torch.randn(n, 16, 16, 16, device=dev, generator=g)
reps = int(sys.argv[1]) if len(sys.argv) > 1 else 2
for _ in range(reps):
    flush.zero_(); meta = ops.pack_cells(mask, action, cells)
    packed = ops.pack_actor_weight(weight, bias, cells)
    flush.zero_(); _, _, _, logits = ops.actor_head_fwd(hidden, packed, meta, want_logits="alive")
    flush.zero_(); ops.actor_head_list_fwd(hidden, packed, meta, want_logits="alive")
    flush.zero_(); ops.cells_bwd(hidden, weight, logits, meta, gl, ge)
    flush.zero_(); ops.head_sample(logits_dense, mask, u)
    flush.zero_(); ops.conv3x3_fwd(x32, w32, b32, wmode=1)
    flush.zero_(); ops.conv3x3_fwd(x16, w16, b16, wmode=1)
    flush.zero_(); ops.conv1_bits_fwd(obs_bits, w1, b1)
    flush.zero_(); ops.conv1_bits_wgrad(obs_bits, dy1)
    flush.zero_(); ops.vtrace_loss(vt["lp_old"], vt["lp_new"], vt["values"], vt["bootstrap"], vt["rewards"], vt["done"], vt["entropy"])
torch.cuda.synchronize()
ops.check_status(dev)
print("profile_r2 ok")
