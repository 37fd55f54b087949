"""Launches the encoder-side kernels (csrc/convgrad.cu, csrc/gemm.cu) and the cell-list kernels once each at
BASELINE config 2 sizes -- the command ncu wraps for profiles/r01_ncu_encoder_kernels.csv."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import torch.nn.functional as F
from microbeast_b200 import ops, synth
dev = torch.device("cuda", 0)
n = 20736
g = torch.Generator(device=dev).manual_seed(2)
reps = int(sys.argv[1]) if len(sys.argv) > 1 else 2
flush = torch.empty(512 << 20, dtype=torch.uint8, device=dev)
shapes = [(27, 16, 16), (16, 16, 8), (16, 32, 8), (32, 32, 4), (32, 32, 2)]
data = []
for cin, cout, h in shapes:
    x = torch.relu(torch.randn(n, cin, h, h, device=dev, generator=g))
    dy = torch.randn(n, cout, h, h, device=dev, generator=g)
    wk = (torch.randn(cin, cout, 9, device=dev, generator=g) * 0.1).contiguous()
    b = torch.randn(cout, device=dev, generator=g)
    data.append((x, dy, wk, b))
hid = torch.randn(n, 256, device=dev, generator=g)
w = torch.randn(19968, 256, device=dev, generator=g) * 0.02
bias = torch.randn(19968, device=dev, generator=g)
px = torch.randn(n, 16, 16, 16, device=dev, generator=g)
py, pidx = F.max_pool2d(px, 3, 2, 1, return_indices=True)
mask = synth.make_masks(n, 256, "dense", gen=g, device=dev)
for _ in range(reps):
    for x, dy, wk, b in data:
        flush.zero_(); ops.conv3x3_wgrad(x, dy)
        flush.zero_(); ops.conv3x3_fwd(x, wk, b)
    flush.zero_(); ops.linear_tf32x3(hid, w, bias)
    flush.zero_(); ops.maxpool3s2_bwd(py, pidx, 16, 16)
    flush.zero_(); ops.active_cells(mask, 256)
torch.cuda.synchronize()
ops.check_status(dev)
print("profile_encoder_kernels ok")
