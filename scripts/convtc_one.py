"""One tensor-core convolution launch per shape, for ncu."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from microbeast_b200 import ops
dev = torch.device("cuda", 0)
n = 20736
for cin, cout, h in [(32, 32, 4), (16, 16, 8)]:
    x, wt, b = torch.randn(n, cin, h, h, device=dev), torch.randn(cout, cin, 3, 3, device=dev) * 0.2, torch.randn(cout, device=dev)
    for _ in range(2):
        ops.conv3x3_fwd(x, wt, b, wmode=1)
torch.cuda.synchronize()
