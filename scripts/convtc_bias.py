"""Signed error of the tensor-core convolution against float64: is it biased towards zero (truncating accumulation)?"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import torch.nn.functional as F
from microbeast_b200 import ops
from microbeast_b200._lib import lib
dev = torch.device("cuda", 0)
g = torch.Generator(device="cuda").manual_seed(0)
for cin, cout, h in [(32, 32, 4), (16, 16, 8)]:
    x = torch.relu(torch.randn(4096, cin, h, h, generator=g, device=dev))
    wt, b = torch.randn(cout, cin, 3, 3, generator=g, device=dev) * 0.1, torch.randn(cout, generator=g, device=dev) * 0.1
    want = F.conv2d(x.double(), wt.double(), b.double(), padding=1)
    big = want.abs() > 0.25 * want.abs().max()
    for tc in (1, 0):
        lib.mb_set_tuning(b"conv_tc", tc)
        got = ops.conv3x3_fwd(x, wt, b, wmode=1).double()
        rel = ((got - want) / want)[big]                     # > 0: magnitude too large, < 0: too small (towards zero)
        print(f"cin={cin} cout={cout} h={h} {'tc ' if tc else 'fma'}: mean signed rel err {float(rel.mean()):+.3e}  rms {float(rel.pow(2).mean().sqrt()):.3e}  (2^-24 = 5.96e-8)")
lib.mb_set_tuning(b"conv_tc", 1)
