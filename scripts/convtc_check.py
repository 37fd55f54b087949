"""Tensor-core 3x3 convolution (csrc/convtc.cu) against float64 conv2d, then against the FMA kernel on time."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import torch.nn.functional as F
from microbeast_b200 import ops
from microbeast_b200._lib import lib
dev = torch.device("cuda", 0)
g = torch.Generator(device="cuda").manual_seed(0)


def rnd(*shape):
    return torch.randn(shape, generator=g, device=dev)


worst = 0.0
for cin, cout, h in [(32, 32, 4), (16, 16, 8), (16, 32, 8), (32, 16, 8), (32, 32, 2), (16, 16, 4), (16, 32, 4), (32, 16, 4), (32, 32, 8)]:
    for n in (1, 5, 301):
        x, wt, b = rnd(n, cin, h, h), rnd(cout, cin, 3, 3) * 0.2, rnd(cout)
        want = F.conv2d(x.double(), wt.double(), b.double(), padding=1)
        lib.mb_set_tuning(b"conv_tc", 1)
        got = ops.conv3x3_fwd(x, wt, b, wmode=1)
        torch.cuda.synchronize()
        err = float((got.double() - want).abs().max() / want.abs().max())
        lib.mb_set_tuning(b"conv_tc", 0)
        old = ops.conv3x3_fwd(x, wt, b, wmode=1)
        err_old = float((old.double() - want).abs().max() / want.abs().max())
        # data gradient form + fused tails
        wt2 = rnd(cin, cout, 3, 3) * 0.2                      # weight of a conv cout -> cin whose input gradient we take
        msk, add = rnd(n, cout, h, h), rnd(n, cout, h, h)
        want2 = F.conv_transpose2d(torch.relu(x).double(), wt2.double(), padding=1)
        want2 = torch.where(msk > 0, want2, torch.zeros_like(want2)) + add.double()
        lib.mb_set_tuning(b"conv_tc", 1)
        got2 = ops.conv3x3_fwd(x, wt2, None, wmode=2, relu_in=True, mask_src=msk, add_src=add)
        torch.cuda.synchronize()
        err2 = float((got2.double() - want2).abs().max() / want2.abs().max())
        worst = max(worst, err, err2)
        print(f"cin={cin} cout={cout} h={h} n={n}: tc {err:.2e} (fma {err_old:.2e})  dgrad+tails {err2:.2e}", flush=True)
print("worst", worst)
assert worst < 2e-6, worst
n = 20736
for cin, cout, h in [(32, 32, 4), (16, 16, 8), (16, 32, 8), (32, 16, 8), (32, 32, 2)]:
    x, wt, b = rnd(n, cin, h, h), rnd(cout, cin, 3, 3) * 0.2, rnd(cout)
    msk, add = rnd(n, cout, h, h), rnd(n, cout, h, h)
    line = f"cin={cin} cout={cout} h={h} n={n}:"
    for tc in (1, 0):
        lib.mb_set_tuning(b"conv_tc", tc)
        for fused in (False, True):
            kw = dict(relu_in=True, mask_src=msk, add_src=add) if fused else {}
            for _ in range(3):
                ops.conv3x3_fwd(x, wt, b, wmode=1, **kw)
            s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            s.record()
            for _ in range(20):
                ops.conv3x3_fwd(x, wt, b, wmode=1, **kw)
            e.record(); torch.cuda.synchronize()
            line += f"  {'tc ' if tc else 'fma'}{'+tails' if fused else ''} {s.elapsed_time(e) / 20 * 1e3:.1f} us"
    print(line, flush=True)
lib.mb_set_tuning(b"conv_tc", 1)
