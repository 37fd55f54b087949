"""Which role bounds csrc/convtc.cu?  Times each shape with one role's work skipped (results are wrong by design)."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from microbeast_b200 import ops
from microbeast_b200._lib import lib
dev = torch.device("cuda", 0)
n = 20736
for cin, cout, h in [(32, 32, 4), (16, 16, 8), (16, 32, 8), (32, 16, 8), (32, 32, 2)]:
    x, wt, b = torch.randn(n, cin, h, h, device=dev), torch.randn(cout, cin, 3, 3, device=dev) * 0.2, torch.randn(cout, device=dev)
    line = f"cin={cin} cout={cout} h={h}:"
    for name, bits in (("all", 0), ("no-epilogue", 1), ("no-convert", 2), ("no-mma", 4), ("no-epi-no-conv", 3), ("nothing", 7)):
        lib.mb_set_tuning(b"conv_tc_debug", bits)
        for _ in range(3):
            ops.conv3x3_fwd(x, wt, b, wmode=1)
        ops.start_kernel_timing()
        for _ in range(20):
            ops.conv3x3_fwd(x, wt, b, wmode=1)
        kt = ops.stop_kernel_timing()
        ms = sorted(v for vs in kt.values() for v in vs)
        line += f"  {name} {ms[len(ms) // 2] * 1e3:.1f}"
    print(line, flush=True)
lib.mb_set_tuning(b"conv_tc_debug", 0)
