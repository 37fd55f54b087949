"""Fixed cost of a csrc/convtc.cu launch: kernel time against the number of frames (event pairs around each launch)."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from microbeast_b200 import ops
from microbeast_b200._lib import lib
dev = torch.device("cuda", 0)
for cin, cout, h in [(32, 32, 4), (16, 16, 8)]:
    for tc in (1, 0):
        lib.mb_set_tuning(b"conv_tc", tc)
        line = f"cin={cin} cout={cout} h={h} {'tc ' if tc else 'fma'}:"
        for n in (1, 64, 1184, 4736, 20736, 82944):
            x, wt, b = torch.randn(n, cin, h, h, device=dev), torch.randn(cout, cin, 3, 3, device=dev) * 0.2, torch.randn(cout, device=dev)
            for _ in range(3):
                ops.conv3x3_fwd(x, wt, b, wmode=1)
            ops.start_kernel_timing()
            for _ in range(20):
                ops.conv3x3_fwd(x, wt, b, wmode=1)
            kt = ops.stop_kernel_timing()
            ms = sorted(v for vs in kt.values() for v in vs)
            line += f"  n={n}: {ms[len(ms) // 2] * 1e3:.1f} us"
        print(line, flush=True)
lib.mb_set_tuning(b"conv_tc", 1)
