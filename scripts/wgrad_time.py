"""Time mb_conv3x3_wgrad against cuDNN's fp32 wgrad at the config-2 encoder shapes (N = 20736)."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from microbeast_b200 import ops
torch.backends.cudnn.allow_tf32 = False; torch.backends.cudnn.benchmark = True
n = int(sys.argv[1]) if len(sys.argv) > 1 else 20736

def timeit(fn, reps=10):
    for _ in range(3): fn()
    torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(reps): fn()
    b.record(); torch.cuda.synchronize()
    return a.elapsed_time(b) / reps

for cin, cout, h in [(27, 16, 16), (16, 16, 8), (16, 32, 8), (32, 32, 4), (32, 32, 2)]:
    x = torch.relu(torch.randn(n, cin, h, h, device="cuda"))
    dy = torch.randn(n, cout, h, h, device="cuda")
    w = torch.zeros(cout, cin, 3, 3, device="cuda")
    cudnn = lambda: torch.ops.aten.convolution_backward(dy, x, w, [cout], [1, 1], [1, 1], [1, 1], False, [0, 0], 1, [False, True, True])
    mine = lambda: ops.conv3x3_wgrad(x, dy)
    t_c, t_m = timeit(cudnn), timeit(mine)
    useful = 2.0 * n * cout * cin * (3 * h - 2) ** 2
    print(f"cin={cin} cout={cout} {h}x{h}: cudnn {t_c*1e3:8.1f} us   mb_conv3x3_wgrad+bias {t_m*1e3:8.1f} us   "
          f"({useful / t_m / 1e9:.1f} TFLOP/s useful, x{t_c / t_m:.2f})", flush=True)

print("forward-type convolution (forward and data gradient):")
for cin, cout, h in [(27, 16, 16), (16, 16, 8), (16, 32, 8), (32, 16, 8), (32, 32, 4), (32, 32, 2)]:
    x = torch.relu(torch.randn(n, cin, h, h, device="cuda"))
    w = torch.randn(cout, cin, 3, 3, device="cuda") * 0.1
    b = torch.randn(cout, device="cuda")
    wk = w.permute(1, 0, 2, 3).reshape(cin, cout, 9).contiguous()
    cudnn = lambda: torch.nn.functional.conv2d(x, w, b, padding=1)
    mine = lambda: ops.conv3x3_fwd(x, wk, b)
    t_c, t_m = timeit(cudnn), timeit(mine)
    useful = 2.0 * n * cout * cin * (3 * h - 2) ** 2
    print(f"cin={cin} cout={cout} {h}x{h}: cudnn {t_c*1e3:8.1f} us   mb_conv3x3_fwd {t_m*1e3:8.1f} us   "
          f"({useful / t_m / 1e9:.1f} TFLOP/s useful, x{t_c / t_m:.2f})", flush=True)
