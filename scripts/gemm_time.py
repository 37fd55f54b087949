"""Time mb_linear_tf32x3 against cuBLAS fp32 SGEMM at the config-2 actor Linear (20736 x 256 -> 19968)."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from microbeast_b200 import ops
torch.backends.cuda.matmul.allow_tf32 = False
m, k, n = 20736, 256, 19968
x = torch.randn(m, k, device="cuda"); w = torch.randn(n, k, device="cuda") * 0.05; b = torch.randn(n, device="cuda")

def timeit(fn, reps=10):
    for _ in range(3): fn()
    torch.cuda.synchronize()
    a, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(reps): fn()
    e.record(); torch.cuda.synchronize()
    return a.elapsed_time(e) / reps

t_c = timeit(lambda: torch.addmm(b, x, w.t()))
t_m = timeit(lambda: ops.linear_tf32x3(x, w, b))
ops.start_kernel_timing(); ops.linear_tf32x3(x, w, b); kt = ops.stop_kernel_timing()
print(f"cuBLAS fp32 addmm {t_c:.3f} ms; mb_linear_tf32x3 incl. splits {t_m:.3f} ms (x{t_c / t_m:.2f}); kernels: "
      + ", ".join(f"{k_}={sum(v)/len(v):.3f} ms" for k_, v in kt.items()))
ref = x[:512].double() @ w.double().t() + b.double()
got = ops.linear_tf32x3(x, w, b)[:512]
print("max err / max|ref|:", float((got.double() - ref).abs().max() / ref.abs().max()),
      " cuBLAS:", float((torch.addmm(b, x, w.t())[:512].double() - ref).abs().max() / ref.abs().max()))

import microbeast_b200
for gm in (0, 4, 8, 16, 32):
    for ss in (0, 1):
        microbeast_b200.set_tuning("gemm_group_m", gm); microbeast_b200.set_tuning("gemm_stream_store", ss)
        a_hi, a_lo = ops.split_tf32(x); b_hi, b_lo = ops.split_tf32(w)
        ops.start_kernel_timing()
        for _ in range(6): ops.linear_tf32x3(x, w, b)
        kt = ops.stop_kernel_timing()["linear_tf32x3"]
        ok = torch.equal(ops.linear_tf32x3(x, w, b)[:512], got)
        print(f"group_m={gm:2d} stream_store={ss}: {sorted(kt)[len(kt)//2]:.3f} ms  same result: {ok}", flush=True)
