"""C1 parity (8x8, T=80, 32 rows, golden N(0, 0.15^2) weights) with the convolutions on / off the tensor cores."""
import os, sys, json
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests"))
import torch
torch.backends.cudnn.allow_tf32 = False; torch.backends.cuda.matmul.allow_tf32 = False
import test_gpu_parity as tp
from oracle import parity
from microbeast_b200._lib import lib
for seed in (41, 42, 45):
    mine, ref = tp._pair(8, 8, seed=seed)
    batch = tp._batch(ref, 8, 8, 80, 32, seed=1234)
    for tc in (1, 0):
        lib.mb_set_tuning(b"conv_tc", tc)
        rep = parity.step_report(tp.gpu_step(mine, batch, 32, 80), ref, batch, 32, 80)
        g, c = rep["gpu_vs_fp64"], rep["cpu_fp32_vs_fp64"]
        print("seed", seed, "tc" if tc else "fma", {k: (f"{g[k]:.2e}", f"{c[k]:.2e}") for k in ("logprobs", "entropy", "baseline", "vs", "pg_advantages", "grad_max")}, flush=True)
lib.mb_set_tuning(b"conv_tc", 1)
