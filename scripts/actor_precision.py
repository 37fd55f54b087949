"""Which actor-Linear path is closest to exactly rounded logits?  (dense tcgen05 3xTF32, dense cuBLAS fp32, cell-list bmm)"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests"))
import torch
from microbeast_b200 import model as mm, ops, synth
from conftest import close_rms
from test_gpu_learner import _pair
torch.backends.cuda.matmul.allow_tf32 = False
for preset, h in [("dense", 8), ("realistic", 8), ("all_valid", 8), ("dense", 16)]:
    mine, _ = _pair(h, h, seed=31)
    batch = synth.make_rollout_batch(h, h, 6, 5, seed=12, preset=preset, full_keys=False)
    dev = {k: v.cuda() for k, v in batch.items()}
    with torch.no_grad():
        hidden = mine.forward(dev, (), True, ())
        l64 = (hidden.double() @ mine.actor.weight.double().t() + mine.actor.bias.double())
        lp_x, en_x = ops.head_eval_fwd(l64.float().contiguous(), dev["action_mask"], dev["action"])
        res = {}
        mm.TENSOR_CORE_ACTOR = True
        res["tcgen05"] = mine.get_action(dev, learning=True)[0]
        mm.TENSOR_CORE_ACTOR = False
        res["cublas"] = mine.get_action(dev, learning=True)[0]
        mm.TENSOR_CORE_ACTOR = True
        mine.sparse_head = True
        res["cells_bmm"] = mine.get_action(dev, learning=True)[0]
        mine.sparse_head = False
    print(f"{preset}-{h}: max|logit| {float(l64.abs().max()):.1f}  rms(entropy) {float(en_x.pow(2).mean().sqrt()):.3f}")
    for name, out in res.items():
        e_lp = close_rms(out["logprobs"], lp_x, 1e-5)[1]
        e_en = close_rms(out["entropy"], en_x, 1e-5)[1]
        lg = out["policy_logits"]
        e_lg = float((lg.double() - l64).abs().max() / l64.abs().max()) if lg is not None else float("nan")
        print(f"   {name:10s} logits err/max {e_lg:.2e}   logp worst {e_lp:.2f} x1e-5   entropy worst {e_en:.2f} x1e-5")
