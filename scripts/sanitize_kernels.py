"""Tiny invocations of every kernel (all tile shapes, ragged tiles, TMA and generic V-trace) for
compute-sanitizer:  compute-sanitizer --tool memcheck|racecheck python scripts/sanitize_kernels.py"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import microbeast_b200
from microbeast_b200 import buffers as mbuf, ops, synth

dev = torch.device("cuda", 0)
g = torch.Generator().manual_seed(0)
for cells, n in ((64, 5), (100, 3), (9, 4), (256, 2)):
    logits = synth.make_logits(n, cells, gen=g).to(dev)
    mask = (synth.make_masks(n, cells, "all_valid", gen=g) * (torch.rand(n, 78 * cells, generator=g) < 0.5)).to(dev)
    act = synth.make_valid_actions(mask.cpu(), gen=g).to(dev)
    u = synth.make_uniforms(n, cells, gen=g).to(dev)
    u[0, :7] = torch.tensor([0.0, 0.25, 0.5, 0.75, 1 / 3, 1 / 7, 0.999999])
    gl = torch.randn(n, generator=g).to(dev)
    for variant in (-1, 0, 1, 3, 5, 7):
        microbeast_b200.set_tuning("head_variant", variant)
        ops.head_eval_fwd(logits, mask, act)
        ops.head_eval_bwd(logits, mask, act, gl, gl)
        ops.head_sample(logits, mask, u)
    microbeast_b200.set_tuning("head_variant", -1)
    m2, a2 = ops.unpack_rollout(mbuf.pack_mask(mask.cpu()).to(dev), act.to(torch.uint8), cells)
    assert torch.equal(m2, (mask != 0).to(torch.uint8)) and torch.equal(a2, act)
for rows, T in ((70, 80), (33, 32), (5, 48), (40, 20), (3, 7)):
    t = {k: v.to(dev) for k, v in synth.make_vtrace_inputs(rows, T, gen=g).items()}
    for variant in (-1, 0, 1, 3, 4):
        microbeast_b200.set_tuning("vtrace_variant", variant)
        ops.vtrace_loss(t["lp_old"], t["lp_new"], t["values"], t["bootstrap"], t["rewards"], t["done"], t["entropy"])
    microbeast_b200.set_tuning("vtrace_variant", -1)
torch.cuda.synchronize()
ops.check_status(dev)
print("sanitize_kernels ok")
