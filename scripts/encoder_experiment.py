"""Times the PyTorch encoder + actor Linear (forward + backward) at BASELINE config 2 under
different memory formats / math modes, to choose the host-side configuration of the step."""
import sys, os, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from microbeast_b200 import synth
from microbeast_b200.model import Agent

dev = torch.device("cuda", 0)
h = w = 16; T = 80; rows = 256; cells = h * w; n = rows * (T + 1)
g = torch.Generator(device=dev).manual_seed(0)
obs = synth.make_obs(n, h, w, gen=g, device=dev, dtype=torch.uint8)

def run(name, channels_last, tf32, benchmark=True, nchw_copy=False, cudnn=True, deterministic=False, fast_bwd=False):
    import microbeast_b200.model as mm
    mm.FAST_CONV_BACKWARD = fast_bwd
    torch.backends.cudnn.enabled = cudnn
    torch.backends.cudnn.deterministic = deterministic
    torch.backends.cudnn.allow_tf32 = tf32
    torch.backends.cuda.matmul.allow_tf32 = tf32
    torch.backends.cudnn.benchmark = benchmark
    torch.manual_seed(0)
    m = Agent((h, w, 27), [6, 4, 4, 4, 4, 7, 49] * cells, cells).to(dev)
    if channels_last:
        m = m.to(memory_format=torch.channels_last)
    def step():
        x = obs.float().permute(0, 3, 1, 2)
        if nchw_copy:
            x = x.contiguous()
        hid = m.network(x)
        logits = m.actor(hid); v = m.critic(hid)
        (logits.sum() * 1e-6 + v.sum()).backward()
        for p in m.parameters():
            p.grad = None
    for _ in range(3):
        step()
    torch.cuda.synchronize()
    s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    s.record()
    for _ in range(5):
        step()
    e.record(); torch.cuda.synchronize()
    print(f"{name:40s} {s.elapsed_time(e) / 5:8.2f} ms/iter", flush=True)

run("nchw-copy fp32, dgrad as forward conv", False, False, nchw_copy=True, fast_bwd=True)
run("channels_last fp32, dgrad as forward conv", True, False, fast_bwd=True)
run("channels_last fp32", True, False)
run("nchw-copy + nchw weights fp32", False, False, nchw_copy=True)
run("nhwc view + nchw weights fp32", False, False)
run("nchw-copy fp32 (no cudnn benchmark)", False, False, benchmark=False, nchw_copy=True)
run("channels_last tf32 (info only)", True, True)
run("nchw-copy tf32 (info only)", False, True, nchw_copy=True)
