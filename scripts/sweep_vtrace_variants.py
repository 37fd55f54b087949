"""V-trace kernel: every pipeline shape x the bandwidth-bound shapes of BASELINE config 5 (one JSON line each).

    python scripts/sweep_vtrace_variants.py [out.jsonl] [variants, e.g. 2,4,9,12]

L2 is flushed between launches (a 512 MB memset); median of 10 after 3 warm-ups; algorithmic bytes
37 B / element + 4 B / row against MEASURED_PEAKS.json.
"""
import json
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import microbeast_b200                       # noqa: E402
from microbeast_b200 import ops, synth       # noqa: E402

NAMES = {-1: "default", 0: "2x32", 1: "3x32", 2: "4x32", 3: "3x16", 4: "4x16", 5: "5x16", 6: "6x16", 7: "8x16", 8: "6x32",
         9: "12x16", 10: "8x32", 11: "4x32 r16", 12: "6x32 r16", 13: "8x32 r16", 14: "8x16 r16", 15: "6x16 r16"}


def main():
    out_path = sys.argv[1] if len(sys.argv) > 1 else ""
    variants = [int(v) for v in sys.argv[2].split(",")] if len(sys.argv) > 2 else sorted(NAMES)
    try:
        peak = float(json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))["hbm_gbs"])
    except Exception:
        peak = 6650.0
    dev = torch.device("cuda", 0)
    g = torch.Generator(device=dev).manual_seed(1)
    flush = torch.empty(512 << 20, dtype=torch.uint8, device=dev)
    shapes = [(512, 65536), (256, 65536), (80, 65536), (512, 16384), (256, 16384), (80, 16384), (512, 4096), (80, 256)]
    lines = []
    for T, R in shapes:
        nbytes = 37 * R * T + 4 * R
        # sustained: back-to-back launches over rotating buffer sets whose total footprint is >= 1 GB (>> 126 MB L2),
        # so every launch reads cold inputs and the only dirty lines in L2 are the previous launch's own outputs
        set_bytes = 2 * 41 * R * T                       # 21 B inputs + 16 B outputs + 4 B entropy... per element, rounded up
        nsets = max(4, min(64, -(-(1 << 30) // set_bytes)))
        sets = [synth.make_vtrace_inputs(R, T, gen=g, device=dev) for _ in range(nsets)]
        t = sets[0]
        for v in variants:
            microbeast_b200.set_tuning("vtrace_variant", v)
            fn = lambda: ops.vtrace_loss(t["lp_old"], t["lp_new"], t["values"], t["bootstrap"], t["rewards"],
                                         t["done"], t["entropy"])
            ts = []
            for i in range(13):
                # the flush also keeps the GPU busy (~0.3 ms) while the host enqueues the call: the start event must not
                # be reached before the launch has been queued, or host launch latency is timed as kernel time
                flush.zero_(); flush.zero_(); flush.zero_()
                s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                s.record(); fn(); e.record()
                torch.cuda.synchronize()
                if i >= 3:
                    ts.append(s.elapsed_time(e))
            ts.sort()
            med = ts[len(ts) // 2]
            gbs = nbytes / (med * 1e-3) / 1e9
            # rotating sets, called straight through ctypes with pre-built argument tuples and pre-allocated outputs:
            # the host side of a call is then ~3 us and the timed region holds back-to-back kernel launches only
            import ctypes
            from microbeast_b200._lib import lib
            calls = []
            for x in sets:
                o = {k: torch.empty((R, T), dtype=torch.float32, device=dev) for k in ("vs", "adv", "dlp", "dv")}
                o["loss4"] = torch.empty(4, dtype=torch.float32, device=dev)
                ws = ops.vtrace_workspace(dev, R)
                P = lambda z: ctypes.c_void_p(z.data_ptr())
                calls.append(((P(x["lp_old"]), P(x["lp_new"]), P(x["values"]), P(x["bootstrap"]), P(x["rewards"]),
                               P(x["done"]), P(x["entropy"]), R, T, T, T, 0.99, 1.0, 1.0, 1.0, 0.01, 0.0,
                               P(o["vs"]), P(o["adv"]), P(o["loss4"]), P(o["dlp"]), P(o["dv"]), P(ws), ws.numel(),
                               ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)), o))
            reps = max(2 * nsets, 16)
            def burst(n):
                for i in range(n):
                    rc = lib.mb_vtrace_loss(*calls[i % nsets][0])
                    assert rc == 0
            burst(nsets)
            flush.zero_(); flush.zero_(); flush.zero_()
            s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            s.record(); burst(reps); e.record()
            torch.cuda.synchronize()
            del calls
            sus = s.elapsed_time(e) / reps
            gbs_sus = nbytes / (sus * 1e-3) / 1e9
            line = {"kernel": "vtrace_loss", "variant": v, "shape": NAMES.get(v, str(v)), "T": T, "rows": R,
                    "median_us": round(med * 1e3, 2), "min_us": round(ts[0] * 1e3, 2), "gbs": round(gbs, 1),
                    "frac_of_measured": round(gbs / peak, 4), "frac_of_8TBs": round(gbs / 8000.0, 4), "peak_gbs": peak,
                    "sustained_us": round(sus * 1e3, 2), "sustained_gbs": round(gbs_sus, 1),
                    "sustained_frac_of_measured": round(gbs_sus / peak, 4), "rotating_sets": nsets,
                    "timing": "median_us: one launch after a 512 MB write-flush of L2 (its dirty lines are written back "
                              "during the launch); sustained_us: back-to-back launches over rotating input sets >= 1 GB"}
            lines.append(line)
            print(json.dumps(line), flush=True)
        del t, sets
    microbeast_b200.set_tuning("vtrace_variant", -1)
    if out_path:
        with open(out_path, "w") as f:
            for line in lines:
                f.write(json.dumps(line) + "\n")


if __name__ == "__main__":
    main()
