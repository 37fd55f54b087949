#!/usr/bin/env python
"""Benchmark of the learner hot path (BASELINE.json metric: learner env-frames/s).

    python bench.py --gpus 1 --steps 20 --warmup 5                 # our arm, one JSON line
    python bench.py --impl reference --steps 3 --warmup 1          # reference CPU arm (the staged reference itself)
    torchrun --nnodes=1 --nproc-per-node N ... bench.py --gpus N   # data-parallel, weak scaling
    python bench.py --sweep vtrace | --sweep head                  # kernel-only tables (profiles/)

A "step" is one learner update (reference libs/utils.py:223-342: policy evaluation on the
batch, V-trace targets, the three losses, backward, Adam) on one synthetic batch
(SURVEY.md 8d).  Workload at N GPUs: BASELINE config 2 per GPU -- 16x16 map, T=80, 256 rows
per rank (N=4 is config 3's global batch of 1024 rows).  frames/step = rows*T as the
reference counts them (microbeast.py:230).

value   = whole-job frames/s with the batch already resident in HBM.
e2e     = the same step through the public calls BatchPrefetcher.submit(...) + PPO_learn(...) with the batch in
          pinned HOST memory: host->device copies of that step's inputs (staged one step ahead on a side stream)
          and the device->host read of the four losses are inside the timed region.
parity  = the gate run before timing: one lr = 0 step on a 32-row sample of the benchmarked workload, GPU fp32 and
          CPU fp32 (the reference's arithmetic) both against a float64 evaluation (oracle/parity.py); the bench exits 3
          instead of printing a number when the GPU is further off than max(1e-5, 4 x the CPU's error).
roofline= the dominant kernel of ours -- the actor head's backward over the (frame, cell) pairs with a valid action
          (csrc/cells.cu), an fp32 FMA-pipe kernel: useful TFLOP/s, timed live with CUDA events on the launching stream
          inside the timed region, against 148 SMs x 128 FMA lanes x the sampled SM clock; the dominant tensor-pipe
          kernel (the fused actor Linear + policy head, dense and list form) sits under info.roofline_tensor, the
          dominant HBM-bound one under info.roofline_hbm.
cpu_baseline = the reference's own PPO_learn + Agent (byte copies staged into oracle/_ref by oracle/stage_ref.py;
          the oracle port when they are missing) timed on the host cores on a bounded sample of the same workload
          (rank 0, N=1 only).  --impl reference runs only that arm and never imports the package.
"""
from __future__ import annotations

import argparse
import json
import os
import statistics
import subprocess
import sys
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

METRIC = "learner_env_frames_per_sec"
UNIT = "frames/s"
FALLBACK_HBM_GBS = 6650.0          # B200_PROFILING.md fallback when MEASURED_PEAKS.json is absent

WORKLOADS = {
    # name: (h, w, T, rows per GPU)
    "c1": (8, 8, 80, 32),
    "c2": (16, 16, 80, 256),
    "c4": (32, 32, 128, 64),
    # configs[2]: B=1024 rows in total, split over the ranks (strong scaling); rows per GPU = 1024 / N
    "c3": (16, 16, 80, 1024),
}


def parse_args():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--workload", default="c2", choices=sorted(WORKLOADS))
    ap.add_argument("--rows", type=int, default=0, help="rows per GPU (default: the workload's)")
    ap.add_argument("--preset", default="dense")
    ap.add_argument("--cpu-rows", type=int, default=32, help="rows of the CPU-baseline sample")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--no-parity", action="store_true", help="skip the parity gate that runs before any timing")
    ap.add_argument("--parity-rows", type=int, default=32, help="rows of the parity-gate sample")
    ap.add_argument("--no-graph-info", action="store_true")
    ap.add_argument("--sparse-head", action="store_true",
                    help="headline with Agent.sparse_head (actor logits only for cells with a valid action)")
    ap.add_argument("--no-tf32-info", dest="tf32_info", action="store_false",
                    help="skip the informational leg that times the step with TF32 allowed in the PyTorch encoder")
    ap.add_argument("--wire", default="compact", choices=["compact", "plain"],
                    help="host batch format of the e2e leg (compact = the rollout store's wire format)")
    ap.add_argument("--sweep", default="", choices=["", "vtrace", "head", "variants"])
    ap.add_argument("--out", default="", help="also append the JSON line(s) to this file")
    return ap.parse_args()


def hbm_peak():
    try:
        with open(os.path.join(ROOT, "MEASURED_PEAKS.json")) as f:
            return float(json.load(f)["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
    except Exception:
        return FALLBACK_HBM_GBS, "fallback (B200_PROFILING.md)"


class ClockSampler:
    """nvidia-smi clocks / throttle reasons sampled DURING the timed region."""
    Q = ("clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap,timestamp")

    def __init__(self, index: int):
        self.index, self.proc = index, None
        self.t_begin = self.t_end = None

    # nvidia-smi needs ~0.1 s to come up and the timed region may be shorter than that, so the sampler is started
    # before the warm-up steps (the same load) and the samples are filtered by time stamp afterwards
    def mark_begin(self):
        self.t_begin = time.time()

    def mark_end(self):
        self.t_end = time.time()

    def start(self):
        try:
            self.proc = subprocess.Popen(
                ["nvidia-smi", "-i", str(self.index), f"--query-gpu={self.Q}", "--format=csv,noheader,nounits",
                 "-lms", "20"], stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
        except Exception:
            self.proc = None

    def stop(self):
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.15)
        self.proc.terminate()
        try:
            out, _ = self.proc.communicate(timeout=5)
        except Exception:
            self.proc.kill()
            out = ""
        import datetime
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        rows = []
        for line in out.strip().splitlines():
            f = [x.strip() for x in line.split(",")]
            if len(f) < 6:
                continue
            try:
                rec = (float(f[0]), float(f[1]), [n for n, v in zip(names, f[2:6]) if v.lower().startswith("active")])
            except ValueError:
                continue
            ts = None
            if len(f) > 6:
                try:
                    ts = datetime.datetime.strptime(f[6], "%Y/%m/%d %H:%M:%S.%f").timestamp()
                except ValueError:
                    ts = None
            rows.append((ts, rec))
        inside = [rec for ts, rec in rows if ts is not None and self.t_begin is not None and self.t_end is not None
                  and self.t_begin - 0.02 <= ts <= self.t_end + 0.02]
        window = "timed region"
        if not inside:            # region shorter than the sampling period: the warm-up steps ran the same load
            inside, window = [rec for _, rec in rows], "warm-up + timed region"
        sm = [r[0] for r in inside]; mx = [r[1] for r in inside]
        reasons = sorted({n for r in inside for n in r[2]})
        return {"sm_mhz": statistics.median(sm) if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "samples": len(sm), "window": window, "reasons": reasons}


# ------------------------------------------------------------------------------------------
# reference arm / cpu baseline: the oracle port of the reference's CPU learner step
# ------------------------------------------------------------------------------------------
def load_synth():
    """``microbeast_b200/synth.py`` loaded from its file, WITHOUT importing the package: importing
    ``microbeast_b200`` dlopens ``libmbeast_sm100.so`` and the reference arm must not load any of the product's
    native code (the generators are pure torch)."""
    import importlib.util
    spec = importlib.util.spec_from_file_location("_mb_synth_standalone",
                                                  os.path.join(ROOT, "microbeast_b200", "synth.py"))
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    return mod


def time_cpu_reference(h, w, T, rows, steps, warmup, preset, threads=None, budget_s=None):
    """Times the reference's CPU learner step on ``rows`` rows.  When the unmodified reference can be imported
    (``/root/reference`` here, ``oracle/_ref`` staged by ``oracle/stage_ref.py`` on the GPU box) it is the REAL
    ``libs.utils.PPO_learn`` driving the REAL ``model.Agent`` (``kind: "reference"``, stdout swallowed because
    ``PPO_learn`` prints tensors, ``utils.py:321-340``); else the oracle's restatement (``kind: "port"``).
    Returns ``(frames_per_s, s_per_step, threads, kind)``."""
    import contextlib
    import io
    import torch
    synth = load_synth()
    from oracle import NVEC_CELL, refload
    cores = threads or os.cpu_count() or 1
    torch.set_num_threads(cores)
    torch.manual_seed(1234)
    nvec = list(NVEC_CELL) * (h * w)
    if refload.available():
        ref_model, ref_utils = refload.load()
        agent, kind = ref_model.Agent((h, w, 27), nvec, h * w), "reference"

        def one_step(b):
            with contextlib.redirect_stdout(io.StringIO()):
                return ref_utils.PPO_learn(agent, agent, "bench_ref", b, opt, rows, T, 1)
    else:
        from oracle.learner import RefAgent, ref_learn_step
        agent, kind = RefAgent((h, w, 27), nvec, h * w), "port"

        def one_step(b):
            return ref_learn_step(agent, b, opt, rows, T, 1)
    with torch.no_grad():
        agent.actor.weight.normal_(0, 0.01)          # reference init is all-zero
    opt = torch.optim.Adam(agent.parameters(), lr=2.5e-4, eps=1e-5)        # microbeast.py:200
    batch = synth.make_rollout_batch(h, w, T, rows, seed=1234, preset=preset)
    with torch.no_grad(), contextlib.redirect_stdout(io.StringIO()):
        o, _ = agent.get_action(batch, learning=True)
    batch["logprobs"] = o["logprobs"] + 0.3 * torch.randn(o["logprobs"].shape)
    times = []
    t_start = time.perf_counter()
    for i in range(warmup + steps):
        t0 = time.perf_counter()
        one_step({k: v.clone() for k, v in batch.items()})
        if i >= warmup:
            times.append(time.perf_counter() - t0)
        if budget_s and times and time.perf_counter() - t_start > budget_s:
            break                                    # bounded sample: stop once the time budget is spent
    sec = sum(times) / max(len(times), 1)
    return rows * T / sec, sec, cores, kind


def run_reference(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    h, w, T, _ = WORKLOADS[args.workload]
    rows = args.cpu_rows
    fps, sec, cores, kind = time_cpu_reference(h, w, T, rows, args.steps, args.warmup, args.preset, budget_s=240)
    what = ("the unmodified reference (libs/utils.py PPO_learn + model.py Agent)" if kind == "reference"
            else "oracle port of libs/utils.py:223-342 + model.py:165-216")
    sample = (f"{rows} of {workload_shape(args)[3]} rows per step ({rows * (T + 1)} frames), "
              f"full reference learner step: {what}, torch CPU, {cores} threads")
    cfg = workload_config(args)
    line = {
        "impl": "reference", "metric": METRIC, "value": fps, "unit": UNIT, "n_gpus": args.gpus,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": sec * 1e3, "higher_is_better": True,
        "scaling": "strong" if args.workload == "c3" else "weak", "vs_baseline": None, "dtype": "f32",
        "data": "synthetic",
        "config": cfg,
        "cpu_baseline": {"value": fps, "unit": UNIT, "cores": cores, "kind": kind, "sample": sample},
        "e2e": {"value": fps, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    emit(line, args)


def workload_shape(args):
    """(h, w, T, rows per GPU): c3 is strong scaling -- its 1024 rows are split over the ranks."""
    h, w, T, rows = WORKLOADS[args.workload]
    if args.workload == "c3":
        rows = rows // max(args.gpus, 1)
    return h, w, T, (args.rows or rows)


def workload_config(args, rows_override=None):
    h, w, T, rows = workload_shape(args)
    rows = rows_override or rows
    return {"workload": f"{args.workload}: {h}x{w} basesWorkers-shaped synthetic rollouts, T={T}, "
                        f"{rows} rows per GPU, full learner step",
            "map": f"{h}x{w}", "T": T, "rows_per_gpu": rows, "frames_per_step_per_gpu": rows * T,
            "mask_preset": args.preset, "l2_policy": "inputs larger than L2 (logits alone are "
            f"{rows * (T + 1) * h * w * 78 * 4 / 1e6:.0f} MB per GPU vs 126 MB L2)",
            "parallelism": f"dp{args.gpus}",
            # both arms print this same config; the CPU arm's step is a bounded sample of it and says so here
            "reference_arm_rows_per_step": args.cpu_rows,
            "reference_arm_note": f"--impl reference (and cpu_baseline) time {args.cpu_rows} of the {rows} rows per "
                                  "step on the host cores; its frames/s does not rise with rows (measured 8/32/96)",
            "arithmetic": "fp32 throughout (TF32 off in cuDNN / cuBLAS); the actor Linear forward runs on tcgen05 as three "
                          "TF32 products of hi/lo-split operands with fp32 accumulation (error 5e-7 of max|logit|, "
                          "cuBLAS fp32: 5e-7); its backward runs over the cells with a valid action only (the others "
                          "have exactly zero gradient)"}


def emit(line, args):
    text = json.dumps(line)
    print(text, flush=True)
    if args.out:
        with open(args.out, "a") as f:
            f.write(text + "\n")


# ------------------------------------------------------------------------------------------
# parity gate: runs BEFORE any timing
# ------------------------------------------------------------------------------------------
PARITY_FLOOR, PARITY_SLACK, PARITY_GRAD_CAP = 1e-5, 4.0, 5e-2


def parity_gate(model, h, w, T, rows, preset, timed_rows):
    """One lr = 0 learner step on a `rows`-row sample of the benchmarked workload, on the GPU (the product, same weights,
    same code path as the timed step) and on the CPU through the oracle's fp32 restatement of the reference and
    through a float64 evaluation (oracle/parity.py).  Returns the `parity` object of the JSON line: errors of
    log-prob, entropy, baseline, V-trace targets, advantages, the four losses and the parameter gradients, GPU and
    CPU-fp32 each against float64, and GPU against CPU-fp32.  pass = the GPU is within 1e-5 of float64, or at most
    PARITY_SLACK times as far from it as the reference's own fp32 arithmetic is (gradients capped at 5e-2)."""
    import torch
    from microbeast_b200.learner import LEARNER_KEYS, learner_losses
    from microbeast_b200 import synth
    from oracle import NVEC_CELL, parity
    from oracle.learner import RefAgent
    t_start = time.perf_counter()
    cells = h * w
    ref = RefAgent((h, w, 27), list(NVEC_CELL) * cells, cells)
    ref.load_state_dict({k: v.detach().cpu() for k, v in model.state_dict().items()})
    batch = synth.make_rollout_batch(h, w, T, rows, seed=4321, preset=preset)
    with torch.no_grad():
        o, _ = ref.get_action(batch, learning=True)
    batch["logprobs"] = o["logprobs"] + 0.3 * torch.randn(o["logprobs"].shape, generator=torch.Generator().manual_seed(9))
    dev = next(model.parameters()).device
    dev_batch = {k: batch[k].to(dev) for k in LEARNER_KEYS}
    keep = model.sparse_backward
    # the timed step takes the cell-list backward when it has >= 2^20 (frame, cell) pairs: the gate must too
    if keep == "auto" and timed_rows * (T + 1) * cells >= (1 << 20):
        model.sparse_backward = True
    try:
        model.zero_grad(set_to_none=True)
        loss4, aux = learner_losses(model, dev_batch, rows, T)
        loss4[3].backward()
        torch.cuda.synchronize()
    finally:
        model.sparse_backward = keep
    ours = dict(logprobs=aux["logprobs"].detach().flatten(), entropy=aux["entropy"].detach().flatten(),
                baseline=aux["baseline"].detach().flatten(), vs=aux["vs"].detach(),
                pg_advantages=aux["pg_advantages"].detach(), loss4=loss4.detach(),
                grads={n: p.grad.detach().clone() for n, p in model.named_parameters()})
    model.zero_grad(set_to_none=True)
    rep = parity.step_report(ours, ref, batch, rows, T)
    ok = True
    for key in ("logprobs", "entropy", "baseline", "vs", "pg_advantages"):
        ok &= rep["gpu_vs_fp64"][key] <= max(PARITY_FLOOR, PARITY_SLACK * rep["cpu_fp32_vs_fp64"][key])
    for i in range(4):
        ok &= rep["gpu_vs_fp64"]["loss4_rel"][i] <= max(PARITY_FLOOR, PARITY_SLACK * rep["cpu_fp32_vs_fp64"]["loss4_rel"][i])
    ok &= rep["gpu_vs_fp64"]["grad_max"] <= min(PARITY_GRAD_CAP,
                                                max(PARITY_FLOOR, PARITY_SLACK * rep["cpu_fp32_vs_fp64"]["grad_max"]))
    rep["pass"] = bool(ok)
    rep["rule"] = (f"gpu_vs_fp64 <= max({PARITY_FLOOR:g}, {PARITY_SLACK:g} x cpu_fp32_vs_fp64) for every value and loss; "
                   f"gradients additionally <= {PARITY_GRAD_CAP:g}")
    rep["sample"] = f"{rows} rows ({rows * (T + 1)} frames) of the benchmarked workload, lr = 0, same weights on both sides"
    rep["seconds"] = round(time.perf_counter() - t_start, 1)
    return rep


# ------------------------------------------------------------------------------------------
# our arm
# ------------------------------------------------------------------------------------------
def run_ours(args):
    import torch
    import torch.distributed as dist
    from microbeast_b200 import ops, synth
    from microbeast_b200.learner import LEARNER_KEYS, PPO_learn
    from microbeast_b200.model import Agent
    from microbeast_b200.parallel import GradAllReducer

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device; the product path has no CPU fallback")
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    bound_cores = None
    if world > 1:
        # one process per GPU, on the host cores (and, by first touch, the pinned memory) of the GPU's own socket
        from microbeast_b200.parallel import bind_to_gpu_numa
        bound_cores = bind_to_gpu_numa(local)
        dist.init_process_group("nccl", device_id=dev)
    torch.backends.cudnn.allow_tf32 = False          # dtype f32 means f32: no TF32 anywhere
    torch.backends.cuda.matmul.allow_tf32 = False
    torch.backends.cudnn.benchmark = True

    h, w, T, rows = workload_shape(args)
    cells = h * w
    n = rows * (T + 1)
    nvec = [6, 4, 4, 4, 4, 7, 49] * cells
    torch.manual_seed(1234)                          # identical replica on every rank
    model = Agent((h, w, 27), nvec, cells).to(dev)
    with torch.no_grad():
        model.actor.weight.normal_(0, 0.01)
    model.sparse_head = bool(args.sparse_head)
    opt = torch.optim.Adam(model.parameters(), lr=2.5e-4, eps=1e-5, fused=True)
    reducer = GradAllReducer() if world > 1 else None

    batch = synth.make_rollout_batch(h, w, T, rows, seed=1234 + rank, device=dev, preset=args.preset,
                                     obs_dtype=torch.uint8, full_keys=False)
    with torch.no_grad():
        o, _ = model.get_action(batch, learning=True)
        batch["logprobs"] = o["logprobs"] + 0.3 * torch.randn_like(o["logprobs"])
        del o
    dev_batch = {k: batch[k] for k in LEARNER_KEYS}
    if args.wire == "compact":
        # resident batch in the rollout store's observation format: one int32 per cell (bit c = plane c),
        # expanded to NCHW fp32 by mb_unpack_obs inside the step
        shifts = torch.arange(27, device=dev, dtype=torch.int32)
        dev_batch["obs"] = (batch["obs"].to(torch.int32) << shifts).sum(-1, dtype=torch.int32)

    # ---- parity gate: nothing is timed (and no number is printed) unless the step agrees with the oracle -------
    parity = None
    if not args.no_parity and rank == 0:
        parity = parity_gate(model, h, w, T, args.parity_rows, args.preset, rows)
        if not parity["pass"]:
            emit({"metric": METRIC, "value": None, "unit": UNIT, "error": "parity gate failed: no timing reported",
                  "parity": parity, "config": workload_config(args)}, args)
            raise SystemExit(3)

    def step(b, sync, prefetch=None):
        return PPO_learn(model, model, "bench", b, opt, rows, T, 1, reducer=reducer, sync=sync, prefetch=prefetch)

    def barrier():
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    # kernels that can be the dominant one of the step: only these carry live CUDA-event pairs inside the timed region
    # (an event pair around each of the step's ~120 launches costs 0.3 ms per step; the rest of the `kernels` table
    # comes from a separate, untimed pass)
    LIVE = ("actor_head_fwd", "actor_head_list_fwd", "pack_cells", "cells_bwd", "head_eval_fwd", "head_eval_bwd", "vtrace_loss")

    def timed(fn, steps, warmup, sample_clocks=False, kernel_timing=False):
        sampler = ClockSampler(local) if sample_clocks else None
        if sampler:
            sampler.start()
        for _ in range(warmup):
            fn()
        barrier()
        if sampler:
            sampler.mark_begin()
        launches0 = ops.total_launches()
        if kernel_timing:
            ops.start_kernel_timing(only=LIVE if kernel_timing == "live" else None)
        t0, t1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        t0.record()
        for _ in range(steps):
            fn()
        t1.record()
        barrier()
        if sampler:
            sampler.mark_end()
        ms = t0.elapsed_time(t1)
        kt = ops.stop_kernel_timing() if kernel_timing else {}
        clocks = sampler.stop() if sampler else None
        if world > 1:
            tt = torch.tensor([ms], device=dev)
            dist.all_reduce(tt, op=dist.ReduceOp.MAX)
            ms = float(tt.item())
        return ms / steps, ops.total_launches() - launches0, kt, clocks

    # ---- value: batch resident in HBM -------------------------------------------------
    ms_step, launches, ktimes, clocks = timed(lambda: step(dev_batch, False), args.steps, max(args.warmup, 3),
                                              sample_clocks=True, kernel_timing="live")
    frames = rows * T * world
    value = frames / (ms_step * 1e-3)
    # every other kernel of the step, for the `kernels` table: three more steps, outside the timed region
    _, _, kt_all, _ = timed(lambda: step(dev_batch, False), 3, 1, kernel_timing=True)
    live_names = set(ktimes)
    for name, ts in kt_all.items():
        ktimes.setdefault(name, ts)

    # ---- roofline of the dominant kernel (live CUDA events from the timed region) -------
    peak, peak_src = hbm_peak()
    alg = {
        "head_eval_fwd": 446 * cells * n + 8 * n,
        "head_eval_bwd": 758 * cells * n + 8 * n,
        "vtrace_loss": 37 * rows * T + 4 * rows,
    }
    for name, note in ops.launch_notes.items():       # kernels whose problem size depends on the data / layer
        if name == "active_cells":
            alg[name] = 79 * note["cells"]
        elif name == "head_eval_cells_fwd":
            alg[name] = (446 + 16) * note["cells"]
        elif name == "head_eval_cells_bwd":
            alg[name] = (758 + 8) * note["cells"]
        elif name == "pack_cells":
            alg[name] = (78 + 56 + 16) * note["cells"]          # mask + i64 actions in, one 16-B record out per cell
    sm_mhz = (clocks or {}).get("sm_mhz") or 1965.0
    fp32_peak = 148 * 128 * 2 * sm_mhz * 1e6 / 1e12            # TFLOP/s: 128 FMA lanes per SM at the sampled clock
    kernels = {}
    for name, ts in ktimes.items():
        avg = sum(ts) / len(ts)
        gbs = alg[name] / (avg * 1e-3) / 1e9 if name in alg else None
        kernels[name] = {"launches": len(ts), "avg_ms": avg, "alg_bytes": alg.get(name),
                         "gbs": gbs, "frac_of_peak": (gbs / peak if gbs else None),
                         "timed": "live, inside the timed region" if name in live_names else "separate pass after it"}
        if name.startswith("conv3x3_wgrad") and name in ops.launch_notes:
            nt = ops.launch_notes[name]
            flops = 2.0 * nt["n"] * nt["cout"] * nt["cin"] * (3 * nt["h"] - 2) ** 2      # taps inside the image only
            kernels[name].update({"bound": "fp32 fma", "useful_flops": flops, "tflops": flops / (avg * 1e-3) / 1e12,
                                  "frac_of_fp32_peak": flops / (avg * 1e-3) / 1e12 / fp32_peak,
                                  "fp32_peak_tflops": fp32_peak})
    # the roofline object is for the path's own dominant HBM-bound kernel: the dense policy-head backward when the
    # step launches it, else the dense policy-head forward (the backward then runs on the list of active cells)
    dom = max((k for k in ("head_eval_bwd", "head_eval_fwd", "pack_cells") if k in kernels),
              key=lambda k: kernels[k]["avg_ms"], default=None)
    roof = None
    if dom in kernels:
        roof = {"kernel": dom, "bound": "hbm", "achieved": kernels[dom]["gbs"], "peak": peak, "unit": "GB/s",
                "frac": kernels[dom]["gbs"] / peak, "traffic": None, "peak_source": peak_src,
                "alg_bytes_per_launch": alg[dom], "avg_launch_ms": kernels[dom]["avg_ms"],
                "kernel_share_of_step": kernels[dom]["avg_ms"] / ms_step}
        tr = os.path.join(ROOT, "profiles", "traffic.json")       # dram bytes per launch from ncu --set full
        if os.path.isfile(tr):
            try:
                roof["traffic"] = json.load(open(tr)).get(dom, {}).get(args.workload)
            except Exception:
                pass

    # the fused actor Linear + policy head is tensor-bound: three TF32 products per fp32 product (3xTF32 split precision)
    if "actor_head_fwd" in kernels and "actor_head_fwd" in ops.launch_notes:
        nt = ops.launch_notes["actor_head_fwd"]
        alg_flops = 2.0 * nt["m"] * (78 * nt["cells"]) * nt["k"]
        issued = 3.0 * 2.0 * nt["m"] * (80 * nt["cells"]) * nt["k"]
        sec = kernels["actor_head_fwd"]["avg_ms"] * 1e-3
        try:
            bf16 = float(json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))["bf16_tflops"])
        except Exception:
            bf16 = 1646.0
        kernels["actor_head_fwd"].update({
            "bound": "tensor", "algorithmic_tflops_fp32": alg_flops / sec / 1e12, "issued_tf32_tflops": issued / sec / 1e12,
            "tf32_peak_tflops": bf16 / 2, "frac_of_tf32_peak_issued": issued / sec / 1e12 / (bf16 / 2),
            "note": "fp32-accurate product as three TF32 tensor-core products of hi/lo-split operands; TF32 peak taken as "
                    "half the measured dense bf16 rate of MEASURED_PEAKS.json"})

    roof_hbm = roof
    try:
        bf16_peak = float(json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))["bf16_tflops"])
    except Exception:
        bf16_peak = 1646.0
    roof_tensor = None
    if "cells_bwd" in kernels or "actor_head_list_fwd" in kernels:
        # the (frame, cell) pairs with a valid action: the units of the two list kernels of the actor head
        alive = (dev_batch["action_mask"].view(n, cells, 78) != 0).any(dim=2)
        pairs = int(alive.sum().item())
        tiles = int(((alive.sum(dim=0) + 127) // 128).sum().item())
        if "actor_head_list_fwd" in kernels:
            kf = kernels["actor_head_list_fwd"]
            issued = 3.0 * 2.0 * tiles * 128 * 80 * 256            # three TF32 products, 128-entry tiles of one cell, N = 80
            kf.update({"bound": "tensor", "live_pairs": pairs, "tiles": tiles, "issued_tf32_tflops": issued / (kf["avg_ms"] * 1e-3) / 1e12,
                       "frac_of_tf32_peak_issued": issued / (kf["avg_ms"] * 1e-3) / 1e12 / (bf16_peak / 2),
                       "note": "the call's time includes building the cell lists, zeroing cell_out and the per-frame sums; "
                               "the dense kernel it replaces (actor_head_fwd, info.roofline_tensor) issues 16x the flops"})
        if "cells_bwd" in kernels:
            kb = kernels["cells_bwd"]
            useful = 2.0 * 2.0 * pairs * 78 * 256                  # d hidden and d weight: two [78 x 256] products per live pair
            kb.update({"bound": "fma", "live_pairs": pairs, "useful_flops": useful, "tflops": useful / (kb["avg_ms"] * 1e-3) / 1e12,
                       "fp32_peak_tflops": fp32_peak, "frac_of_fp32_peak": useful / (kb["avg_ms"] * 1e-3) / 1e12 / fp32_peak})
            if roof is None or kb["avg_ms"] > roof["avg_launch_ms"]:
                # The dominant kernel of the step: the actor head's backward over the live pairs, two fp32 FMA products per
                # pair.  It is bound by the fp32 FMA pipe -- neither of the contract's "hbm" | "tensor" --
                # so the peak is 148 SMs x 128 FMA lanes x 2 x the SM clock sampled during the timed region.
                roof = {"kernel": "cells_bwd", "bound": "fma", "achieved": kb["tflops"], "peak": fp32_peak, "unit": "TFLOP/s",
                        "frac": kb["frac_of_fp32_peak"], "traffic": None,
                        "peak_source": "148 SMs x 128 fp32 FMA lanes x 2 flop x sampled SM clock (nvidia-smi, timed region)",
                        "alg_flops_per_launch": useful, "live_pairs": pairs, "avg_launch_ms": kb["avg_ms"],
                        "kernel_share_of_step": kb["avg_ms"] / ms_step,
                        "note": "fp32 FMA-pipe bound (ncu: FMA pipe 48 % active, profiles/r02_ncu_kernels.csv); the dominant "
                                "tensor-pipe and HBM-bound kernels are under info.roofline_tensor / info.roofline_hbm"}
                try:
                    roof["traffic"] = json.load(open(os.path.join(ROOT, "profiles", "traffic.json"))).get("cells_bwd", {}).get(args.workload)
                except Exception:
                    pass
    if "actor_head_fwd" in kernels and "issued_tf32_tflops" in kernels["actor_head_fwd"] and \
            (roof is None or kernels["actor_head_fwd"]["avg_ms"] > roof["avg_launch_ms"]):
        kf = kernels["actor_head_fwd"]
        # The dominant kernel of the path is now the fused actor Linear + policy head, which is bound by the tensor
        # pipe.  Its arithmetic is DEFINED as three TF32 products per fp32 product: algorithmic flops
        # per launch = 3 * 2 * frames * (80 * cells) * 256, against the TF32 dense peak = half the measured bf16 rate.
        roof = {"kernel": "actor_head_fwd", "bound": "tensor", "achieved": kf["issued_tf32_tflops"],
                "peak": kf["tf32_peak_tflops"], "unit": "TFLOP/s", "frac": kf["frac_of_tf32_peak_issued"], "traffic": None,
                "peak_source": "MEASURED_PEAKS.json bf16_tflops / 2 (TF32 runs at half the bf16 rate)",
                "alg_flops_per_launch": 3.0 * 2.0 * ops.launch_notes["actor_head_fwd"]["m"] *
                                        (80 * ops.launch_notes["actor_head_fwd"]["cells"]) * ops.launch_notes["actor_head_fwd"]["k"],
                "fp32_equivalent_tflops": kf["algorithmic_tflops_fp32"],
                "frac_of_bf16_peak": kf["issued_tf32_tflops"] / (2 * kf["tf32_peak_tflops"]),
                "avg_launch_ms": kf["avg_ms"], "kernel_share_of_step": kf["avg_ms"] / ms_step}
        try:        # DRAM bytes of one launch from the ncu --set full capture (profiles/): operands are L2-resident
            roof["traffic"] = json.load(open(os.path.join(ROOT, "profiles", "traffic.json"))).get("actor_head_fwd", {}).get(args.workload)
        except Exception:
            pass

    # ---- e2e: host batch through the public call ---------------------------------------
    e2e = None
    e2e_other = None
    if not args.no_e2e:
        from microbeast_b200 import buffers as mbuf
        wire = {k: v.cpu() for k, v in dev_batch.items()}
        wire["obs"] = batch["obs"].cpu()
        if args.wire == "compact":
            wire = mbuf.pack_batch(wire)        # rollout-store wire format: bit-packed obs / masks, u8 actions
        host = {k: torch.empty(v.shape, dtype=v.dtype, pin_memory=True).copy_(v) for k, v in wire.items()}
        h2d = sum(v.numel() * v.element_size() for v in host.values())
        e2e_steps = max(3, min(args.steps, 10))
        # serial: PPO_learn(host batch) copies, then computes.  pipelined (the headline): learner.BatchPrefetcher stages
        # batch k+1 on a side stream while step k runs -- every step still pays its own host->device copy and the
        # device->host read of its four losses inside the timed region, they just overlap the previous step's kernels
        ms_serial, _, _, _ = timed(lambda: step(host, True), max(3, e2e_steps // 2), 3)
        from microbeast_b200.learner import BatchPrefetcher
        pf = BatchPrefetcher(dev)
        state = {"next": pf.submit(host)}

        def submit_next():
            state["next"] = pf.submit(host)
            return state["next"]

        def e2e_step():
            # step k is enqueued, batch k+1 submitted behind it (its copies overlap step k's kernels; the slot it
            # overwrites was released by step k-1), then step k's four losses are read on the host
            # measured (scripts/e2e_placement.py, N = 1 / 8): submitting once the whole step is enqueued 6.86 / 8.62 ms,
            # from PPO_learn's prefetch hook after the forward pass 7.01 / 8.85, before the step 7.69 / 11.0 -- right
            # after a loss read the forward pass is host-bound, so the submit's host time is cheapest at the end
            out = step(state["next"], False)
            submit_next()
            return out.tolist()
        ms_e2e, _, _, _ = timed(e2e_step, e2e_steps, 3)
        torch.cuda.synchronize()

        def e2e_step_deferred():
            # for the record, not the headline: the same loop with step k's losses read while step k+1 runs, so the
            # device queue never drains (a logger that is one update late); every step still pays its H2D and its D2H
            out = step(state["next"], False)
            submit_next()
            slot = state["turn"] = (state.get("turn", 0) + 1) % 2
            pinned4[slot].copy_(out, non_blocking=True)       # D2H of this step's four losses, stream-ordered after it
            ev = torch.cuda.Event()
            ev.record()
            prev, state["pending"] = state.get("pending"), (pinned4[slot], ev)
            if prev is None:
                return None
            prev[1].synchronize()                             # waits for step k-1 only; step k keeps the device busy
            return prev[0].tolist()
        pinned4 = [torch.empty(4, dtype=torch.float32, pin_memory=True) for _ in range(2)]
        ms_def, _, _, _ = timed(e2e_step_deferred, e2e_steps, 3)
        torch.cuda.synchronize()
        del state, pf
        e2e = {"value": frames / (ms_e2e * 1e-3), "unit": UNIT, "h2d_bytes_per_step": h2d * world,
               "d2h_bytes_per_step": 16 * world, "ms_per_step": ms_e2e, "steps": e2e_steps,
               "call": "out = PPO_learn(model, model, name, cur, adam, rows, T, 1, sync=False); cur = "
                       "BatchPrefetcher.submit(pinned_host_batch); losses = out.tolist()  (microbeast_b200.learner)",
               "deferred_read": {"value": frames / (ms_def * 1e-3), "ms_per_step": ms_def,
                                 "note": "losses of step k read on the host while step k+1 runs; not the headline"},
               "serial": {"value": frames / (ms_serial * 1e-3), "ms_per_step": ms_serial,
                          "call": "PPO_learn(model, model, name, pinned_host_batch, adam, rows, T, 1): copy, then compute"},
               "wire_format": ("obs one int32 per cell (27 binary planes), action_mask bit-packed, action u8 -- widened "
                               "on the device by mb_unpack_obs / mb_unpack_rollout; reward/logprobs f32, done bool") if args.wire == "compact"
               else "obs u8, action_mask u8, action i64, reward/logprobs f32, done bool"}
        del host
        # the other wire format, for the record: compact is an opt-in of create_buffers
        other = {k: v.cpu() for k, v in dev_batch.items()}
        other["obs"] = batch["obs"].cpu()
        if args.wire != "compact":
            other = mbuf.pack_batch(other)
        host2 = {k: torch.empty(v.shape, dtype=v.dtype, pin_memory=True).copy_(v) for k, v in other.items()}
        ms_o, _, _, _ = timed(lambda: step(host2, True), 3, 2)
        e2e_other = {"wire": "plain" if args.wire == "compact" else "compact", "value": frames / (ms_o * 1e-3), "unit": UNIT,
                     "ms_per_step": ms_o, "h2d_bytes_per_step": sum(v.numel() * v.element_size() for v in host2.values()) * world}
        del host2, other

    # ---- informational only: the same step with PyTorch's TF32 convolutions / matmuls allowed ----
    # (the encoder is PyTorch's; `value` above keeps it in strict fp32 so that "dtype": "f32" holds)
    info = None
    if args.tf32_info:
        torch.backends.cudnn.allow_tf32 = True
        torch.backends.cuda.matmul.allow_tf32 = True
        ms_tf32, _, _, _ = timed(lambda: step(dev_batch, False), max(3, min(args.steps, 5)), 3)
        torch.backends.cudnn.allow_tf32 = False
        torch.backends.cuda.matmul.allow_tf32 = False
        info = {"encoder_tf32": {"value": frames / (ms_tf32 * 1e-3), "unit": UNIT, "ms_per_step": ms_tf32,
                                 "note": "not the headline: TF32 tensor-core math in the PyTorch encoder only; "
                                         "the policy-head and V-trace kernels are unchanged (fp32)"}}

    if not args.sparse_head:
        # informational: the same step with the sparse actor head (logits only for active cells)
        model.sparse_head = True
        ms_sp, _, _, _ = timed(lambda: step(dev_batch, False), max(3, min(args.steps, 5)), 3)
        model.sparse_head = False
        info = dict(info or {})
        info["sparse_head"] = {"value": frames / (ms_sp * 1e-3), "unit": UNIT, "ms_per_step": ms_sp,
                               "note": "opt-in Agent.sparse_head: fp32, same losses and gradients; policy_logits "
                                       "is not produced for fully masked cells"}

    if not args.sparse_head and not args.no_graph_info:
        # informational: the same dense step replayed from a CUDA graph (one launch per step)
        from microbeast_b200.graphed import GraphedLearner
        try:
            gopt = torch.optim.Adam(model.parameters(), lr=2.5e-4, eps=1e-5, fused=True, capturable=True)
            gl = GraphedLearner(model, gopt, rows, T, 1, reducer=reducer)
            ms_g, _, _, _ = timed(lambda: gl.step(dev_batch, sync=False), max(3, min(args.steps, 10)), 3)
            info = dict(info or {})
            info["graphed"] = {"value": frames / (ms_g * 1e-3), "unit": UNIT, "ms_per_step": ms_g,
                               "kernels_per_step": gl.launches_per_step,
                               "note": "microbeast_b200.graphed.GraphedLearner: the same kernels replayed from "
                                       "a CUDA graph, batch resident in HBM"}
            del gl, gopt
        except Exception as exc:        # report, do not hide: the headline above does not depend on it
            info = dict(info or {})
            info["graphed"] = {"error": f"{type(exc).__name__}: {exc}"[:300]}

    info = dict(info or {})
    if "actor_head_list_fwd" in kernels and "actor_head_fwd" not in kernels:
        # the dense fused actor head (what Agent.get_action(policy_logits=True) runs): timed in a pass of its own
        from microbeast_b200 import model as mmodel
        mmodel.LIST_ACTOR_HEAD = False
        try:
            for _ in range(2):
                step(dev_batch, False)
            ops.start_kernel_timing(only=("actor_head_fwd",))
            for _ in range(3):
                step(dev_batch, False)
            ts = ops.stop_kernel_timing().get("actor_head_fwd", [])
        finally:
            mmodel.LIST_ACTOR_HEAD = True
        if ts:
            sec = sum(ts) / len(ts) * 1e-3
            issued = 3.0 * 2.0 * n * (80 * cells) * 256
            info["roofline_tensor"] = {"kernel": "actor_head_fwd", "bound": "tensor", "achieved": issued / sec / 1e12,
                                       "peak": bf16_peak / 2, "unit": "TFLOP/s", "frac": issued / sec / 1e12 / (bf16_peak / 2),
                                       "avg_launch_ms": sec * 1e3, "alg_flops_per_launch": issued,
                                       "peak_source": "MEASURED_PEAKS.json bf16_tflops / 2 (TF32 runs at half the bf16 rate)",
                                       "note": "dense fused actor Linear + policy head over every (frame, cell) pair; the learner "
                                               "step runs the list form instead (kernels.actor_head_list_fwd), which issues "
                                               "1/16 of these flops; timed in a separate pass with model.LIST_ACTOR_HEAD = False",
                                       "list_form": {k: kernels["actor_head_list_fwd"].get(k) for k in
                                                     ("avg_ms", "issued_tf32_tflops", "frac_of_tf32_peak_issued", "tiles", "live_pairs")}}
    if world > 1:
        info["host_binding"] = {"cores_per_rank": len(bound_cores) if bound_cores else None,
                                "note": "each rank restricted to the host cores NVML lists next to its GPU "
                                        "(parallel.bind_to_gpu_numa); None = NVML / sched_setaffinity unavailable"}
    if e2e_other is not None:
        info["e2e_other_wire"] = e2e_other
    if roof_hbm is not None and roof_hbm is not roof:
        info["roofline_hbm"] = roof_hbm
    if rank == 0 and not args.no_e2e:
        # what the compact store costs the ACTOR side: the slot write of one env step of the
        # reference's 6 envs per actor (microbeast.py:73-100) packs the mask to bits and the observation planes to one
        # int32 per cell on the host; no other number in this line contains that cost
        from microbeast_b200 import buffers as mbuf
        g6 = torch.Generator().manual_seed(0)
        m6 = (torch.rand(6, 78 * cells, generator=g6) < 0.06).to(torch.uint8)
        o6 = (torch.rand(6, h, w, 27, generator=g6) < 0.2).to(torch.uint8)
        for fn, arg in ((mbuf.pack_mask, m6), (mbuf.pack_obs, o6)):
            for _ in range(20):
                fn(arg)
        t0 = time.perf_counter()
        for _ in range(200):
            mbuf.pack_mask(m6)
        t1 = time.perf_counter()
        for _ in range(200):
            mbuf.pack_obs(o6)
        t2 = time.perf_counter()
        info["host_packing"] = {"mask_us_per_slot_write": (t1 - t0) / 200 * 1e6, "obs_us_per_slot_write": (t2 - t1) / 200 * 1e6,
                                "envs_per_write": 6, "note": "numpy packing an actor pays per env step with "
                                "create_buffers(compact=True); one host thread, outside every timed region above"}

    # ---- acting mode: Agent.get_action (encoder + mb_head_sample) latency at small frame counts ------------------
    if rank == 0 and not args.no_graph_info:
        acting = {}
        for n_act in (6, 60, 600):
            env_out = dict(obs=batch["obs"][:n_act].float().view((1, 1, n_act) + tuple(batch["obs"].shape[1:])),
                           action_mask=batch["action_mask"][:n_act].view(1, n_act, -1))
            u = torch.rand((n_act, 7 * cells), device=dev)
            with torch.no_grad():
                for _ in range(5):
                    model.get_action(env_out, uniforms=u)
                torch.cuda.synchronize()
                t0 = time.perf_counter()
                for _ in range(20):
                    out_a, _ = model.get_action(env_out, uniforms=u)
                    out_a["action"].cpu()                    # an actor needs the actions on the host (microbeast.py:91)
                wall = (time.perf_counter() - t0) / 20
                ops.start_kernel_timing()
                for _ in range(10):
                    model.get_action(env_out, uniforms=u)
                kt = ops.stop_kernel_timing()
            acting[str(n_act)] = {"get_action_wall_us": wall * 1e6,
                                  "head_sample_kernel_us": 1e3 * sum(kt.get("head_sample", [0])) / max(len(kt.get("head_sample", [1])), 1)}
        info["acting_mode"] = {"frames": acting, "note": "Agent.get_action(learning=False) end to end on the host clock "
                               "(encoder + actor Linear + mb_head_sample + action D2H) and the sampling kernel alone"}

    # ---- configs[2] as written: 1024 rows in total, split over the ranks (strong scaling) ---------------------------
    if args.workload == "c2" and not args.no_graph_info:
        try:
            rows3 = 1024 // world
            b3 = synth.make_rollout_batch(h, w, T, rows3, seed=4321 + rank, device=dev, preset=args.preset,
                                          obs_dtype=torch.uint8, full_keys=False)
            with torch.no_grad():
                o3, _ = model.get_action(b3, learning=True, policy_logits=False)
                b3["logprobs"] = o3["logprobs"] + 0.3 * torch.randn_like(o3["logprobs"])
                del o3
            d3 = {k: b3[k] for k in LEARNER_KEYS}
            ms3, _, _, _ = timed(lambda: PPO_learn(model, model, "bench", d3, opt, rows3, T, 1, reducer=reducer, sync=False),
                                 max(3, min(args.steps, 5)), 3)
            info["c3_strong"] = {"rows_total": 1024, "rows_per_gpu": rows3, "n_gpus": world, "ms_per_step": ms3,
                                 "value": 1024 * T / (ms3 * 1e-3), "unit": UNIT,
                                 "note": "BASELINE configs[2]: 16x16, T=80, B=1024 rows in total over the ranks, batch resident"}
            del b3, d3
        except Exception as exc:
            info["c3_strong"] = {"error": f"{type(exc).__name__}: {exc}"[:300]}

    # ---- cpu baseline (rank 0, N=1) ------------------------------------------------------
    cpu = None
    if rank == 0 and world == 1 and not args.no_cpu_baseline:
        crow = args.cpu_rows
        fps, sec, cores, kind = time_cpu_reference(h, w, T, crow, 2, 1, args.preset)
        cpu = {"value": fps, "unit": UNIT, "cores": cores, "kind": kind,
               "sample": f"{crow} of {rows} rows ({crow * (T + 1)} frames) per step, 2 timed steps after 1 "
                         f"warm-up, full reference learner step on torch CPU ({sec:.2f} s/step)"}
        # the reference's nominal setting is OMP_NUM_THREADS=1 (microbeast.py:16): one timed step on a quarter sample
        fps1, sec1, _, _ = time_cpu_reference(h, w, T, max(crow // 4, 1), 1, 1, args.preset, threads=1)
        cpu["single_thread"] = {"value": fps1, "unit": UNIT, "cores": 1,
                                "sample": f"{max(crow // 4, 1)} rows per step, 1 timed step ({sec1:.2f} s/step)"}
        torch.set_num_threads(os.cpu_count() or 1)

    if rank == 0:
        line = {
            "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps,
            "warmup": max(args.warmup, 3), "ms_per_step": ms_step, "higher_is_better": True,
            "scaling": "strong" if args.workload == "c3" else "weak",
            "vs_baseline": None, "dtype": "f32", "data": "synthetic", "config": workload_config(args),
            "clocks": clocks, "e2e": e2e, "gpu_launches": launches, "roofline": roof, "cpu_baseline": cpu,
            "parity": parity,
            "kernels": kernels, "info": info,
        }
        emit(line, args)
    if world > 1:
        dist.destroy_process_group()


# ------------------------------------------------------------------------------------------
# kernel-only sweeps (profiles/): BASELINE config 5 and the head kernels at configs 1-4
# ------------------------------------------------------------------------------------------
def run_sweep(args):
    import torch
    from microbeast_b200 import ops, synth
    dev = torch.device("cuda", 0)
    peak, peak_src = hbm_peak()
    g = torch.Generator(device=dev).manual_seed(1)
    flush = torch.empty(512 << 20, dtype=torch.uint8, device=dev)        # > 126 MB L2

    def bench(fn, iters):
        ts = []
        for i in range(3 + iters):
            flush.zero_()
            s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            s.record(); fn(); e.record()
            torch.cuda.synchronize()
            if i >= 3:
                ts.append(s.elapsed_time(e))
        ts.sort()
        return ts[len(ts) // 2], ts[max(0, len(ts) // 10)], ts[min(len(ts) - 1, (9 * len(ts)) // 10)]

    rows_out = []
    if args.sweep == "vtrace":
        for T in (20, 80, 256, 512):
            for R in (256, 1024, 4096, 16384, 65536):
                t = synth.make_vtrace_inputs(R, T, gen=g, device=dev)
                fn = lambda: ops.vtrace_loss(t["lp_old"], t["lp_new"], t["values"], t["bootstrap"], t["rewards"],
                                             t["done"], t["entropy"])
                med, p10, p90 = bench(fn, 20)
                nbytes = 37 * R * T + 4 * R
                gbs = nbytes / (med * 1e-3) / 1e9
                cpu_ms = None
                if R * T <= (1 << 21):
                    from oracle import vtrace as ovt
                    c = {k: v.cpu() for k, v in t.items()}
                    t0 = time.perf_counter()
                    ovt.vtrace_losses(c["lp_old"], c["lp_new"], c["values"], c["bootstrap"], c["rewards"], c["done"],
                                      c["entropy"])
                    cpu_ms = (time.perf_counter() - t0) * 1e3
                tt = {k: (v.t().contiguous() if v.dim() == 2 else v) for k, v in t.items()}
                fn_tm = lambda: ops.vtrace_loss(tt["lp_old"], tt["lp_new"], tt["values"], tt["bootstrap"], tt["rewards"],
                                                tt["done"], tt["entropy"], time_major=True)
                med_tm, p10_tm, p90_tm = bench(fn_tm, 20)
                gbs_tm = nbytes / (med_tm * 1e-3) / 1e9
                rows_out.append({"kernel": "vtrace_loss_tmajor", "T": T, "rows": R, "elements": R * T,
                                 "median_us": med_tm * 1e3, "p10_us": p10_tm * 1e3, "p90_us": p90_tm * 1e3,
                                 "alg_bytes": nbytes, "gbs": gbs_tm, "frac_of_peak": gbs_tm / peak,
                                 "regime": "hbm" if R * T >= 8_000_000 else "latency-bound (L2 resident / launch)"})
                del tt
                rows_out.append({"kernel": "vtrace_loss", "T": T, "rows": R, "elements": R * T, "median_us": med * 1e3,
                                 "cpu_reference_ms": cpu_ms,
                                 "p10_us": p10 * 1e3, "p90_us": p90 * 1e3, "alg_bytes": nbytes, "gbs": gbs,
                                 "frac_of_peak": gbs / peak,
                                 "regime": "hbm" if R * T >= 8_000_000 else "latency-bound (L2 resident / launch)"})
                del t
    elif args.sweep == "variants":
        import microbeast_b200
        h, w, T, rows = WORKLOADS["c2"]
        cells, n = h * w, rows * (T + 1)
        logits = synth.make_logits(n, cells, gen=g, device=dev)
        mask = synth.make_masks(n, cells, args.preset, gen=g, device=dev)
        act = synth.make_valid_actions(mask, gen=g)
        u = synth.make_uniforms(n, cells, gen=g, device=dev)
        gl = torch.randn(n, generator=g, device=dev)
        out = torch.empty_like(logits)
        names = {0: "64x3", 1: "64x2", 2: "64x4", 3: "32x3", 4: "32x5", 5: "128x3", 6: "128x2", 7: "32x2"}
        for variant, vname in names.items():
            for per_sm in (0,):
                microbeast_b200.set_tuning("head_variant", variant)
                microbeast_b200.set_tuning("head_ctas_per_sm", per_sm)
                for kname, fn, per_cell in (
                        ("head_eval_fwd", lambda: ops.head_eval_fwd(logits, mask, act), 446),
                        ("head_eval_bwd", lambda: ops.head_eval_bwd(logits, mask, act, gl, gl, out=out), 758),
                        ("head_sample", lambda: ops.head_sample(logits, mask, u), 474)):
                    med, p10, p90 = bench(fn, 6)
                    nbytes = per_cell * cells * n + 8 * n
                    gbs = nbytes / (med * 1e-3) / 1e9
                    rows_out.append({"kernel": kname, "variant": vname, "ctas_per_sm": per_sm or "max",
                                     "median_us": med * 1e3, "gbs": gbs, "frac_of_peak": gbs / peak})
        microbeast_b200.set_tuning("head_variant", -1)
        microbeast_b200.set_tuning("head_ctas_per_sm", 0)
        vnames = {0: "2x32", 1: "3x32", 2: "4x32", 3: "3x16", 4: "4x16", 5: "5x16", 6: "6x16", 7: "8x16"}
        for variant, vname in vnames.items():
            microbeast_b200.set_tuning("vtrace_variant", variant)
            for T, R in ((512, 65536), (256, 65536), (80, 65536), (512, 4096), (80, 256)):
                t = synth.make_vtrace_inputs(R, T, gen=g, device=dev)
                fn = lambda: ops.vtrace_loss(t["lp_old"], t["lp_new"], t["values"], t["bootstrap"], t["rewards"],
                                             t["done"], t["entropy"])
                med, p10, p90 = bench(fn, 10)
                nbytes = 37 * R * T + 4 * R
                gbs = nbytes / (med * 1e-3) / 1e9
                rows_out.append({"kernel": "vtrace_loss", "variant": vname, "T": T, "rows": R,
                                 "median_us": med * 1e3, "gbs": gbs, "frac_of_peak": gbs / peak})
                del t
        microbeast_b200.set_tuning("vtrace_variant", -1)
    else:
        for name, (h, w, T, rows) in WORKLOADS.items():
            cells, n = h * w, rows * (T + 1)
            logits = synth.make_logits(n, cells, gen=g, device=dev)
            mask = synth.make_masks(n, cells, args.preset, gen=g, device=dev)
            act = synth.make_valid_actions(mask, gen=g)
            u = synth.make_uniforms(n, cells, gen=g, device=dev)
            gl = torch.randn(n, generator=g, device=dev)
            out = torch.empty_like(logits)
            for kname, fn, per_cell, per_frame in (
                    ("head_eval_fwd", lambda: ops.head_eval_fwd(logits, mask, act), 446, 8),
                    ("head_eval_bwd", lambda: ops.head_eval_bwd(logits, mask, act, gl, gl, out=out), 758, 8),
                    ("head_sample", lambda: ops.head_sample(logits, mask, u), 474, 8)):
                med, p10, p90 = bench(fn, 10)
                nbytes = per_cell * cells * n + per_frame * n
                gbs = nbytes / (med * 1e-3) / 1e9
                rows_out.append({"kernel": kname, "workload": name, "frames": n, "cells": cells,
                                 "median_us": med * 1e3, "p10_us": p10 * 1e3, "p90_us": p90 * 1e3,
                                 "alg_bytes": nbytes, "gbs": gbs, "frac_of_peak": gbs / peak})
            del logits, mask, act, u, out
    for r in rows_out:
        r["peak_gbs"] = peak
        r["peak_source"] = peak_src
        emit(r, args)


def main():
    args = parse_args()
    if args.sweep:
        run_sweep(args)
    elif args.impl == "reference":
        run_reference(args)
    else:
        run_ours(args)


if __name__ == "__main__":
    main()
