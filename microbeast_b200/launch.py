"""Multi-process launcher for ``train()`` -- reference ``microbeast.py:109-264,271-278``.

    python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 --master-port P \\
        -m microbeast_b200.launch --exp_name runs/exp --env_size 16 --total_steps 200000

One process per GPU.  Every rank runs its own actors and rollout store (``train.train``) on its own GPU and learns
from its own batches; the ranks hold identical model replicas and sum their gradients through
``parallel.GradAllReducer`` (NCCL, the actor-head slice overlapped with the encoder backward), so one update is
the reference's update on a batch of ``world_size * B * n_envs`` rows.  Rank 0 writes ``<exp_name>.csv`` and
``<exp_name>Losses.csv`` exactly as the reference does (``microbeast.py:130-139,233-239``; the logged losses are the
global means); the other ranks write ``<exp_name>.rank<r>*.csv`` so that their episode logs are kept too.
Without ``torchrun`` (no ``RANK`` in the environment) it runs a single process, like ``python microbeast.py``.
The reference takes ``--exp_name`` (``parser.py:10``) and keeps its hyper-parameters as literals in ``train()``
(``microbeast.py:113-121``); they are flags here with those literals as defaults.
"""
from __future__ import annotations

import argparse
import os
import sys


def parse_args(argv=None):
    ap = argparse.ArgumentParser(description=__doc__, formatter_class=argparse.RawDescriptionHelpFormatter)
    ap.add_argument("--exp_name", required=True, help="prefix of the CSV logs (parser.py:10)")
    ap.add_argument("--n_actors", type=int, default=10)          # microbeast.py:113
    ap.add_argument("--n_envs", type=int, default=6)             # microbeast.py:114
    ap.add_argument("--env_size", type=int, default=8)           # microbeast.py:115
    ap.add_argument("--T", type=int, default=64)                 # microbeast.py:116
    ap.add_argument("--B", type=int, default=2)                  # microbeast.py:117
    ap.add_argument("--total_steps", type=int, default=100_000, help="env frames to learn from, summed over the ranks")
    ap.add_argument("--lr", type=float, default=2.5e-4)          # microbeast.py:200
    ap.add_argument("--seed", type=int, default=0)
    ap.add_argument("--backend", default="nccl")
    ap.add_argument("--plain_wire", action="store_true", help="rollout store in the reference's dtypes instead of the compact format")
    return ap.parse_args(argv)


def main(argv=None) -> int:
    args = parse_args(argv)
    import torch
    import torch.distributed as dist

    from .parallel import GradAllReducer
    from .train import train

    distributed = "RANK" in os.environ and int(os.environ.get("WORLD_SIZE", "1")) > 1
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    if not torch.cuda.is_available():
        raise SystemExit("microbeast_b200.launch: no CUDA device; the learner has no CPU fallback")
    torch.cuda.set_device(local)
    device = torch.device("cuda", local)
    if distributed:
        from .parallel import bind_to_gpu_numa
        bind_to_gpu_numa(local)                 # host cores (and first-touch pinned memory) of the GPU's own socket
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group(args.backend, device_id=device if args.backend == "nccl" else None)
    reducer = GradAllReducer() if distributed else None
    exp = args.exp_name if rank == 0 else f"{args.exp_name}.rank{rank}"
    os.makedirs(os.path.dirname(os.path.abspath(exp)), exist_ok=True)
    try:
        rows = train(exp, n_actors=args.n_actors, n_envs=args.n_envs, env_size=args.env_size, T=args.T, B=args.B,
                     total_steps=max(args.total_steps // world, 1), device=str(device), lr=args.lr,
                     seed=args.seed, rank_seed=rank, reducer=reducer, compact=not args.plain_wire, log=(rank == 0))
        if distributed:
            # replicas must still be identical: compare a checksum of the weights across the ranks
            check = torch.tensor([float(rows[-1][4])], device=device, dtype=torch.float64)
            lo, hi = check.clone(), check.clone()
            dist.all_reduce(lo, op=dist.ReduceOp.MIN)
            dist.all_reduce(hi, op=dist.ReduceOp.MAX)
            if rank == 0:
                print(f"microbeast_b200.launch: {len(rows)} updates on {world} ranks; last total loss {rows[-1][4]:.6f} "
                      f"(identical on every rank: {bool(lo.item() == hi.item())})")
        elif rank == 0:
            print(f"microbeast_b200.launch: {len(rows)} updates; last total loss {rows[-1][4]:.6f}")
    finally:
        if distributed:
            dist.destroy_process_group()
    return 0


if __name__ == "__main__":
    sys.exit(main())
