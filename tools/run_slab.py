"""Slab-decomposed cost-to-go field over N GPUs (BASELINE.json config 5).

    python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 \
        --master-port 29511 tools/run_slab.py --size 1536 [--outer-window 64] [--verify]

One rank per GPU; rank p holds planes z in [pZ/N, (p+1)Z/N) plus ghost planes.  Prints one JSON
line (rank 0): time (max over ranks, CUDA events), GTEPS, halo bytes, certificate result.
"""
import argparse
import json
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--size", dest="n", type=int, default=768)
    ap.add_argument("--outer-window", type=float, default=512.0)
    ap.add_argument("--steps", type=int, default=2)
    ap.add_argument("--warmup", type=int, default=1)
    ap.add_argument("--verify", action="store_true")
    args = ap.parse_args()
    import torch
    import torch.distributed as dist
    from python_motion_planning_b200 import slab, workloads
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local_rank)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))
    n = args.n
    blk, shape = workloads.replicated_boxes3d(base=128, reps=max(1, n // 128))
    h0, h1 = slab.ghost_range(n, rank, world)
    occ = workloads.replicated_slab(blk, (n, n, n), h0, h1)
    eng = slab.CudaSlabEngine(slab.build_local_env(occ, "6"), "u32")
    del occ
    drv = slab.SlabDriver(eng, n, rank, world, dist if world > 1 else None, outer_window=args.outer_window)
    src = (1, 1, 1)
    times = []
    for i in range(args.warmup + args.steps):
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        its = drv.solve(src)
        e1.record()
        torch.cuda.synchronize()
        if i >= args.warmup:
            times.append(e0.elapsed_time(e1))
    bad, reached, pend = drv.verify(src) if args.verify else (None, None, None)
    if reached is None:
        _, reached, _ = eng.verify(None, drv.local_z(drv.z0), drv.local_z(drv.z1))
        if world > 1:
            t = torch.tensor([reached], dtype=torch.int64, device="cuda")
            dist.all_reduce(t)
            reached = int(t.item())
    ms = float(np.mean(times))
    t = torch.tensor([ms], dtype=torch.float64, device="cuda")
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    ms = float(t.item())
    if rank == 0:
        print(json.dumps({"workload": f"c5 slab: {n}^3 6-connected unit weights (u32), {world} slab(s)",
                          "n_gpus": world, "ms": ms, "iterations": its, "reached_cells": reached,
                          "gteps": reached * 6 / ms / 1e6, "algorithmic_GBps": reached * 8.125 / ms / 1e6,
                          "halo_bytes_sent_rank0": drv.bytes_sent, "outer_window": args.outer_window,
                          "certificate": None if bad is None else {"violations": bad, "pending_flags": pend}}))
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
