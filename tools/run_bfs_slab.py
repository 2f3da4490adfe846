"""One rank of the z-slab field solver with device-side halo exchange (pmp_bfs_slab_*).

    torchrun --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 --master-port P \
        tools/run_bfs_slab.py --size 768 [--steps 3] [--verify] [--reached R]

The grid is BASELINE.json config 5's: the 128^3 box block replicated to n^3, 6-connected unit
weights, source (1,1,1).  Rank 0 prints one JSON line: device time per search (max over ranks),
GTEPS, and with --verify the global certificate (violations, reached cells, pending flags).
"""
import argparse
import json
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--size", dest="n", type=int, default=768)
    ap.add_argument("--steps", type=int, default=3)
    ap.add_argument("--verify", action="store_true")
    ap.add_argument("--reached", type=int, default=0, help="expected number of reached cells (0: do not check)")
    args = ap.parse_args()
    import torch
    import torch.distributed as dist
    from python_motion_planning_b200 import slab, workloads
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    d = None
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
        d = dist
    n = args.n
    blk, _ = workloads.replicated_boxes3d(base=128, reps=max(1, n // 128))
    solver = slab.BfsSlabSolver(lambda z0, z1: workloads.replicated_packed(blk, (n, n, n), z0, z1), (n, n, n), rank, world, d, packed=True)
    src = (1, 1, 1)
    solver.solve(src)     # warm-up
    solver.check()
    ms = []
    for _ in range(args.steps):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        if d is not None:
            d.barrier()
        torch.cuda.synchronize()
        e0.record()
        solver.solve(src)
        e1.record()
        torch.cuda.synchronize()
        ms.append(e0.elapsed_time(e1))
    solver.check()
    t = torch.tensor([sum(ms) / len(ms)], dtype=torch.float64, device="cuda")
    st = solver.slab.stats.clone()
    per_rank = None
    if d is not None:        # every rank's counters of the last search, for the per-slab breakdown
        mine = torch.from_numpy(solver.slab.stats_host().copy()).to("cuda")
        allst = [torch.empty_like(mine) for _ in range(world)]
        d.all_gather(allst, mine)
        per_rank = [{"rank": r, "first_ms": (a[18].item() - a[9].item()) / 1e6, "busy_ms": (a[16].item() - a[9].item()) / 1e6, "activations": a[1].item(), "idle": a[13].item(),
                     "lowered": a[2].item(), "remote_pushes": a[7].item(), "nvlink_bytes_read": a[17].item() * 128}
                    for r, a in enumerate(allst)]
    if d is not None:
        d.all_reduce(t, op=d.ReduceOp.MAX)
        d.all_reduce(st, op=d.ReduceOp.SUM)
    out = {"n": n, "world": world, "ms_per_search": float(t[0]), "steps": args.steps}
    s = st.cpu().tolist()
    out.update(tile_activations=s[1], idle_activations=s[13], cells_lowered=s[2], levels=s[12], requeues=s[5], remote_pushes=s[7])
    ok = True
    if args.verify:
        bad, reached, pend = solver.verify(src)
        out.update(violations=bad, reached=reached, pending=pend,
                   gteps=6.0 * reached / (float(t[0]) * 1e-3) / 1e9)
        ok = bad == 0 and pend == 0 and (args.reached == 0 or reached == args.reached)
    out["ok"] = ok
    if per_rank:
        out["per_rank"] = per_rank
    if rank == 0:
        print(json.dumps(out), flush=True)
    if d is not None:
        d.destroy_process_group()
    return 0 if ok else 1


if __name__ == "__main__":
    sys.exit(main())
