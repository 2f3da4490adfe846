"""Latency of incremental replanning vs planning from scratch (row f1), on the c4 workload
(512^3 26-connected voxel grid, random boxes, A* euclidean) or a smaller copy (--size).

    python tools/bench_replan.py [--size 512] [--changes 20]

Prints one JSON line: first plan, then per change (a cell of the current path gets blocked, every
third change frees a previously blocked one) the device time of pmp_replan_update + path
extraction, and the time of a fresh AStar3D plan on the same map for comparison.
"""
import argparse
import json
import os
import sys

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--size", type=int, default=512)
    ap.add_argument("--changes", type=int, default=20)
    ap.add_argument("--block", type=int, default=1, help="cells blocked per change (a cube of this edge)")
    args = ap.parse_args()
    import torch
    import python_motion_planning_b200 as pmp
    from python_motion_planning_b200 import workloads
    n = args.size
    occ = workloads.boxes3d(n, seed=0)
    env = pmp.Grid3D(n, n, n, walls=False)
    env.update(occ)
    start, goal = (1, 1, 1), (n - 2, n - 2, n - 2)

    def timed(fn):
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        out = fn()
        e1.record()
        torch.cuda.synchronize()
        return out, e0.elapsed_time(e1)

    fresh = pmp.AStar3D(start, goal, env)
    for _ in range(2):
        fresh.plan()
    (c_fresh, _, _), ms_fresh = timed(fresh.plan)
    lpa = pmp.LPAStar3D(start, goal, env)
    (cost, path, _), ms_first = timed(lpa.plan)
    rng = np.random.default_rng(0)
    rows = []
    blocked_hist = []
    for i in range(args.changes):
        if i % 3 == 2 and blocked_hist:
            cells, flags = blocked_hist.pop(0), False
        else:
            c = path[int(rng.integers(len(path) // 4, 3 * len(path) // 4))]
            e = args.block
            cells = [(c[0] + a, c[1] + b, c[2] + d) for a in range(e) for b in range(e) for d in range(e)]
            cells = [p for p in cells if p != start and p != goal and max(p) < n - 1]
            flags = True
            blocked_hist.append(cells)
        (cost, path, _), ms = timed(lambda: lpa.apply_changes(cells, [flags] * len(cells)))
        st = [int(v) for v in lpa.stats]
        rows.append({"blocked": bool(flags), "cells": len(cells), "ms": round(ms, 3), "cost": cost,
                     "reset_cells": lpa.info["reset_cells"], "waves": lpa.info["waves"],
                     "fallback": lpa.info["fallback"], "rounds": st[0], "activations": st[1]})
    occ_now = env.occupancy()
    env2 = pmp.Grid3D(n, n, n, walls=False)
    env2.update(occ_now)
    chk = pmp.AStar3D(start, goal, env2)
    chk.plan()
    (c_chk, _, _), ms_chk = timed(chk.plan)
    ms = [r["ms"] for r in rows]
    print(json.dumps({"workload": f"f1: LPAStar3D on {n}^3 26-connected boxes, {args.changes} changes of {args.block}^3 cells",
                      "first_plan_ms": round(ms_first, 3), "astar_from_scratch_ms": round(ms_fresh, 3),
                      "replan_ms_median": float(np.median(ms)), "replan_ms_max": float(np.max(ms)),
                      "replan_ms_min": float(np.min(ms)), "final_cost": cost, "from_scratch_cost_on_final_map": c_chk,
                      "from_scratch_ms_on_final_map": round(ms_chk, 3), "changes": rows}))
    assert abs(cost - c_chk) <= 1e-9 * c_chk, (cost, c_chk)


if __name__ == "__main__":
    main()
