"""Row f2 numbers: line-of-sight throughput and the cost of shortcutting the c4 path.

    python tools/bench_anyangle.py [--size 512] [--pairs 1000000]
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
    ap.add_argument("--pairs", type=int, default=1000000)
    args = ap.parse_args()
    import torch
    import python_motion_planning_b200 as pmp
    from python_motion_planning_b200 import _lib, workloads, anyangle, engine
    n = args.size
    occ = workloads.boxes3d(n, seed=0)
    env = pmp.Grid3D(n, n, n, walls=False)
    env.update(occ)
    dg = env.device_grid()
    rng = np.random.default_rng(0)
    a = rng.integers(1, n - 1, size=(args.pairs, 3)).astype(np.int32)
    b = np.clip(a + rng.integers(-64, 65, size=(args.pairs, 3)), 1, n - 2).astype(np.int32)
    ad, bd = torch.from_numpy(a).cuda(), torch.from_numpy(b).cuda()
    out = torch.empty(args.pairs, dtype=torch.uint8, device="cuda")
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    for _ in range(3):
        _lib.check(dg.L.pmp_line_of_sight(dg.handle, args.pairs, ad.data_ptr(), bd.data_ptr(), out.data_ptr(), dg.stream()), "los")
    torch.cuda.synchronize()
    e0.record()
    for _ in range(5):
        _lib.check(dg.L.pmp_line_of_sight(dg.handle, args.pairs, ad.data_ptr(), bd.data_ptr(), out.data_ptr(), dg.stream()), "los")
    e1.record()
    torch.cuda.synchronize()
    los_ms = e0.elapsed_time(e1) / 5
    vis = float(out.float().mean())
    steps = float(np.abs(b - a).max(axis=1).mean())
    th = pmp.ThetaStar3D((1, 1, 1), (n - 2, n - 2, n - 2), env)
    for _ in range(2):
        cost, path, _ = th.plan()
    ws = th.last_field
    torch.cuda.synchronize()
    e0.record()
    for _ in range(10):
        anyangle.shortcut(ws)
    e1.record()
    torch.cuda.synchronize()
    sc_ms = e0.elapsed_time(e1) / 10
    print(json.dumps({"workload": f"f2 on the {n}^3 boxes map", "los_pairs": args.pairs, "los_ms": los_ms,
                      "los_Mpairs_per_s": args.pairs / los_ms / 1e3, "mean_ray_cells": steps, "visible_frac": vis,
                      "grid_path_cells": int(ws.path_len[0].item()), "grid_cost": th.grid_cost, "any_angle_cost": cost,
                      "waypoints": len(path), "shortcut_ms": sc_ms}))


if __name__ == "__main__":
    main()
