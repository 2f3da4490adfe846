"""Throughput of the reference-order batch (plan_batch_reference: one warp per query) on the README
map (51 x 31, README.md:62-76) and the 3d_example map size, random free start / goal pairs.
    python tools/bench_reference_batch.py [nq]"""
import json
import os
import sys
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch

import python_motion_planning_b200 as pmp

nq = int(sys.argv[1]) if len(sys.argv) > 1 else 2048
rng = np.random.default_rng(0)
env = pmp.Grid(51, 31)
obs = env.obstacles
for i in range(10, 21):
    obs.add((i, 15))
for i in range(15):
    obs.add((20, i))
for i in range(15, 30):
    obs.add((30, i))
for i in range(16):
    obs.add((40, i))
env.update(obs)
env3 = pmp.Grid3D(26, 20, 16)
out = {"nq": nq}
for name, e in (("readme_51x31", env), ("grid3d_26x20x16", env3)):
    free = np.argwhere(~e.occupancy())
    s = free[rng.integers(0, len(free), nq)]
    g = free[rng.integers(0, len(free), nq)]
    for algo in ("a_star", "dijkstra", "gbfs", "theta_star", "jps"):
        pmp.plan_batch_reference(e, s[:64], g[:64], algo=algo, want_paths=False)
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        res = pmp.plan_batch_reference(e, s, g, algo=algo, want_paths=False)
        torch.cuda.synchronize()
        dt = time.perf_counter() - t0
        out[f"{name}/{algo}"] = {"ms": round(dt * 1e3, 2), "queries_per_s": round(nq / dt, 1),
                                 "mean_visited": round(float(res.n_expand.mean()), 1)}
print(json.dumps(out))
