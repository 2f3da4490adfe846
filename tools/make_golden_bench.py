"""Known answers at the BENCHMARKED sizes, written by the pinned oracle
(oracle/pmp_oracle.c, itself held to the reference's vectors by tests/test_oracle_golden.py):

  c4   one A* (euclidean, 26-connected) on workloads.boxes3d(n, seed=0), (1,1,1) -> (n-2,n-2,n-2),
       n = 256 and 512: cost, path length, expansions
  c3   the first 64 of the 4096 seeded Dijkstra queries on workloads.random2d(1024): costs

-> tests/golden/bench_sizes.json.  bench.py asserts its check values against this file and
tests/test_gpu_parity.py compares the CUDA path with it.  Run here (CPU): python tools/make_golden_bench.py
"""
import json
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from oracle import oracle   # noqa: E402
from python_motion_planning_b200 import workloads   # noqa: E402

out = {}
for n in (256, 512):
    occ = workloads.boxes3d(n, seed=0)
    t0 = time.time()
    r = oracle.plan(occ, (1, 1, 1), (n - 2, n - 2, n - 2), algo=oracle.ASTAR, want_expand=False)
    out[f"c4_{n}"] = {"cost": float(r.cost), "cost_hex": float(r.cost).hex(), "path_len": len(r.path), "n_expand": int(r.n_expand),
                      "reached": bool(r.reached)}
    print(n, out[f"c4_{n}"], f"{time.time() - t0:.1f} s", flush=True)
occ, starts, goals = workloads.random2d(1024, nq=4096, seed=0)
t0 = time.time()
cost, nexp, plen, flags = oracle.plan_batch(occ, starts[:64], goals[:64], algo=oracle.DIJKSTRA)
out["c3_1024_first64"] = {"cost_hex": [float(c).hex() for c in cost], "n_expand": [int(v) for v in nexp]}
print("c3", np.isfinite(cost).sum(), "reachable of 64", f"{time.time() - t0:.1f} s")
json.dump(out, open(os.path.join(ROOT, "tests", "golden", "bench_sizes.json"), "w"), indent=1)
