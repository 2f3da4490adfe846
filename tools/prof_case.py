"""One small search for ncu (python tools/prof_case.py [field3d|batch2d] [size])."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
import python_motion_planning_b200 as pmp
from python_motion_planning_b200 import workloads

case = sys.argv[1] if len(sys.argv) > 1 else "field3d"
n = int(sys.argv[2]) if len(sys.argv) > 2 else 128
if case == "field3d":
    occ = workloads.boxes3d(n, seed=0)
    env = pmp.Grid3D(n, n, n, walls=False); env.update(occ)
    for _ in range(2):
        ws, st = pmp.cost_field(env, (1, 1, 1), to_host=False)
    torch.cuda.synchronize()
    print("stats", st.cpu().numpy()[:13].tolist())
else:
    occ, starts, goals = workloads.random2d(n, nq=int(sys.argv[3]) if len(sys.argv) > 3 else 32)
    env = pmp.Grid(n, n); env.update(occ)
    for _ in range(2):
        res = pmp.plan_batch(env, starts, goals, want_paths=False)
    print("stats", res.stats[:13].tolist())
