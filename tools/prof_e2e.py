"""Where the host time of the c4 end-to-end step goes (cProfile + wall clock per piece)."""
import cProfile, pstats, sys, os, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
import python_motion_planning_b200 as pmp
from python_motion_planning_b200 import workloads
n = 512
occ = workloads.boxes3d(n, seed=0)
host = torch.from_numpy(occ.view(np.uint8)).pin_memory()
env = pmp.Grid3D(n, n, n, walls=False)
def step():
    env.update(host)
    return pmp.AStar3D((1, 1, 1), (n - 2, n - 2, n - 2), env).plan()
for _ in range(3): step()
torch.cuda.synchronize(); t0 = time.perf_counter()
for _ in range(10): step()
torch.cuda.synchronize(); print("e2e ms", (time.perf_counter() - t0) / 10 * 1e3)
# pieces
dg = env.device_grid(); torch.cuda.synchronize()
t0 = time.perf_counter()
for _ in range(10): dg.set_dense(host); torch.cuda.synchronize()
print("set_dense (H2D + pack) ms", (time.perf_counter() - t0) / 10 * 1e3)
p = pmp.AStar3D((1, 1, 1), (n - 2, n - 2, n - 2), env)
env._device_version = env._version
t0 = time.perf_counter()
for _ in range(10): p.plan()
torch.cuda.synchronize(); print("plan() alone ms", (time.perf_counter() - t0) / 10 * 1e3)
pr = cProfile.Profile(); pr.enable()
for _ in range(10): step()
pr.disable(); pstats.Stats(pr).sort_stats("cumulative").print_stats(18)
