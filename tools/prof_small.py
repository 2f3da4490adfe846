import sys, time, cProfile, pstats
sys.path.insert(0, "/root/repo")
import numpy as np
import python_motion_planning_b200 as pmp
from tests import golden_io as gio
z = gio.load("csv_replay.npz")
dims = tuple(int(v) for v in z["dims"])
occ = gio.unpack_occ(z["occ_bits"][250], dims)
s, g = tuple(int(v) for v in z["starts"][250]), tuple(int(v) for v in z["goals"][250])
obs = set(map(tuple, np.argwhere(occ).tolist()))
def one(cls):
    env = pmp.Grid3D(*occ.shape, walls=False); env.update(obs)
    return cls(start=s, goal=g, env=env).plan()
for cls in (pmp.AStar3D, pmp.GBFS3D):
    for _ in range(20): one(cls)
    t0 = time.time()
    for _ in range(500): one(cls)
    print(cls.__name__, "ms per fresh-env plan", (time.time() - t0) / 500 * 1e3)
env = pmp.Grid3D(*occ.shape, walls=False); env.update(obs)
p = pmp.AStar3D(start=s, goal=g, env=env)
for _ in range(20): p.plan()
t0 = time.time()
for _ in range(500): p.plan()
print("same-env plan ms", (time.time() - t0) / 500 * 1e3)
pr = cProfile.Profile(); pr.enable()
for _ in range(300): one(pmp.AStar3D)
pr.disable()
pstats.Stats(pr).sort_stats("cumulative").print_stats(28)
