"""Phase split of the unit-weight tile solver (needs a -DPMP_BFSSTAT build of the library:
PMP_B200_LIB=.../libpmp_b200_bfs.so python tools/bfsstat.py)."""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
import python_motion_planning_b200 as pmp
from python_motion_planning_b200 import workloads, engine, _lib
n = int(sys.argv[1]) if len(sys.argv) > 1 else 768
blk, shape = workloads.replicated_boxes3d(base=128, reps=max(1, n // 128))
occ = workloads.replicated_slab(blk, (n, n, n), 0, n)
env = pmp.Grid3D(n, n, n, walls=False); env.update(occ)
env.motions = [m for m in env.motions if sum(abs(v) for v in m.current) == 1]
dg = env.device_grid()
from python_motion_planning_b200.environment import motion_table
dg.set_motions(*motion_table(env, 3))
ws = engine.get_workspace(dg, 1, _lib.W_U32)
engine.upload_queries(ws, [(1, 1, 1)], None, 3)
for _ in range(3):
    engine.sssp(ws, _lib.DIJKSTRA, 0, full_field=True, have_goals=False)
torch.cuda.synchronize()
st = [int(v) for v in ws.stats.cpu()]
mhz = 1965.0
act = st[1]
print("kernel_ms", st[8] / 1e6, "activations", act, "levels with a frontier", st[12])
print("per activation us: load %.2f settle %.2f publish %.2f" % (st[5] / act / mhz, st[6] / act / mhz, st[7] / act / mhz))
print("inside settle, per activation us: reduce+clear+bucket %.2f sweep %.2f owners %.2f" % (st[13] / act / mhz, st[14] / act / mhz, st[15] / act / mhz))
