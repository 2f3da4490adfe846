"""Ring pop latency of k_sssp on c4 (needs a -DPMP_POPSTAT build of the library:
    make -C python_motion_planning_b200/csrc EXTRA=-DPMP_POPSTAT   (or nvcc ... -DPMP_POPSTAT -o libpmp_b200_pop.so)
    PMP_B200_BANDS=1 PMP_B200_LIB=.../libpmp_b200_pop.so python tools/popstat.py
The counters cover the plain FIFO ring only.)"""
import sys, os
sys.path.insert(0, "/root/repo")
import numpy as np, torch
import python_motion_planning_b200 as pmp
from python_motion_planning_b200 import workloads
occ = workloads.boxes3d(512, seed=0)
env = pmp.Grid3D(512, 512, 512, walls=False); env.update(occ)
for _ in range(3):
    p = pmp.AStar3D((1, 1, 1), (510, 510, 510), env); c, path, ex = p.plan()
st = [int(v) for v in p.stats]
print("cost", c, "activations", st[1], "kernel_ms", st[8] / 1e6)
print("pops", st[15], "immediate", st[14], "avg pop us", st[13] / max(st[15], 1) / 1965.0, "total pop CTA-ms", st[13] / 1965.0 / 1e3)
