"""one unit-weight field search (c5 block replicated to n^3) for ncu: python tools/prof_bfs.py [n] [reps]"""
import sys
sys.path.insert(0, ".")
import python_motion_planning_b200 as pmp   # noqa: E402
from python_motion_planning_b200 import workloads   # noqa: E402
import torch   # noqa: E402

n = int(sys.argv[1]) if len(sys.argv) > 1 else 512
reps = int(sys.argv[2]) if len(sys.argv) > 2 else 2
blk, _ = workloads.replicated_boxes3d(base=128, reps=n // 128)
env = pmp.Grid3D(n, n, n, walls=False)
env.motions = [m for m in env.motions if m.g == 1]
env.update_packed(workloads.replicated_packed(blk, (n, n, n), 0, n))
for _ in range(reps):
    ws, st = pmp.cost_field(env, (1, 1, 1), weight="u32", to_host=False)
torch.cuda.synchronize()
print("kernel ms", ws.stats.cpu()[8].item() / 1e6, "status", ws.status.cpu().tolist())
