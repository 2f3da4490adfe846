"""several z-slabs of the config-5 grid on ONE GPU in one launch (k_bfs_group, free-running warps):
the slab protocol without NVLink latency.   python tools/gpu_slab_group.py [n] [nslab ...]"""
import sys
sys.path.insert(0, ".")
import numpy as np   # noqa: E402
import torch   # noqa: E402
from python_motion_planning_b200 import slab, workloads   # noqa: E402

n = int(sys.argv[1]) if len(sys.argv) > 1 else 768
counts = [int(a) for a in sys.argv[2:]] or [1, 2, 4]
blk, _ = workloads.replicated_boxes3d(base=128, reps=n // 128)
for ns in counts:
    slabs = []
    for p in range(ns):
        z0, z1 = slab.slab_range8(n, p, ns)
        slabs.append(slab.BfsSlab(workloads.replicated_packed(blk, (n, n, n), z0, z1), p, ns, (n, n, z1 - z0)))
    for rep in range(2):
        slab.solve_group(slabs, n, (1, 1, 1))
    torch.cuda.synchronize()
    sts = [s.stats_host() for s in slabs]
    st = np.sum(sts, axis=0)
    ms = max(float(a[8]) for a in sts) / 1e6
    reached = sum(int((s.field != -1).sum().item()) for s in slabs)
    print(f"{n}^3 in {ns} slab(s), one launch: kernel {ms:.2f} ms, reached {reached}, activations {st[1]} idle {st[13]} lowered {st[2]} "
          f"requeues {st[5]} remote pushes {st[7]}; busy until " + " ".join(f"{(a[16]-a[9])/1e6:.1f}" for a in sts), flush=True)
    for s in slabs:
        s.close()
