"""e2e of the c5 field with the streamed occupancy upload: where does the time go?"""
import sys, time, os
sys.path.insert(0, ".")
import numpy as np, torch
import python_motion_planning_b200 as pmp
from python_motion_planning_b200 import workloads
n = int(sys.argv[1]) if len(sys.argv) > 1 else 1536
blk, _ = workloads.replicated_boxes3d(base=128, reps=n // 128)
rows = torch.from_numpy(workloads.replicated_packed(blk, (n, n, n), 0, n).view(np.int32)).pin_memory()
env = pmp.Grid3D(n, n, n, walls=False)
env.motions = [m for m in env.motions if m.g == 1]
for mode in ("streamed", "plain", "streamed"):
    if mode == "plain":
        os.environ["PMP_B200_NO_STREAMED_UPLOAD"] = "1"
    else:
        os.environ.pop("PMP_B200_NO_STREAMED_UPLOAD", None)
    for rep in range(4):
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        env.update_packed(rows)
        ws, st = pmp.cost_field(env, (1, 1, 1), weight="u32", to_host=False)
        t1 = time.perf_counter()
        v = int(ws.g[0][-1].item())
        t2 = time.perf_counter()
        s = ws.stats.cpu().numpy()
    print(f"{mode}: host enqueue {1e3*(t1-t0):.2f} ms, total {1e3*(t2-t0):.2f} ms, k_bfs {s[8]/1e6:.2f} ms, activations {s[1]}, "
          f"wait/act {s[4]/max(s[1],1)/1965:.1f} us setup/act {s[6]/max(s[1],1)/1965:.1f} us", flush=True)
# upload alone
dev = torch.empty_like(rows, device="cuda")
torch.cuda.synchronize(); t0 = time.perf_counter(); dev.copy_(rows, non_blocking=True); torch.cuda.synchronize()
print(f"plain H2D of {rows.numel()*4/1e6:.0f} MB: {1e3*(time.perf_counter()-t0):.2f} ms")
