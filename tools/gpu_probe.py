"""Quick on-GPU probe: prints results + kernel counters for a few cases (debug aid)."""
import sys, time, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch
import python_motion_planning_b200 as pmp
from oracle import oracle

def fmt(st):
    st = [int(v) for v in st]
    mhz = 1900.0
    return (f"rounds {st[0]} tiles {st[1]} cells {st[2]} edges {st[3]} wlmax {st[4]} | per-tile us load {st[5]/max(st[1],1)/mhz:.1f} "
            f"settle {st[6]/max(st[1],1)/mhz:.1f} publish {st[7]/max(st[1],1)/mhz:.1f} iters/tile {st[12]/max(st[1],1):.1f} | kernel {st[8]/1e6:.2f} ms: "
            f"scan {st[9]/1e6:.2f} proc {st[10]/1e6:.2f} wait {st[11]/1e6:.2f}")

def t_ms(fn, n=3):
    torch.cuda.synchronize(); out = None; ts = []
    for _ in range(n):
        t0 = time.perf_counter(); out = fn(); torch.cuda.synchronize(); ts.append((time.perf_counter() - t0) * 1e3)
    return out, min(ts)

env = pmp.Grid(51, 31)
obs = env.obstacles
for i in range(10, 21): obs.add((i, 15))
for i in range(15): obs.add((20, i))
for i in range(15, 30): obs.add((30, i))
for i in range(16): obs.add((40, i))
env.update(obs)
for cls in (pmp.AStar, pmp.Dijkstra, pmp.GBFS):
    p = cls((5, 5), (45, 25), env)
    (c, path, ex), ms = t_ms(p.plan)
    print(cls.__name__, repr(c), len(path), len(ex), f"{ms:.2f} ms", "" if p.stats is None else fmt(p.stats), flush=True)

rng = np.random.default_rng(0)
n = int(os.environ.get("PROBE_N", "1024"))
occ = rng.random((n, n)) < 0.30
occ[0, :] = occ[-1, :] = True; occ[:, 0] = occ[:, -1] = True
free = np.argwhere(~occ)
idx = np.random.default_rng(1).integers(0, len(free), size=(64, 2))
starts, goals = free[idx[:, 0]], free[idx[:, 1]]
e2 = pmp.Grid(n, n); e2.update(occ)
for w in (0.0, 12.0, 48.0):
    res, ms = t_ms(lambda: pmp.plan_batch(e2, starts, goals, algo="dijkstra", want_paths=False, window=w), 2)
    print(f"2D {n}^2 batch64 dijkstra window={w}: {ms:.1f} ms  {fmt(res.stats)} reached {int(np.isfinite(res.costs).sum())}", flush=True)
ref_cost, *_ = oracle.plan_batch(occ, starts[:8], goals[:8], algo=oracle.DIJKSTRA)
print("oracle vs gpu (first 8):", np.max(np.abs(np.where(np.isfinite(ref_cost), res.costs[:8] - ref_cost, 0))), flush=True)

m = int(os.environ.get("PROBE_M", "256"))
occ3 = np.zeros((m, m, m), dtype=bool)
for ax in range(3):
    sl = [slice(None)] * 3; sl[ax] = 0; occ3[tuple(sl)] = True; sl[ax] = -1; occ3[tuple(sl)] = True
r = np.random.default_rng(0)
for _ in range(m * 2):
    c = r.integers(1, m - 1, size=3); e = r.integers(4, 33, size=3)
    occ3[c[0]:c[0] + e[0], c[1]:c[1] + e[1], c[2]:c[2] + e[2]] = True
occ3[1:4, 1:4, 1:4] = False; occ3[-4:-1, -4:-1, -4:-1] = False
for ax in range(3):
    sl = [slice(None)] * 3; sl[ax] = 0; occ3[tuple(sl)] = True; sl[ax] = -1; occ3[tuple(sl)] = True
e3 = pmp.Grid3D(m, m, m, walls=False); e3.update(occ3)
for cls in (pmp.AStar3D, pmp.Dijkstra3D):
    for w in (0.0, 4.0, 24.0):
        os.environ["PMP_B200_WINDOW"] = str(w) if w else ""
        p = cls((1, 1, 1), (m - 2, m - 2, m - 2), e3)
        (c, path, ex), ms = t_ms(p.plan, 2)
        print(f"3D {m}^3 {cls.__name__} window={w}: cost {c!r} len {len(path)} {ms:.1f} ms {fmt(p.stats)}", flush=True)
os.environ["PMP_B200_WINDOW"] = ""
for sub in (2, 4):
    os.environ["PMP_B200_SUB"] = str(sub)
    res, ms = t_ms(lambda: pmp.plan_batch(e2, starts, goals, algo="dijkstra", want_paths=False), 2)
    print(f"2D batch64 sub={sub}: {ms:.1f} ms  {fmt(res.stats)}", flush=True)
    p = pmp.Dijkstra3D((1, 1, 1), (m - 2, m - 2, m - 2), e3)
    (c, path, ex), ms = t_ms(p.plan, 2)
    print(f"3D Dijkstra3D sub={sub}: {ms:.1f} ms {fmt(p.stats)}", flush=True)
os.environ["PMP_B200_SUB"] = ""
for wt in ("f64", "u32"):
    e3.motions = [mm for mm in e3.motions if mm.g == 1] if wt == "u32" else e3.motions
    (wsst), ms = t_ms(lambda: pmp.cost_field(e3, (1, 1, 1), weight=wt, to_host=False), 2)
    print(f"3D {m}^3 field {wt} ({len(e3.motions)}-conn): {ms:.1f} ms {fmt(wsst[1].cpu().numpy())}", flush=True)
