"""Golden vectors for the any-angle row, written from the UNMODIFIED
reference in the build container (needs /root/reference):

  * lineOfSight answers: ThetaStar.lineOfSight (theta_star.py:110-171) on 2D grids and
    ThetaStar3D.lineOfSight (theta_star3d.py:139-215) on the benchmark's 3D scenario grids, for
    seeded random ordered pairs (both directions, blocked endpoints and a == b included);
  * the Theta* / lazy Theta* `Distance` column of the reference's 3d_pathfinding_results.csv in
    the row order of tests/golden/csv_replay.npz, and ThetaStar (2D) costs on the 2D grids.

    python tools/make_golden_los.py        ->  tests/golden/los.npz
"""
import csv
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tools"))
import ref_harness  # noqa: E402
from tests import golden_io as gio  # noqa: E402

OUT = os.path.join(ROOT, "tests", "golden")


def pack(occ):
    return np.packbits(np.ascontiguousarray(occ.T).reshape(-1), bitorder="little")


def env_from_occ(pmp, occ):
    if occ.ndim == 2:
        env = pmp.Grid(*occ.shape)
    else:
        env = pmp.Grid3D(*occ.shape)
    env.update(set(map(tuple, np.argwhere(occ).tolist())))
    return env


def main():
    pmp = ref_harness.load_reference()
    ref_harness.fast_grid_map(pmp)
    rng = np.random.default_rng(3)
    out = {}
    grids = []
    # 2D: the README grid and two random grids
    c2 = gio.Cases("cases2d.npz", 2)
    for occ in c2.grids[:3]:
        grids.append(occ)
    # 3D: scenario grids of the benchmark (one per scenario, from the CSV replay fixture)
    z = gio.load("csv_replay.npz")
    dims = tuple(int(v) for v in z["dims"])
    scen = z["scen_idx"]
    for s in np.unique(scen):
        k = int(np.flatnonzero(scen == s)[7])
        grids.append(gio.unpack_occ(z["occ_bits"][k], dims))
    out["n_grids"] = np.array(len(grids))
    for gi, occ in enumerate(grids):
        ndim = occ.ndim
        env = env_from_occ(pmp, occ)
        free = np.argwhere(~occ)
        start, goal = tuple(int(v) for v in free[0]), tuple(int(v) for v in free[-1])
        planner = (pmp.ThetaStar if ndim == 2 else pmp.ThetaStar3D)(start=start, goal=goal, env=env)
        node = pmp.Node if ndim == 2 else pmp.Node3D
        n = 1500
        a = np.stack([rng.integers(0, s, size=n) for s in occ.shape], axis=1)
        b = np.stack([rng.integers(0, s, size=n) for s in occ.shape], axis=1)
        # a third of the pairs: short segments, where most of the True answers are
        near = rng.random(n) < 0.35
        off = rng.integers(-6, 7, size=(n, ndim))
        b[near] = np.clip(a[near] + off[near], 0, np.array(occ.shape) - 1)
        # and free endpoints for half of them
        pick = rng.random(n) < 0.5
        a[pick] = free[rng.integers(0, len(free), size=int(pick.sum()))]
        pick = rng.random(n) < 0.5
        b[pick] = free[rng.integers(0, len(free), size=int(pick.sum()))]
        a[:5] = b[:5]
        ans = np.zeros(n, dtype=np.uint8)
        for i in range(n):
            ans[i] = bool(planner.lineOfSight(node(tuple(int(v) for v in a[i]), None, 0, 0),
                                              node(tuple(int(v) for v in b[i]), None, 0, 0)))
        out[f"g{gi}_shape"] = np.array(occ.shape, dtype=np.int32)
        out[f"g{gi}_bits"] = pack(occ)
        out[f"g{gi}_a"] = a.astype(np.int32)
        out[f"g{gi}_b"] = b.astype(np.int32)
        out[f"g{gi}_los"] = ans
        print("grid", gi, occ.shape, "pairs", n, "visible", int(ans.sum()))
    # Theta* distances of the reference's CSV, aligned with csv_replay.npz (scenario-major, seed)
    rows = {}
    with open(ref_harness.REFERENCE_ROOT + "/3d_pathfinding_results.csv") as f:
        for r in csv.DictReader(f):
            if r["Algorithm"] in ("theta_star", "lazy_theta_star"):
                rows.setdefault((r["Scenario"], r["Algorithm"], int(r["Seed"])), set()).add(float(r["Distance"]))
    names = [str(s) for s in z["scen_names"]]
    for algo in ("theta_star", "lazy_theta_star"):
        d = np.zeros(len(scen))
        for k in range(len(scen)):
            v = rows[(names[int(scen[k])], algo, int(z["seeds"][k]))]
            assert len(v) == 1
            d[k] = v.pop()
        out[f"{algo}_csv_dist"] = d
    # ThetaStar (2D) on the 2D grids
    q2 = []
    for gi, occ in enumerate(grids[:3]):
        env = env_from_occ(pmp, occ)
        free = np.argwhere(~occ)
        for _ in range(12):
            i, j = rng.integers(0, len(free), size=2)
            s, g = tuple(int(v) for v in free[i]), tuple(int(v) for v in free[j])
            if s == g:
                continue
            cost, path, _ = pmp.ThetaStar(start=s, goal=g, env=env).plan()
            q2.append((gi, s[0], s[1], g[0], g[1], float(cost) if path else float("inf")))
    out["theta2d_queries"] = np.array(q2, dtype=np.float64)
    np.savez_compressed(os.path.join(OUT, "los.npz"), **out)
    print("wrote", os.path.join(OUT, "los.npz"), "2D theta queries", len(q2))


if __name__ == "__main__":
    main()
