"""Golden vectors for the reference-order any-angle planners (pmp_bestfirst, algo THETA / LAZY_THETA),
written from the UNMODIFIED reference in the build container (needs /root/reference):

  * anyangle2d.npz / anyangle3d.npz -- ThetaStar / LazyThetaStar (theta_star.py, lazy_theta_star.py)
    and ThetaStar3D / LazyThetaStar3D (theta_star3d.py, lazy_theta_star3d.py) on the grids of
    cases2d.npz and on seeded 3D box maps (26- and 18-connected): cost, path, CLOSED in insertion
    order with g, h and parent -- same layout as cases2d.npz (tests/golden_io.py:Cases),
    algo id 3 = Theta*, 4 = Lazy Theta*;
  * csv_anyangle.npz -- the Distance / Visited Nodes columns of the theta_star and lazy_theta_star
    rows of the reference's 3d_pathfinding_results.csv for the 500 instances of csv_replay.npz,
    next to what the reference returns when run here.

  * jpscases2d.npz / jpscases3d.npz -- JPS (jps.py) and JPS3D (jps3d.py) in the same layout, algo id 5:
    the jump-point path, CLOSED order, g, h, parent.

    python tools/make_golden_anyangle.py
"""
import csv
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tools"))
import ref_harness  # noqa: E402
import make_golden as mg  # noqa: E402
from tests import golden_io as gio  # noqa: E402

THETA, LAZY, JPS = 3, 4, 5


def env2d_from(pmp, occ):
    env = pmp.Grid(*occ.shape)
    env.update(set(map(tuple, np.argwhere(occ).tolist())))
    return env


def make_2d(pmp):
    cw = mg.CaseWriter(2)
    classes = {THETA: pmp.ThetaStar, LAZY: pmp.LazyThetaStar}
    c2 = gio.Cases("cases2d.npz", 2)
    env = mg.readme_env(pmp)
    gi = cw.add_grid("readme", mg.dense(env, 2))
    for a, cls in classes.items():
        for heur in ("euclidean", "manhattan"):
            res = mg.run(cls(start=(5, 5), goal=(45, 25), env=env, heuristic_type=heur), (51, 31))
            cw.add(gi, a, mg.HEUR_ID[heur], 8, (5, 5), (45, 25), res)
            print("readme", a, heur, repr(res["cost"]), len(res["path"]), len(res["e_cell"]))
        cw.add(gi, a, 0, 8, (5, 5), (5, 5), mg.run(cls(start=(5, 5), goal=(5, 5), env=env), (51, 31)))
    for name, occ in zip(c2.grid_names[1:5], c2.grids[1:5]):
        env = env2d_from(pmp, occ)
        gi = cw.add_grid(name, occ)
        rng = np.random.default_rng(900 + gi)
        pts = mg.free_cells(occ, rng, 10)
        for qi in range(5):
            s, g = pts[2 * qi], pts[2 * qi + 1]
            for a, cls in classes.items():
                res = mg.run(cls(start=s, goal=g, env=env), occ.shape)
                cw.add(gi, a, 0, 8, s, g, res)
    # 4-connected: env.motions filtered before the planner aliases it (graph_search.py:26)
    occ = c2.grids[2]
    env = env2d_from(pmp, occ)
    env.motions = [m for m in env.motions if m.g == 1]
    gi = cw.add_grid(c2.grid_names[2] + "_4conn", occ)
    pts = mg.free_cells(occ, np.random.default_rng(950), 8)
    for qi in range(4):
        s, g = pts[2 * qi], pts[2 * qi + 1]
        for a, cls in classes.items():
            cw.add(gi, a, 0, 4, s, g, mg.run(cls(start=s, goal=g, env=env), occ.shape))
    cw.save("anyangle2d.npz")
    print("anyangle2d:", len(cw.meta), "cases;", sum(cw.found), "with a path")


def make_3d(pmp):
    ref_harness.fast_grid_map(pmp)
    cw = mg.CaseWriter(3)
    classes = {THETA: pmp.ThetaStar3D, LAZY: pmp.LazyThetaStar3D}
    for mset, dims, seed in ((26, (18, 14, 10), 31), (26, (16, 16, 12), 32), (18, (16, 13, 10), 33),
                             (6, (14, 12, 10), 34)):
        env = mg.boxes_env3d(pmp, dims, 14, 4, seed)
        if mset == 18:
            env.motions = [m for m in env.motions if m.g < 1.5]
        elif mset == 6:
            env.motions = [m for m in env.motions if m.g == 1]
        occ = mg.dense(env, 3)
        gi = cw.add_grid(f"boxes_{mset}conn_{seed}", occ)
        pts = mg.free_cells(occ, np.random.default_rng(400 + seed), 8)
        for qi in range(4):
            s, g = pts[2 * qi], pts[2 * qi + 1]
            for a, cls in classes.items():
                heur = "manhattan" if qi == 3 else "euclidean"
                res = mg.run(cls(start=s, goal=g, env=env, heuristic_type=heur), dims)
                cw.add(gi, a, mg.HEUR_ID[heur], mset, s, g, res)
        for a, cls in classes.items():   # start == goal, goal on an obstacle
            cw.add(gi, a, 0, mset, pts[0], pts[0], mg.run(cls(start=pts[0], goal=pts[0], env=env), dims))
            cw.add(gi, a, 0, mset, pts[0], (0, 0, 0), mg.run(cls(start=pts[0], goal=(0, 0, 0), env=env), dims))
    cw.save("anyangle3d.npz")
    print("anyangle3d:", len(cw.meta), "cases;", sum(cw.found), "with a path")


def make_jps(pmp):
    sys.setrecursionlimit(100000)
    ref_harness.fast_grid_map(pmp)
    cw = mg.CaseWriter(2)
    c2 = gio.Cases("cases2d.npz", 2)
    env = mg.readme_env(pmp)
    gi = cw.add_grid("readme", mg.dense(env, 2))
    for heur in ("euclidean", "manhattan"):
        res = mg.run(pmp.JPS(start=(5, 5), goal=(45, 25), env=env, heuristic_type=heur), (51, 31))
        cw.add(gi, JPS, mg.HEUR_ID[heur], 8, (5, 5), (45, 25), res)
        print("readme jps", heur, repr(res["cost"]), len(res["path"]), len(res["e_cell"]))
    cw.add(gi, JPS, 0, 8, (5, 5), (5, 5), mg.run(pmp.JPS(start=(5, 5), goal=(5, 5), env=env), (51, 31)))
    for name, occ in zip(c2.grid_names[1:5], c2.grids[1:5]):
        env = env2d_from(pmp, occ)
        gi = cw.add_grid(name, occ)
        pts = mg.free_cells(occ, np.random.default_rng(700 + gi), 16)
        for qi in range(8):
            s, g = pts[2 * qi], pts[2 * qi + 1]
            cw.add(gi, JPS, 0, 8, s, g, mg.run(pmp.JPS(start=s, goal=g, env=env), occ.shape))
    occ = c2.grids[2]
    env = env2d_from(pmp, occ)
    env.motions = [m for m in env.motions if m.g == 1]
    gi = cw.add_grid(c2.grid_names[2] + "_4conn", occ)
    pts = mg.free_cells(occ, np.random.default_rng(750), 8)
    for qi in range(4):
        s, g = pts[2 * qi], pts[2 * qi + 1]
        cw.add(gi, JPS, 0, 4, s, g, mg.run(pmp.JPS(start=s, goal=g, env=env), occ.shape))
    cw.save("jpscases2d.npz")
    print("jpscases2d:", len(cw.meta), "cases;", sum(cw.found), "with a path")
    cw = mg.CaseWriter(3)
    for mset, dims, seed, nb in ((26, (18, 14, 10), 41, 14), (26, (16, 16, 12), 42, 24), (18, (16, 13, 10), 43, 14),
                                 (6, (14, 12, 10), 44, 12)):
        env = mg.boxes_env3d(pmp, dims, nb, 4, seed)
        if mset == 18:
            env.motions = [m for m in env.motions if m.g < 1.5]
        elif mset == 6:
            env.motions = [m for m in env.motions if m.g == 1]
        occ = mg.dense(env, 3)
        gi = cw.add_grid(f"boxes_{mset}conn_{seed}", occ)
        pts = mg.free_cells(occ, np.random.default_rng(500 + seed), 12)
        for qi in range(6):
            s, g = pts[2 * qi], pts[2 * qi + 1]
            heur = "manhattan" if qi == 5 else "euclidean"
            res = mg.run(pmp.JPS3D(start=s, goal=g, env=env, heuristic_type=heur), dims)
            cw.add(gi, JPS, mg.HEUR_ID[heur], mset, s, g, res)
        cw.add(gi, JPS, 0, mset, pts[0], pts[0], mg.run(pmp.JPS3D(start=pts[0], goal=pts[0], env=env), dims))
        cw.add(gi, JPS, 0, mset, pts[0], (0, 0, 0), mg.run(pmp.JPS3D(start=pts[0], goal=(0, 0, 0), env=env), dims))
    # the benchmark scenarios of examples/scenarios.py (21 x 15 x 11), first 4 seeds of each
    z = gio.load("csv_replay.npz")
    dims = tuple(int(v) for v in z["dims"])
    for k in range(len(z["starts"])):
        if int(z["seeds"][k]) >= 4:
            continue
        occ = gio.unpack_occ(z["occ_bits"][k], dims)
        env = pmp.Grid3D(*dims)
        env.update(set(map(tuple, np.argwhere(occ).tolist())))
        gi = cw.add_grid(f"csv_{k}", occ)
        s, g = tuple(int(v) for v in z["starts"][k]), tuple(int(v) for v in z["goals"][k])
        cw.add(gi, JPS, 0, 26, s, g, mg.run(pmp.JPS3D(start=s, goal=g, env=env), dims))
    cw.save("jpscases3d.npz")
    print("jpscases3d:", len(cw.meta), "cases;", sum(cw.found), "with a path")


def make_csv(pmp):
    ref_harness.fast_grid_map(pmp)
    z = gio.load("csv_replay.npz")
    dims = tuple(int(v) for v in z["dims"])
    scen_names = [str(s) for s in z["scen_names"]]
    n = len(z["starts"])
    rows = {}
    with open(ref_harness.REFERENCE_ROOT + "/3d_pathfinding_results.csv") as f:
        for r in csv.DictReader(f):
            if r["Algorithm"] in ("theta_star", "lazy_theta_star"):
                rows.setdefault((r["Scenario"], r["Algorithm"], int(r["Seed"])), []).append(r)
    out = {}
    for a, cls in (("theta_star", pmp.ThetaStar3D), ("lazy_theta_star", pmp.LazyThetaStar3D)):
        csv_dist, csv_vis = np.zeros(n), np.zeros(n, dtype=np.int64)
        ref_cost, ref_vis = np.zeros(n), np.zeros(n, dtype=np.int64)
        for k in range(n):
            occ = gio.unpack_occ(z["occ_bits"][k], dims)
            start = tuple(int(v) for v in z["starts"][k])
            goal = tuple(int(v) for v in z["goals"][k])
            rr = rows[(scen_names[int(z["scen_idx"][k])], a, int(z["seeds"][k]))]
            assert len({(r["Distance"], r["Visited Nodes"]) for r in rr}) == 1
            assert eval(rr[0]["Start"]) == start and eval(rr[0]["Goal"]) == goal
            env = pmp.Grid3D(*dims)
            env.update(set(map(tuple, np.argwhere(occ).tolist())))
            cost, path, expand = cls(start=start, goal=goal, env=env).plan()
            csv_dist[k] = float(rr[0]["Distance"])
            csv_vis[k] = int(rr[0]["Visited Nodes"])
            ref_cost[k] = float(cost)
            ref_vis[k] = len(expand)
        out[f"{a}_csv_dist"], out[f"{a}_csv_vis"] = csv_dist, csv_vis
        out[f"{a}_ref_cost"], out[f"{a}_ref_vis"] = ref_cost, ref_vis
        fin = np.isfinite(csv_dist)
        print(a, "visited == csv:", int((ref_vis == csv_vis).sum()), "/", n, " max rel dist diff:",
              float(np.max(np.abs(ref_cost[fin] - csv_dist[fin]) / csv_dist[fin])))
    np.savez_compressed(os.path.join(mg.OUT, "csv_anyangle.npz"), **out)


if __name__ == "__main__":
    pmp = ref_harness.load_reference()
    which = sys.argv[1:] or ["2d", "3d", "jps", "csv"]
    if "2d" in which:
        make_2d(pmp)
    if "3d" in which:
        make_3d(pmp)
    if "jps" in which:
        make_jps(pmp)
    if "csv" in which:
        make_csv(pmp)
