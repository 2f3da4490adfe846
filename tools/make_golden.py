"""Generate tests/golden/*.npz by running the UNMODIFIED reference (/root/reference) here.

Run in the build container only (`python tools/make_golden.py`); the fixtures it writes are
committed, this script is the record of how they were made.  The GPU box has no
/root/reference, so tests read only the .npz files.

Fixtures
  csv_replay.npz   the 500 (scenario, seed) queries of examples/benchmark.py:45-89 on
                   Grid3D(21,15,11): occupancy after carve_safety_bubble, start, goal, the
                   CSV's Distance / Visited Nodes for astar/dijkstra/gbfs
                   (3d_pathfinding_results.csv) and the live reference's cost / visited /
                   path / expansion order.
  cases2d.npz      README example (README.md:62-76) and seeded random 2D grids: all three
                   planners, both heuristics; cost, path, expansion order with g and h.
  cases3d.npz      examples/3d_example.py defaults, default Grid3D walls (incl. the z-1 quirk,
                   env3d.py:86-92), 6- and 18-connected motion subsets, random boxes.
  field64.npz      config 5 on a 64^3 crop of its obstacle block (`python tools/make_golden.py field64`):
                   Dijkstra3D, six unit moves, to exhaustion.
  fields.npz       full cost-to-go fields: Dijkstra3D with an unreachable goal
                   (dijkstra3d.py:86-87 returns CLOSED), 26- and 6-connected; 2D Dijkstra to
                   the farthest corner (CLOSED of the settled region).
"""
import csv
import os
import random
import sys

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
import ref_harness  # noqa: E402

OUT = os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", "tests", "golden")


def dense(env, ndim):
    shape = (env.x_range, env.y_range) + ((env.z_range,) if ndim == 3 else ())
    occ = np.zeros(shape, dtype=bool)
    for p in env.obstacles:
        if all(0 <= p[i] < shape[i] for i in range(ndim)):
            occ[p] = True
    return occ


def flat(p, shape):
    if len(shape) == 2:
        return p[1] * shape[0] + p[0]
    return (p[2] * shape[1] + p[1]) * shape[0] + p[0]


def run(planner, shape):
    cost, path, expand = planner.plan()
    ndim = len(shape)
    found = bool(path)
    cost = float(cost) if found else float("inf")
    path_a = np.array(path, dtype=np.int32).reshape(-1, ndim)
    cells = np.array([flat(e.current, shape) for e in expand], dtype=np.int64)
    g = np.array([float(e.g) for e in expand], dtype=np.float64)
    h = np.array([float(e.h) if e.h is not None else 0.0 for e in expand], dtype=np.float64)
    par = np.array([flat(e.parent, shape) for e in expand], dtype=np.int64)
    return dict(cost=cost, found=found, path=path_a, e_cell=cells, e_g=g, e_h=h, e_parent=par)


class Ragged:
    def __init__(self):
        self.items = []

    def add(self, a):
        self.items.append(np.asarray(a))

    def pack(self, prefix, out, dtype=None):
        off = np.zeros(len(self.items) + 1, dtype=np.int64)
        for i, a in enumerate(self.items):
            off[i + 1] = off[i] + len(a)
        if self.items:
            width = max([int(np.prod(a.shape[1:])) for a in self.items if len(a)] + [1])
            cat = np.concatenate([a.reshape(len(a), width) for a in self.items], axis=0)
        else:
            cat = np.zeros((0, 1))
        if dtype is not None:
            cat = cat.astype(dtype)
        out[prefix + "_off"] = off
        out[prefix + "_dat"] = cat


def make_csv_replay(pmp):
    from scenarios import scenarios, carve_safety_bubble
    from algorithms import algorithms
    ref_harness.fast_grid_map(pmp)
    W, H, D = 21, 15, 11
    shape = (W, H, D)
    rows = {}
    with open(ref_harness.REFERENCE_ROOT + "/3d_pathfinding_results.csv") as f:
        for r in csv.DictReader(f):
            if r["Algorithm"] in ("astar", "dijkstra", "gbfs"):
                rows.setdefault((r["Scenario"], r["Algorithm"], int(r["Seed"])), []).append(r)
    scen_names = list(scenarios.keys())
    algos = ["astar", "dijkstra", "gbfs"]
    n = len(scen_names) * 100
    occ_bits = np.zeros((n, (W * H * D + 7) // 8), dtype=np.uint8)
    starts = np.zeros((n, 3), dtype=np.int32)
    goals = np.zeros((n, 3), dtype=np.int32)
    scen_idx = np.zeros(n, dtype=np.int32)
    seeds = np.zeros(n, dtype=np.int32)
    out = {}
    per_algo = {a: dict(csv_dist=np.zeros(n), csv_vis=np.zeros(n, dtype=np.int64),
                        ref_cost=np.zeros(n), ref_vis=np.zeros(n, dtype=np.int64),
                        path=Ragged(), e_cell=Ragged(), e_g=Ragged()) for a in algos}
    k = 0
    for si, scen in enumerate(scen_names):
        for i in range(100):
            env = pmp.Grid3D(W, H, D)
            XR, YR, ZR = env.x_range, env.y_range, env.z_range
            random.seed(i)                                           # benchmark.py:51
            rp = lambda: (random.randint(1, XR - 2), random.randint(1, YR - 2), random.randint(1, ZR - 2))
            start = rp()
            goal = rp()
            while goal == start:
                goal = rp()
            obstacles = scenarios[scen](env)
            carve_safety_bubble(obstacles, start, radius=2)          # benchmark.py:74-75
            carve_safety_bubble(obstacles, goal, radius=2)
            env.update(obstacles)
            occ = dense(env, 3)
            occ_bits[k] = np.packbits(occ.T.reshape(-1).astype(np.uint8), bitorder="little")
            starts[k], goals[k], scen_idx[k], seeds[k] = start, goal, si, i
            for a in algos:
                rr = rows[(scen, a, i)]
                assert len({(r["Distance"], r["Visited Nodes"], r["Start"], r["Goal"]) for r in rr}) == 1
                assert eval(rr[0]["Start"]) == start and eval(rr[0]["Goal"]) == goal
                res = run(algorithms[a](start=start, goal=goal, env=env), shape)
                d = per_algo[a]
                d["csv_dist"][k] = float(rr[0]["Distance"])
                d["csv_vis"][k] = int(rr[0]["Visited Nodes"])
                d["ref_cost"][k] = res["cost"]
                d["ref_vis"][k] = len(res["e_cell"])
                d["path"].add(res["path"])
                keep_expand = a != "dijkstra" or i < 10
                d["e_cell"].add(res["e_cell"] if keep_expand else np.zeros(0, dtype=np.int64))
                d["e_g"].add(res["e_g"] if keep_expand else np.zeros(0))
            k += 1
    out.update(dims=np.array(shape, dtype=np.int32), occ_bits=occ_bits, starts=starts, goals=goals,
               scen_idx=scen_idx, seeds=seeds, scen_names=np.array(scen_names))
    for a in algos:
        d = per_algo[a]
        for key in ("csv_dist", "csv_vis", "ref_cost", "ref_vis"):
            out[f"{a}_{key}"] = d[key]
        d["path"].pack(f"{a}_path", out, np.int16)
        d["e_cell"].pack(f"{a}_ecell", out, np.int32)
        d["e_g"].pack(f"{a}_eg", out, np.float64)
    np.savez_compressed(os.path.join(OUT, "csv_replay.npz"), **out)
    for a in algos:
        d = per_algo[a]
        print(a, "visited == csv:", int((d["ref_vis"] == d["csv_vis"]).sum()), "/", n,
              " max rel dist diff:", float(np.max(np.abs(d["ref_cost"] - d["csv_dist"]) / d["csv_dist"])))


class CaseWriter:
    """A list of (grid, query, planner) cases with the reference's answers, as one npz."""

    def __init__(self, ndim):
        self.ndim = ndim
        self.grids, self.grid_names = [], []
        self.meta = []   # grid idx, algo, heuristic, motion-set id, start, goal
        self.cost, self.found = [], []
        self.path, self.e_cell, self.e_g, self.e_h, self.e_parent = (Ragged() for _ in range(5))

    def add_grid(self, name, occ):
        self.grids.append(occ)
        self.grid_names.append(name)
        return len(self.grids) - 1

    def add(self, gi, algo, heur, mset, start, goal, res):
        s = list(start) + [0] * (3 - self.ndim)
        g = list(goal) + [0] * (3 - self.ndim)
        self.meta.append([gi, algo, heur, mset] + s + g)
        self.cost.append(res["cost"])
        self.found.append(res["found"])
        self.path.add(res["path"])
        self.e_cell.add(res["e_cell"]); self.e_g.add(res["e_g"])
        self.e_h.add(res["e_h"]); self.e_parent.add(res["e_parent"])

    def save(self, fname):
        out = dict(meta=np.array(self.meta, dtype=np.int32), cost=np.array(self.cost),
                   found=np.array(self.found), grid_names=np.array(self.grid_names),
                   n_grids=np.array(len(self.grids)))
        for i, g in enumerate(self.grids):
            out[f"grid{i}_shape"] = np.array(g.shape, dtype=np.int32)
            out[f"grid{i}_bits"] = np.packbits(g.T.reshape(-1).astype(np.uint8), bitorder="little")
        self.path.pack("path", out, np.int32)
        self.e_cell.pack("ecell", out, np.int64)
        self.e_g.pack("eg", out, np.float64)
        self.e_h.pack("eh", out, np.float64)
        self.e_parent.pack("eparent", out, np.int64)
        np.savez_compressed(os.path.join(OUT, fname), **out)


ALGO_ID = {"astar": 0, "dijkstra": 1, "gbfs": 2}
HEUR_ID = {"euclidean": 0, "manhattan": 1}


def readme_env(pmp):
    env = pmp.Grid(51, 31)                      # README.md:65-76
    obstacles = env.obstacles
    for i in range(10, 21):
        obstacles.add((i, 15))
    for i in range(15):
        obstacles.add((20, i))
    for i in range(15, 30):
        obstacles.add((30, i))
    for i in range(16):
        obstacles.add((40, i))
    env.update(obstacles)
    return env


def random_env2d(pmp, n, density, seed):
    rng = np.random.default_rng(seed)
    env = pmp.Grid(n, n)
    obstacles = env.obstacles
    for x, y in np.argwhere(rng.random((n, n)) < density):
        obstacles.add((int(x), int(y)))
    env.update(obstacles)
    return env


def free_cells(occ, rng, k):
    free = np.argwhere(~occ)
    idx = rng.choice(len(free), size=k, replace=False)
    return [tuple(int(v) for v in free[i]) for i in idx]


def make_cases2d(pmp):
    cw = CaseWriter(2)
    classes = {"astar": pmp.AStar, "dijkstra": pmp.Dijkstra, "gbfs": pmp.GBFS}
    env = readme_env(pmp)
    gi = cw.add_grid("readme", dense(env, 2))
    for a, cls in classes.items():
        for heur in ("euclidean", "manhattan"):
            res = run(cls(start=(5, 5), goal=(45, 25), env=env, heuristic_type=heur), (51, 31))
            cw.add(gi, ALGO_ID[a], HEUR_ID[heur], 8, (5, 5), (45, 25), res)
            print("readme", a, heur, repr(res["cost"]), len(res["path"]), len(res["e_cell"]))
    # start == goal, adjacent goal, unreachable goal (inside a sealed pocket)
    res = run(pmp.AStar(start=(5, 5), goal=(5, 5), env=env), (51, 31))
    cw.add(gi, 0, 0, 8, (5, 5), (5, 5), res)
    res = run(pmp.Dijkstra(start=(5, 5), goal=(6, 6), env=env), (51, 31))
    cw.add(gi, 1, 0, 8, (5, 5), (6, 6), res)
    for n, density, seed in ((48, 0.30, 1), (64, 0.30, 2), (64, 0.15, 3), (33, 0.40, 4)):
        env = random_env2d(pmp, n, density, seed)
        occ = dense(env, 2)
        gi = cw.add_grid(f"rand{n}_{density}_{seed}", occ)
        rng = np.random.default_rng(100 + seed)
        pts = free_cells(occ, rng, 12)
        for qi in range(6):
            s, g = pts[2 * qi], pts[2 * qi + 1]
            for a, cls in classes.items():
                heurs = ("euclidean", "manhattan") if a != "dijkstra" and qi < 2 else ("euclidean",)
                for heur in heurs:
                    res = run(cls(start=s, goal=g, env=env, heuristic_type=heur), (n, n))
                    cw.add(gi, ALGO_ID[a], HEUR_ID[heur], 8, s, g, res)
    # 4-connected: env.motions filtered before the planner aliases it (graph_search.py:26)
    env = random_env2d(pmp, 40, 0.25, 7)
    env.motions = [m for m in env.motions if m.g == 1]
    occ = dense(env, 2)
    gi = cw.add_grid("rand40_4conn", occ)
    pts = free_cells(occ, np.random.default_rng(107), 8)
    for qi in range(4):
        s, g = pts[2 * qi], pts[2 * qi + 1]
        for a, cls in classes.items():
            res = run(cls(start=s, goal=g, env=env), (40, 40))
            cw.add(gi, ALGO_ID[a], 0, 4, s, g, res)
    cw.save("cases2d.npz")
    print("cases2d:", len(cw.meta), "cases")


def boxes_env3d(pmp, dims, n_boxes, max_edge, seed, default_walls=False):
    from scenarios import shell_walls
    env = pmp.Grid3D(*dims)
    obs = set(env.obstacles) if default_walls else shell_walls(env)
    rng = np.random.default_rng(seed)
    for _ in range(n_boxes):
        e = rng.integers(1, max_edge + 1, size=3)
        c = [int(rng.integers(1, dims[i] - 1)) for i in range(3)]
        for x in range(c[0], min(dims[0] - 1, c[0] + int(e[0]))):
            for y in range(c[1], min(dims[1] - 1, c[1] + int(e[1]))):
                for z in range(c[2], min(dims[2] - 1, c[2] + int(e[2]))):
                    obs.add((x, y, z))
    env.update(obs)
    return env


def make_cases3d(pmp):
    from scenarios import scenarios, carve_safety_bubble
    ref_harness.fast_grid_map(pmp)
    cw = CaseWriter(3)
    classes = {"astar": pmp.AStar3D, "dijkstra": pmp.Dijkstra3D, "gbfs": pmp.GBFS3D}

    def mk(a, s, g, env, heur="euclidean"):
        if a == "dijkstra":
            return classes[a](start=s, goal=g, env=env)           # dijkstra3d.py:29: no heuristic arg
        return classes[a](start=s, goal=g, env=env, heuristic_type=heur)

    # examples/3d_example.py defaults: 26x20x16 door, (1,1,1) -> (24,18,14), bubble radius 1
    dims = (26, 20, 16)
    env = pmp.Grid3D(*dims)
    s, g = (1, 1, 1), (24, 18, 14)
    obstacles = scenarios["door"](env)
    carve_safety_bubble(obstacles, s, radius=1)
    carve_safety_bubble(obstacles, g, radius=1)
    env.update(obstacles)
    gi = cw.add_grid("3d_example_door", dense(env, 3))
    for a in classes:
        for heur in (("euclidean", "manhattan") if a != "dijkstra" else ("euclidean",)):
            res = run(mk(a, s, g, env, heur), dims)
            cw.add(gi, ALGO_ID[a], HEUR_ID[heur], 26, s, g, res)
            print("3d_example", a, heur, repr(res["cost"]), len(res["path"]), len(res["e_cell"]))
    # default Grid3D walls (side walls written at z-1, env3d.py:86-92) + boxes
    dims = (18, 14, 10)
    env = boxes_env3d(pmp, dims, 14, 4, 11, default_walls=True)
    occ = dense(env, 3)
    gi = cw.add_grid("default_walls_boxes", occ)
    pts = free_cells(occ, np.random.default_rng(211), 8)
    for qi in range(4):
        s, g = pts[2 * qi], pts[2 * qi + 1]
        for a in classes:
            cw.add(gi, ALGO_ID[a], 0, 26, s, g, run(mk(a, s, g, env), dims))
    # start == goal and an unreachable goal (an obstacle cell)
    cw.add(gi, 0, 0, 26, pts[0], pts[0], run(mk("astar", pts[0], pts[0], env), dims))
    cw.add(gi, 0, 0, 26, pts[0], (0, 0, 0), run(mk("astar", pts[0], (0, 0, 0), env), dims))
    # 6- and 18-connected subsets of env.motions
    for mset, keep in ((6, lambda m: m.g == 1), (18, lambda m: m.g < 1.5)):
        dims = (16, 16, 12)
        env = boxes_env3d(pmp, dims, 16, 4, 20 + mset)
        env.motions = [m for m in env.motions if keep(m)]
        assert len(env.motions) == mset
        occ = dense(env, 3)
        gi = cw.add_grid(f"boxes_{mset}conn", occ)
        pts = free_cells(occ, np.random.default_rng(300 + mset), 6)
        for qi in range(3):
            s, g = pts[2 * qi], pts[2 * qi + 1]
            for a in classes:
                heur = "manhattan" if (mset == 6 and a == "astar") else "euclidean"
                cw.add(gi, ALGO_ID[a], HEUR_ID[heur], mset, s, g, run(mk(a, s, g, env, heur), dims))
    cw.save("cases3d.npz")
    print("cases3d:", len(cw.meta), "cases")


def make_fields(pmp):
    ref_harness.fast_grid_map(pmp)
    out = {}
    # 3D: Dijkstra3D with goal = an obstacle cell => exhausts and returns CLOSED (dijkstra3d.py:86-87)
    for name, dims, mset, seed in (("f3d_26", (24, 24, 24), 26, 5), ("f3d_6", (28, 28, 28), 6, 6),
                                   ("f3d_18", (20, 20, 20), 18, 7)):
        env = boxes_env3d(pmp, dims, 40, 6, seed)
        if mset == 6:
            env.motions = [m for m in env.motions if m.g == 1]
        elif mset == 18:
            env.motions = [m for m in env.motions if m.g < 1.5]
        occ = dense(env, 3)
        src = free_cells(occ, np.random.default_rng(seed), 1)[0]
        cost, path, expand = pmp.Dijkstra3D(start=src, goal=(0, 0, 0), env=env).plan()
        assert cost == float("inf") and path == []
        field = np.full(dims, np.inf)
        for e in expand:
            field[e.current] = e.g
        out[name + "_shape"] = np.array(dims, dtype=np.int32)
        out[name + "_bits"] = np.packbits(occ.T.reshape(-1).astype(np.uint8), bitorder="little")
        out[name + "_src"] = np.array(src, dtype=np.int32)
        out[name + "_mset"] = np.array(mset)
        out[name + "_g"] = field
        print(name, "settled", len(expand), "of", int((~occ).sum()), "max g", float(np.max(field[np.isfinite(field)])))
    # 2D: Dijkstra to the far corner; CLOSED holds the settled disc
    for name, n, density, seed in (("f2d_a", 96, 0.30, 8), ("f2d_b", 80, 0.10, 9)):
        env = random_env2d(pmp, n, density, seed)
        occ = dense(env, 2)
        free = np.argwhere(~occ)
        src = tuple(int(v) for v in free[0])
        for cand in free[::-1]:         # farthest-listed free cell that is reachable
            goal = tuple(int(v) for v in cand)
            cost, path, expand = pmp.Dijkstra(start=src, goal=goal, env=env).plan()
            if path:
                break
        field = np.full((n, n), np.inf)
        for e in expand:
            field[e.current] = e.g
        out[name + "_shape"] = np.array((n, n), dtype=np.int32)
        out[name + "_bits"] = np.packbits(occ.T.reshape(-1).astype(np.uint8), bitorder="little")
        out[name + "_src"] = np.array(src, dtype=np.int32)
        out[name + "_goal"] = np.array(goal, dtype=np.int32)
        out[name + "_cost"] = np.array(float(cost) if path else np.inf)
        out[name + "_g"] = field
        print(name, "closed", len(expand), "cost", cost)
    np.savez_compressed(os.path.join(OUT, "fields.npz"), **out)


def make_field64(pmp):
    """BASELINE.json config 5 on a 64^3 crop of its obstacle block: the unmodified Dijkstra3D with
    env.motions filtered to the six unit moves, run to exhaustion (dijkstra3d.py:39-87).  Closes the
    chain reference -> oracle -> kernel at a multi-tile size (tests/test_bfs_gpu.py,
    tests/test_oracle_golden.py)."""
    ref_harness.fast_grid_map(pmp)
    sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), ".."))
    from python_motion_planning_b200 import workloads
    blk, _ = workloads.replicated_boxes3d(base=128, reps=1)
    n = 64
    occ = blk[:n, :n, :n].copy()
    workloads.seal(occ)
    env = pmp.Grid3D(n, n, n)
    env.update(set(map(tuple, np.argwhere(occ).tolist())))
    env.motions = [m for m in env.motions if m.g == 1]
    assert len(env.motions) == 6
    src = (1, 1, 1)
    assert not occ[src]
    cost, path, expand = pmp.Dijkstra3D(start=src, goal=(0, 0, 0), env=env).plan()
    assert cost == float("inf") and path == []
    field = np.full((n, n, n), np.inf)
    for e in expand:
        field[e.current] = e.g
    print("field64: settled", len(expand), "of", int((~occ).sum()), "max g", float(np.max(field[np.isfinite(field)])))
    np.savez_compressed(os.path.join(OUT, "field64.npz"), occ=occ, src=np.array(src, dtype=np.int32), g=field)


if __name__ == "__main__":
    os.makedirs(OUT, exist_ok=True)
    pmp = ref_harness.load_reference()
    which = sys.argv[1:] or ["csv", "2d", "3d", "fields"]
    if "csv" in which:
        make_csv_replay(pmp)
    if "2d" in which:
        make_cases2d(pmp)
    if "3d" in which:
        make_cases3d(pmp)
    if "fields" in which:
        make_fields(pmp)
    if "field64" in which:
        make_field64(pmp)
