"""Golden vectors for the reference-order incremental planners (pmp_lpa_star / pmp_d_star), written
from the UNMODIFIED reference in the build container (needs /root/reference):

  * incr.npz -- LPAStar / LPAStar3D and DStar / DStar3D on seeded maps: plan(), then a sequence of
    occupancy changes through the reference's own entry points (LPAStar3D.apply_change; the body of
    LPAStar.OnPress; DStar3D.apply_dynamic_obstacles; the body of DStar.OnPress), recording after
    every step the cost, the path and the EXPAND list (cells in order).
  * csv_incr.npz -- Distance / Visited Nodes of the lpastar and dstar rows of the reference's
    3d_pathfinding_results.csv for the 500 instances of csv_replay.npz, next to a live run.

    python tools/make_golden_incr.py
"""
import csv
import os
import signal
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tools"))
import ref_harness  # noqa: E402
import make_golden as mg  # noqa: E402
from make_golden_replan import occupancy, pack, toggle_2d  # noqa: E402
from tests import golden_io as gio  # noqa: E402


class Timeout(Exception):
    pass


def _alarm(signum, frame):
    raise Timeout()


def flat(p, shape):
    return mg.flat(p, shape)


def dstar_press_2d(planner, xy):
    """d_star.py:108-127 without the plot calls"""
    planner.obstacles.add(xy)
    planner.env.update(planner.obstacles)
    node = planner.start
    planner.EXPAND, path, cost = [], [], 0
    while node != planner.goal:
        node_parent = planner.map[node.parent]
        if planner.isCollision(node, node_parent):
            planner.modify(node, node_parent)
            continue
        path.append(node.current)
        cost += planner.cost(node, node_parent)
        node = node_parent
    return cost, path


def dlite_press_2d(planner, xy):
    """d_star_lite.py:64-96 without the plot calls"""
    x, y = xy
    cur_start, new_start = planner.start, planner.start
    update_start = True
    cost, count = 0, 0
    path = [planner.start.current]
    planner.EXPAND = []
    while cur_start != planner.goal:
        neighbors = [node_n for node_n in planner.getNeighbor(cur_start)
                     if not planner.isCollision(cur_start, node_n)]
        next_node = min(neighbors, key=lambda n: n.g)
        path.append(next_node.current)
        cost += planner.cost(cur_start, next_node)
        count += 1
        cur_start = next_node
        if update_start:
            update_start = False
            planner.km = planner.h(cur_start, new_start)
            new_start = cur_start
            node_change = planner.map[(x, y)]
            if (x, y) not in planner.obstacles:
                planner.obstacles.add((x, y))
            else:
                planner.obstacles.remove((x, y))
                planner.updateVertex(node_change)
            planner.env.update(planner.obstacles)
            for node_n in planner.getNeighbor(node_change):
                planner.updateVertex(node_n)
            planner.computeShortestPath()
    return cost, path


def record(steps, shape, cost, path, expand):
    steps["cost"].append(float(cost))
    steps["path"].add(np.array(path, dtype=np.int32).reshape(-1, len(shape)))
    steps["expand"].add(np.array([flat(e.current, shape) for e in (expand or []) if e is not None], dtype=np.int64))


def random_map(pmp, rng, dims, density, start, goal):
    env = pmp.Grid(*dims) if len(dims) == 2 else pmp.Grid3D(*dims)
    obs = env.obstacles
    it = np.ndindex(*[d - 2 for d in dims])
    for idx in it:
        c = tuple(int(v) + 1 for v in idx)
        if rng.random() < density and c not in (start, goal):
            obs.add(c)
    env.update(obs)
    return env


def main_incr(pmp):
    ref_harness.fast_grid_map(pmp)
    signal.signal(signal.SIGALRM, _alarm)
    rng = np.random.default_rng(23)
    out = {}
    specs = [("lpa", (31, 21), 0.12, 12), ("lpa", (36, 28), 0.2, 12), ("lpa", (14, 12, 10), 0.1, 10),
             ("lpa", (18, 10, 9), 0.18, 10), ("dstar", (31, 21), 0.12, 8), ("dstar", (36, 28), 0.18, 8),
             ("dstar", (14, 12, 10), 0.1, 8), ("dstar", (18, 10, 9), 0.16, 8),
             ("dlite", (31, 21), 0.12, 10), ("dlite", (36, 28), 0.2, 10)]
    k = 0
    for kind, dims, density, n_changes in specs:
        ndim = len(dims)
        start = (2, 2) if ndim == 2 else (1, 1, 1)
        goal = tuple(d - 3 for d in dims) if ndim == 2 else tuple(d - 2 for d in dims)
        env = random_map(pmp, rng, dims, density, start, goal)
        occ0 = occupancy(env, dims)
        if kind == "lpa":
            planner = (pmp.LPAStar if ndim == 2 else pmp.LPAStar3D)(start=start, goal=goal, env=env)
        elif kind == "dlite":
            planner = pmp.DStarLite(start=start, goal=goal, env=env)
        else:
            planner = (pmp.DStar if ndim == 2 else pmp.DStar3D)(start=start, goal=goal, env=env)
        steps = dict(cost=[], path=mg.Ragged(), expand=mg.Ragged())
        signal.alarm(20)
        try:
            cost, path, _ = planner.plan()
        except (ValueError, Timeout):
            signal.alarm(0)
            print("skipped", kind, dims)
            continue
        signal.alarm(0)
        record(steps, dims, cost, path, planner.EXPAND)
        changes, flags = [], []
        for i in range(n_changes):
            cur = path
            if cur and len(cur) > 4 and i % 2 == 0:
                c = tuple(cur[int(rng.integers(2, len(cur) - 2))])
            else:
                c = tuple(int(rng.integers(1, s - 1)) for s in dims)
            if c in (start, goal):
                continue
            signal.alarm(20)
            try:
                if kind == "dlite":
                    now_blocked = c not in planner.obstacles
                    cost, path = dlite_press_2d(planner, c)
                elif kind == "lpa":
                    now_blocked = c not in planner.obstacles
                    if ndim == 2:
                        cost, path, _ = toggle_2d(planner, c)
                    else:
                        cost, path, _ = planner.apply_change(c)
                else:
                    if c in planner.obstacles:
                        signal.alarm(0)
                        continue
                    now_blocked = True
                    if ndim == 2:
                        cost, path = dstar_press_2d(planner, c)
                    else:
                        cost, path = planner.apply_dynamic_obstacles([c])
            except (ValueError, KeyError, Timeout) as e:
                signal.alarm(0)
                print("  sequence ends:", kind, dims, type(e).__name__)
                break
            signal.alarm(0)
            changes.append(c)
            flags.append(now_blocked)
            record(steps, dims, cost, path, planner.EXPAND)
        out[f"c{k}_kind"] = np.array({"lpa": 0, "dstar": 1, "dlite": 2}[kind])
        out[f"c{k}_shape"] = np.array(dims, dtype=np.int32)
        out[f"c{k}_bits"] = pack(occ0)
        out[f"c{k}_start"] = np.array(start, dtype=np.int32)
        out[f"c{k}_goal"] = np.array(goal, dtype=np.int32)
        out[f"c{k}_changes"] = np.array(changes, dtype=np.int32).reshape(-1, ndim)
        out[f"c{k}_blocked"] = np.array(flags, dtype=np.uint8)
        out[f"c{k}_cost"] = np.array(steps["cost"])
        steps["path"].pack(f"c{k}_path", out, np.int32)
        steps["expand"].pack(f"c{k}_expand", out, np.int64)
        print(kind, dims, "steps", len(steps["cost"]), "costs", np.round(steps["cost"], 3),
              "visited", [len(a) for a in steps["expand"].items])
        k += 1
    out["n_cases"] = np.array(k)
    np.savez_compressed(os.path.join(mg.OUT, "incr.npz"), **out)


def main_csv(pmp):
    ref_harness.fast_grid_map(pmp)
    z = gio.load("csv_replay.npz")
    dims = tuple(int(v) for v in z["dims"])
    scen_names = [str(s) for s in z["scen_names"]]
    n = len(z["starts"])
    rows = {}
    with open(ref_harness.REFERENCE_ROOT + "/3d_pathfinding_results.csv") as f:
        for r in csv.DictReader(f):
            if r["Algorithm"] in ("lpastar", "dstar"):
                rows.setdefault((r["Scenario"], r["Algorithm"], int(r["Seed"])), []).append(r)
    out = {}
    for a, cls in (("lpastar", pmp.LPAStar3D), ("dstar", pmp.DStar3D)):
        csv_dist, csv_vis = np.zeros(n), np.zeros(n, dtype=np.int64)
        ref_cost, ref_vis, ref_plen = np.zeros(n), np.zeros(n, dtype=np.int64), np.zeros(n, dtype=np.int64)
        for k in range(n):
            occ = gio.unpack_occ(z["occ_bits"][k], dims)
            start = tuple(int(v) for v in z["starts"][k])
            goal = tuple(int(v) for v in z["goals"][k])
            rr = rows[(scen_names[int(z["scen_idx"][k])], a, int(z["seeds"][k]))]
            assert len({(r["Distance"], r["Visited Nodes"]) for r in rr}) == 1
            env = pmp.Grid3D(*dims)
            env.update(set(map(tuple, np.argwhere(occ).tolist())))
            cost, path, expand = cls(start=start, goal=goal, env=env).plan()
            csv_dist[k] = float(rr[0]["Distance"])
            csv_vis[k] = int(rr[0]["Visited Nodes"])
            ref_cost[k] = float(cost)
            ref_vis[k] = len(expand)
            ref_plen[k] = len(path)
        out[f"{a}_csv_dist"], out[f"{a}_csv_vis"] = csv_dist, csv_vis
        out[f"{a}_ref_cost"], out[f"{a}_ref_vis"], out[f"{a}_ref_plen"] = ref_cost, ref_vis, ref_plen
        same = (ref_cost == csv_dist) | (np.abs(ref_cost - csv_dist) <= 4e-16 * np.abs(csv_dist))
        print(a, "visited == csv:", int((ref_vis == csv_vis).sum()), "/", n, " distance == csv:", int(same.sum()),
              " inf distances:", int(np.isinf(csv_dist).sum()))
    np.savez_compressed(os.path.join(mg.OUT, "csv_incr.npz"), **out)


if __name__ == "__main__":
    pmp = ref_harness.load_reference()
    which = sys.argv[1:] or ["incr", "csv"]
    if "incr" in which:
        main_incr(pmp)
    if "csv" in which:
        main_csv(pmp)
