"""Golden vectors for incremental replanning, written from the
UNMODIFIED reference: LPAStar (2D) and LPAStar3D run on seeded maps, a sequence of obstacle
toggles applied through the reference's own repair entry points (LPAStar3D.apply_change,
lpa_star3d.py:98-133; 2D: the body of LPAStar.OnPress, lpa_star.py:112-127), cost and path
recorded after every change.  Runs only in the build container (needs /root/reference):

    python tools/make_golden_replan.py        ->  tests/golden/replan.npz
"""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tools"))
import ref_harness  # noqa: E402

OUT = os.path.join(ROOT, "tests", "golden")


def pack(occ):
    """bool [x,y(,z)] -> little-endian bits in z-major (x fastest) order (tests/golden_io.unpack_occ)"""
    return np.packbits(np.ascontiguousarray(occ.T).reshape(-1), bitorder="little")


def occupancy(env, shape):
    occ = np.zeros(shape, dtype=bool)
    for p in env.obstacles:
        if all(0 <= c < s for c, s in zip(p, shape)):
            occ[p] = True
    return occ


def toggle_2d(planner, xy):
    """lpa_star.py:112-127 without the plot calls"""
    planner.EXPAND = []
    node_change = planner.map[xy]
    if xy not in planner.obstacles:
        planner.obstacles.add(xy)
    else:
        planner.obstacles.remove(xy)
        planner.updateVertex(node_change)
    planner.env.update(planner.obstacles)
    for node_n in planner.getNeighbor(node_change):
        planner.updateVertex(node_n)
    return planner.plan()


def pick_changes(rng, path, occ, start, goal, n, ndim):
    """half of the toggles block a cell of the current path, the rest hit random interior cells"""
    out = []
    shape = occ.shape
    for i in range(n):
        if path and len(path) > 4 and i % 2 == 0:
            c = tuple(path[int(rng.integers(2, len(path) - 2))])
        else:
            c = tuple(int(rng.integers(1, s - 1)) for s in shape)
        if c == tuple(start) or c == tuple(goal):
            continue
        out.append(c)
    return out


def main():
    pmp = ref_harness.load_reference()
    rng = np.random.default_rng(7)
    out = {}
    cases = []
    # ---- 2D ----
    for ci, (dims, density, n_changes) in enumerate([((31, 21), 0.12, 14), ((41, 31), 0.2, 16), ((24, 24), 0.05, 10)]):
        env = pmp.Grid(*dims)
        obs = env.obstacles
        start, goal = (2, 2), (dims[0] - 3, dims[1] - 3)
        for x in range(1, dims[0] - 1):
            for y in range(1, dims[1] - 1):
                if rng.random() < density and (x, y) not in (start, goal):
                    obs.add((x, y))
        env.update(obs)
        planner = pmp.LPAStar(start=start, goal=goal, env=env)
        cases.append(("2d", dims, env, planner, start, goal, n_changes))
    # ---- 3D ----
    for dims, density, n_changes in [((14, 12, 10), 0.1, 12), ((18, 10, 9), 0.18, 12)]:
        env = pmp.Grid3D(*dims)
        obs = env.obstacles
        start, goal = (1, 1, 1), (dims[0] - 2, dims[1] - 2, dims[2] - 2)
        for x in range(1, dims[0] - 1):
            for y in range(1, dims[1] - 1):
                for z in range(1, dims[2] - 1):
                    if rng.random() < density and (x, y, z) not in (start, goal):
                        obs.add((x, y, z))
        env.update(obs)
        planner = pmp.LPAStar3D(start=start, goal=goal, env=env)
        cases.append(("3d", dims, env, planner, start, goal, n_changes))

    for k, (kind, dims, env, planner, start, goal, n_changes) in enumerate(cases):
        ndim = len(dims)
        occ0 = occupancy(env, dims)
        try:
            cost, path, _ = planner.plan()
        except ValueError:           # min() of an empty neighbour list: goal unreachable (lpa_star.py:222)
            cost, path = float("inf"), []
        costs = [cost if path else float("inf")]
        changes, states = [], []
        for c in pick_changes(rng, path, occ0, start, goal, n_changes, ndim):
            now_blocked = c not in planner.obstacles
            try:
                if kind == "2d":
                    cost, path, _ = toggle_2d(planner, c)
                else:
                    cost, path, _ = planner.apply_change(c)
            except ValueError:
                cost, path = float("inf"), []
            changes.append(c)
            states.append(now_blocked)
            costs.append(cost if path else float("inf"))
        out[f"c{k}_shape"] = np.array(dims, dtype=np.int32)
        out[f"c{k}_bits"] = pack(occ0)
        out[f"c{k}_start"] = np.array(start, dtype=np.int32)
        out[f"c{k}_goal"] = np.array(goal, dtype=np.int32)
        out[f"c{k}_changes"] = np.array(changes, dtype=np.int32).reshape(-1, ndim)
        out[f"c{k}_blocked"] = np.array(states, dtype=np.uint8)
        out[f"c{k}_ref_cost"] = np.array(costs, dtype=np.float64)
        print(kind, dims, "changes", len(changes), "costs", np.round(costs, 3))
    out["n_cases"] = np.array(len(cases))
    np.savez_compressed(os.path.join(OUT, "replan.npz"), **out)
    print("wrote", os.path.join(OUT, "replan.npz"))


if __name__ == "__main__":
    main()
