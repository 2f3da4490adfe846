"""Any-angle row: lineOfSight parity and path shortcutting.

lineOfSight is a predicate, so the bar is exact equality: the oracle's restatement against the
committed answers of the unmodified reference (tests/golden/los.npz, tools/make_golden_los.py),
the device kernel against both.  Theta* proper (mode="reference") is tested in
tests/test_bestfirst_gpu.py.  The shortcut kernel (mode="shortcut", for large maps) has no reference
counterpart (the reference shortcuts inside the search); it is checked against a plain-Python statement of the same greedy
rule driven by the oracle's lineOfSight (waypoints identical), by properties (every segment
visible under the reference predicate, cost = sum of segment lengths <= grid A* cost), and its
cost is compared with the Theta* column of the reference's own CSV.
"""
import math

import numpy as np
import pytest

from oracle import oracle
from tests import golden_io as gio

REL = 1e-9


def los_grids():
    z = gio.load("los.npz")
    for gi in range(int(z["n_grids"])):
        shape = tuple(int(v) for v in z[f"g{gi}_shape"])
        yield gio.unpack_occ(z[f"g{gi}_bits"], shape), z[f"g{gi}_a"], z[f"g{gi}_b"], z[f"g{gi}_los"].astype(bool)


def test_oracle_line_of_sight_matches_reference():
    n = 0
    for occ, a, b, ref in los_grids():
        got = oracle.line_of_sight(occ, a, b)
        assert np.array_equal(got, ref)
        n += len(ref)
        assert ref.any() and not ref.all()
    assert n >= 12000


def make_env(pmp, occ):
    env = pmp.Grid(*occ.shape) if occ.ndim == 2 else pmp.Grid3D(*occ.shape, walls=False)
    env.update(set(map(tuple, np.argwhere(occ).tolist())))
    return env


def greedy_shortcut(occ, path_goal_first):
    """the rule of k_shortcut_paths, stated with the oracle's lineOfSight"""
    p = [tuple(c) for c in path_goal_first]
    out, i, cost = [p[0]], 0, 0.0
    while i < len(p) - 1:
        j = i + 1
        while j + 1 < len(p) and oracle.line_of_sight(occ, [p[i]], [p[j + 1]])[0]:
            j += 1
        cost += math.sqrt(sum((a - b) ** 2 for a, b in zip(p[i], p[j])))
        out.append(p[j])
        i = j
    return out, cost


def seg_cost(path):
    c = 0.0
    for a, b in zip(path[:-1], path[1:]):
        c += math.sqrt(sum((x - y) ** 2 for x, y in zip(a, b)))
    return c


@pytest.mark.gpu
def test_device_line_of_sight_matches_reference_and_oracle():
    import python_motion_planning_b200 as pmp
    for occ, a, b, ref in los_grids():
        env = make_env(pmp, occ)
        got = pmp.line_of_sight(env, a, b)
        assert np.array_equal(got, ref)
    # a larger grid, long rays: device vs oracle
    rng = np.random.default_rng(2)
    occ = np.zeros((97, 61, 45), dtype=bool)
    for _ in range(160):
        lo = [int(rng.integers(0, s)) for s in occ.shape]
        e = rng.integers(1, 7, size=3)
        occ[lo[0]:lo[0] + e[0], lo[1]:lo[1] + e[1], lo[2]:lo[2] + e[2]] = True
    a = np.stack([rng.integers(-1, s + 1, size=30000) for s in occ.shape], axis=1)
    b = np.stack([rng.integers(-1, s + 1, size=30000) for s in occ.shape], axis=1)
    env = make_env(pmp, occ)
    got = pmp.line_of_sight(env, a, b)
    inside = np.all((a >= 0) & (a < occ.shape) & (b >= 0) & (b < occ.shape), axis=1)
    assert not got[~inside].any()                       # outside the box: blocked
    ref = oracle.line_of_sight(occ, a[inside], b[inside])
    assert np.array_equal(got[inside], ref)
    assert 0.02 < ref.mean() < 0.98
    occ2 = rng.random((300, 200)) < 0.03
    a = np.stack([rng.integers(0, s, size=30000) for s in occ2.shape], axis=1)
    b = np.stack([rng.integers(0, s, size=30000) for s in occ2.shape], axis=1)
    got = pmp.line_of_sight(make_env(pmp, occ2), a, b)
    assert np.array_equal(got, oracle.line_of_sight(occ2, a, b))
    assert got.any() and not got.all()


@pytest.mark.gpu
def test_theta_star_3d_on_benchmark_scenarios():
    import python_motion_planning_b200 as pmp
    z = gio.load("csv_replay.npz")
    lz = gio.load("los.npz")
    dims = tuple(int(v) for v in z["dims"])
    ratios = []
    for k in range(0, len(z["starts"]), 5):             # 100 of the 500 benchmark instances
        occ = gio.unpack_occ(z["occ_bits"][k], dims)
        s, g = tuple(int(v) for v in z["starts"][k]), tuple(int(v) for v in z["goals"][k])
        env = make_env(pmp, occ)
        cost, path, expand = pmp.ThetaStar3D(s, g, env, mode="shortcut").plan()
        a_cost, a_path, _ = pmp.AStar3D(s, g, env).plan()
        assert path[0] == s and path[-1] == g
        # every segment passes the reference predicate in the direction updateVertex uses it
        assert oracle.line_of_sight(occ, path[1:], path[:-1]).all()
        assert abs(cost - seg_cost(path[::-1])) <= 1e-12 * max(cost, 1.0)
        assert cost <= a_cost * (1 + REL)
        want_path, want_cost = greedy_shortcut(occ, a_path[::-1])
        assert path[::-1] == want_path and cost == want_cost
        assert len(expand) > 0
        theta = float(lz["theta_star_csv_dist"][k])
        ratios.append(cost / theta)
        assert cost <= float(z["astar_csv_dist"][k]) * (1 + REL)
    ratios = np.array(ratios)
    # shortcutting the grid-optimal path is close to, not identical with, Theta* proper
    assert ratios.mean() < 1.03 and ratios.max() < 1.25, (ratios.mean(), ratios.max())


@pytest.mark.gpu
def test_theta_star_2d_vs_reference_costs():
    import python_motion_planning_b200 as pmp
    lz = gio.load("los.npz")
    grids = [occ for occ, _, _, _ in los_grids()][:3]
    ratios = []
    for gi, sx, sy, gx, gy, ref_cost in lz["theta2d_queries"]:
        occ = grids[int(gi)]
        s, g = (int(sx), int(sy)), (int(gx), int(gy))
        env = make_env(pmp, occ)
        res = pmp.ThetaStar(s, g, env, mode="shortcut").plan()
        if not math.isfinite(ref_cost):
            assert res == ([], [], [])
            continue
        cost, path, _ = res
        assert path[0] == g and path[-1] == s           # 2D family: goal -> start
        assert oracle.line_of_sight(occ, path[:-1], path[1:]).all()
        a_cost = pmp.AStar(s, g, env).plan()[0]
        assert cost <= a_cost * (1 + REL)
        assert abs(cost - seg_cost(path)) <= 1e-12 * max(cost, 1.0)
        ratios.append(cost / ref_cost)
    ratios = np.array(ratios)
    assert len(ratios) >= 20 and ratios.mean() < 1.05, (ratios.mean(), ratios.max())
