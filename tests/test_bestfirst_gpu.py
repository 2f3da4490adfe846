"""GPU parity of the reference-order mode (pmp_bestfirst; planners with reference_order=True and the
Theta* classes in mode="reference"), through the Python planner API -> ctypes -> C ABI.

The bar is EXACT equality with the unmodified reference on the committed golden vectors: cost, the
path itself (not just its cost), len(expand), and every expand record (current, parent, g, h) in
CLOSED order; plus the Distance / Visited Nodes columns of the reference's own
3d_pathfinding_results.csv for astar, dijkstra, gbfs, theta_star and lazy_theta_star.
"""
import numpy as np
import pytest

import python_motion_planning_b200 as pmp
from tests import golden_io as gio
from tests.test_gpu_parity import make_env

pytestmark = pytest.mark.gpu

CLS = {2: {0: pmp.AStar, 1: pmp.Dijkstra, 2: pmp.GBFS, 3: pmp.ThetaStar, 4: pmp.LazyThetaStar, 5: pmp.JPS},
       3: {0: pmp.AStar3D, 1: pmp.Dijkstra3D, 2: pmp.GBFS3D, 3: pmp.ThetaStar3D, 4: pmp.LazyThetaStar3D,
           5: pmp.JPS3D}}


def make_planner(ndim, algo, start, goal, env, heur="euclidean"):
    cls = CLS[ndim][algo]
    if algo >= 3:
        return cls(start, goal, env, heur, mode="reference")
    if cls is pmp.Dijkstra3D:
        return cls(start, goal, env, reference_order=True)
    return cls(start, goal, env, heur, reference_order=True)


def flat(coords, shape):
    c = np.asarray(coords, dtype=np.int64).reshape(-1, len(shape))
    if len(shape) == 2:
        return c[:, 1] * shape[0] + c[:, 0]
    return (c[:, 2] * shape[1] + c[:, 1]) * shape[0] + c[:, 0]


@pytest.mark.parametrize("fname,ndim", [("cases2d.npz", 2), ("cases3d.npz", 3),
                                        ("anyangle2d.npz", 2), ("anyangle3d.npz", 3),
                                        ("jpscases2d.npz", 2), ("jpscases3d.npz", 3)])
def test_reference_order_equals_reference(fname, ndim):
    cases = gio.Cases(fname, ndim)
    n_found = 0
    for i in range(len(cases)):
        c = cases.case(i)
        env = make_env(c["occ"], c["mset"])
        heur = "manhattan" if c["heur"] == 1 else "euclidean"
        cost, path, expand = make_planner(ndim, c["algo"], c["start"], c["goal"], env, heur).plan()
        tag = (cases.grid_names[c["gi"]], i, c["algo"], c["heur"], c["start"], c["goal"])
        if not c["found"]:
            if ndim == 2:
                assert (cost, path, expand) == ([], [], []), tag       # a_star.py:83
                continue
            assert cost == float("inf") and path == [], tag            # a_star3d.py:77-78
        else:
            assert cost == c["cost"], tag
            assert path == c["path"], tag
            n_found += 1
        shape = c["occ"].shape
        assert len(expand) == len(c["e_cell"]), tag
        nodes = list(expand)
        assert np.array_equal(flat([n.current for n in nodes], shape), c["e_cell"]), tag
        assert np.array_equal(flat([n.parent for n in nodes], shape), c["e_parent"]), tag
        assert np.array_equal(np.array([n.g for n in nodes]), c["e_g"]), tag
        assert np.array_equal(np.array([n.h for n in nodes]), c["e_h"]), tag
    assert n_found >= 25


@pytest.mark.parametrize("name,algo,fname,step", [("astar", 0, "csv_replay.npz", 1), ("dijkstra", 1, "csv_replay.npz", 4),
                                                  ("gbfs", 2, "csv_replay.npz", 2),
                                                  ("theta_star", 3, "csv_anyangle.npz", 1),
                                                  ("lazy_theta_star", 4, "csv_anyangle.npz", 1)])
def test_csv_known_answers_reference_order(name, algo, fname, step):
    """Distance AND Visited Nodes of the reference's published result file"""
    z = gio.load("csv_replay.npz")
    a = gio.load(fname)
    dims = tuple(int(v) for v in z["dims"])
    n = 0
    for k in range(0, len(z["starts"]), step):
        occ = gio.unpack_occ(z["occ_bits"][k], dims)
        s, g = tuple(int(v) for v in z["starts"][k]), tuple(int(v) for v in z["goals"][k])
        cost, path, expand = make_planner(3, algo, s, g, make_env(occ)).plan()
        assert len(expand) == int(a[f"{name}_csv_vis"][k]), (name, k)
        assert cost == pytest.approx(float(a[f"{name}_csv_dist"][k]), rel=4e-16, abs=0), (name, k)
        assert cost == float(a[f"{name}_ref_cost"][k]), (name, k)
        assert path[0] == s and path[-1] == g
        n += 1
    assert n >= 125


def test_reference_order_switches():
    """constructor keyword, attribute and module default select the same mode; the default stays
    the wavefront"""
    c = gio.Cases("cases2d.npz", 2).case(0)
    env = make_env(c["occ"])
    assert pmp.AStar(c["start"], c["goal"], env).reference_order is False
    p = pmp.AStar(c["start"], c["goal"], env)
    p.reference_order = True
    cost, path, expand = p.plan()
    assert cost == c["cost"] and path == c["path"] and len(expand) == len(c["e_cell"])
    pmp.set_reference_order(True)
    try:
        q = pmp.Dijkstra(c["start"], c["goal"], env)
        assert q.reference_order is True
    finally:
        pmp.set_reference_order(False)
    wave_cost = pmp.AStar(c["start"], c["goal"], env).plan()[0]
    assert wave_cost == pytest.approx(cost, rel=1e-9)


def test_reference_order_batch_equals_single_queries():
    """plan_batch_reference: all queries of a (grid, algorithm) group in one launch, one warp each"""
    checked = 0
    for fname, ndim in (("cases2d.npz", 2), ("cases3d.npz", 3), ("anyangle2d.npz", 2), ("jpscases3d.npz", 3)):
        cases = gio.Cases(fname, ndim)
        groups = {}
        for i in range(len(cases)):
            c = cases.case(i)
            groups.setdefault((c["gi"], c["algo"], c["heur"], c["mset"]), []).append(c)
        names = {0: "a_star", 1: "dijkstra", 2: "gbfs", 3: "theta_star", 4: "lazy_theta_star", 5: "jps"}
        for (gi, algo, heur, mset), cs in groups.items():
            if len(cs) < 2:
                continue
            env = make_env(cs[0]["occ"], mset)
            res = pmp.plan_batch_reference(env, [c["start"] for c in cs], [c["goal"] for c in cs], algo=names[algo],
                                           heuristic_type="manhattan" if heur == 1 else "euclidean")
            for q, c in enumerate(cs):
                tag = (fname, gi, algo, q)
                assert int(res.n_expand[q]) == len(c["e_cell"]) or (ndim == 2 and not c["found"]), tag
                if c["found"]:
                    assert res.costs[q] == c["cost"], tag
                    want = c["path"] if ndim == 2 else c["path"][::-1]      # batch paths are goal -> start
                    assert [tuple(int(v) for v in p) for p in res.paths[q]] == want, tag
                else:
                    assert np.isinf(res.costs[q]) and res.path_len[q] == 0, tag
                checked += 1
    assert checked >= 100


@pytest.mark.parametrize("ndim", [2, 3])
def test_reference_order_equals_oracle_on_random_maps(ndim):
    """the kernel against the pinned oracle on seeded random maps (4/8- and 6/18/26-connected, both
    heuristics, unreachable goals): cost, path and len(CLOSED) of every query, batched per map"""
    from oracle import oracle
    from tests.test_bestfirst_core import random_cases
    names = {0: "a_star", 1: "dijkstra", 2: "gbfs"}
    n = 0
    for occ, mset, heur, pairs in random_cases(11 + ndim, 60, ndim):
        motions, mg = gio.motions_for(ndim, mset)
        env = make_env(occ, mset)
        for algo in (0, 1, 2):
            res = pmp.plan_batch_reference(env, [s for s, _ in pairs], [g for _, g in pairs], algo=names[algo],
                                           heuristic_type="manhattan" if heur == 1 else "euclidean")
            for q, (s, g) in enumerate(pairs):
                want = oracle.plan(occ, s, g, algo=algo, heuristic=heur, motions=motions, motion_g=mg)
                tag = (occ.shape, mset, heur, algo, s, g)
                assert int(res.n_expand[q]) == want.n_expand, tag
                if want.reached:
                    assert res.costs[q] == want.cost, tag
                    path = [tuple(int(v) for v in p) for p in res.paths[q]]           # goal -> start
                    assert path == (want.path if ndim == 2 else want.path[::-1]), tag
                else:
                    assert np.isinf(res.costs[q]) and res.path_len[q] == 0, tag
                n += 1
    assert n >= 600
