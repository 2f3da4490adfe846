"""Pins the C oracle (oracle/pmp_oracle.c) to the reference.

(1) the reference's own result file 3d_pathfinding_results.csv: 1 500 known answers
    (Distance, Visited Nodes) for astar/dijkstra/gbfs on Grid3D(21,15,11), and
(2) outputs of the unmodified reference executed in the build container
    (tools/make_golden.py): cost, path, CLOSED order with g / h / parent.
All comparisons are EXACT (==) -- the oracle restates the reference's arithmetic and heap order.
"""
import numpy as np
import pytest

from oracle import oracle
from tests import golden_io as gio


@pytest.fixture(scope="module")
def csv_replay():
    return gio.load("csv_replay.npz")


@pytest.mark.parametrize("algo_name,algo", [("astar", oracle.ASTAR), ("dijkstra", oracle.DIJKSTRA),
                                            ("gbfs", oracle.GBFS)])
def test_csv_known_answers(csv_replay, algo_name, algo):
    z = csv_replay
    dims = tuple(int(v) for v in z["dims"])
    path_r = gio.Ragged(z, f"{algo_name}_path")
    ecell_r = gio.Ragged(z, f"{algo_name}_ecell")
    eg_r = gio.Ragged(z, f"{algo_name}_eg")
    n = len(z["starts"])
    assert n == 500
    for k in range(n):
        occ = gio.unpack_occ(z["occ_bits"][k], dims)
        r = oracle.plan(occ, z["starts"][k], z["goals"][k], algo=algo)
        assert not r.escaped
        # the CSV (reference's published result file)
        assert r.n_expand == int(z[f"{algo_name}_csv_vis"][k]), (algo_name, k)
        assert r.cost == pytest.approx(float(z[f"{algo_name}_csv_dist"][k]), rel=4e-16, abs=0)
        # the live reference
        assert r.cost == float(z[f"{algo_name}_ref_cost"][k]), (algo_name, k)
        assert r.path == [tuple(int(v) for v in p) for p in path_r[k]], (algo_name, k)
        ec = ecell_r[k]
        if len(ec):
            assert np.array_equal(r.expand_cell, ec[:, 0].astype(np.int64)), (algo_name, k)
            assert np.array_equal(r.expand_g, eg_r[k][:, 0]), (algo_name, k)


def _check_cases(cases):
    for i in range(len(cases)):
        c = cases.case(i)
        motions, mg = gio.motions_for(cases.ndim, c["mset"])
        r = oracle.plan(c["occ"], c["start"], c["goal"], algo=c["algo"], heuristic=c["heur"],
                        motions=motions, motion_g=mg)
        tag = (cases.grid_names[c["gi"]], i, c["algo"], c["heur"])
        assert not r.escaped, tag
        assert r.reached == c["found"], tag
        if c["found"]:
            assert r.cost == c["cost"], tag
            assert r.path == c["path"], tag
        if cases.ndim == 2 and not c["found"]:
            continue   # the 2D planners drop CLOSED when no path exists (a_star.py:83)
        assert np.array_equal(r.expand_cell, c["e_cell"]), tag
        assert np.array_equal(r.expand_g, c["e_g"]), tag
        assert np.array_equal(r.expand_parent, c["e_parent"]), tag
        if not (cases.ndim == 2 and c["algo"] == oracle.ASTAR):
            assert np.array_equal(r.expand_h, c["e_h"]), tag
        else:
            # 2D A*: Node(start).h stays 0 in the reference (planner.py:15); compare the rest
            assert np.array_equal(r.expand_h[1:], c["e_h"][1:]), tag


def test_cases2d_exact():
    cases = gio.Cases("cases2d.npz", 2)
    assert len(cases) > 100
    _check_cases(cases)


def test_cases3d_exact():
    cases = gio.Cases("cases3d.npz", 3)
    assert len(cases) > 30
    _check_cases(cases)


def test_readme_example_values():
    """SURVEY section 6 probe values for README.md:62-81."""
    cases = gio.Cases("cases2d.npz", 2)
    c = cases.case(0)
    assert cases.grid_names[c["gi"]] == "readme" and c["algo"] == oracle.ASTAR
    r = oracle.plan(c["occ"], (5, 5), (45, 25), algo=oracle.ASTAR)
    assert (r.cost, len(r.path), r.n_expand) == (54.04163056034261, 48, 579)
    r = oracle.plan(c["occ"], (5, 5), (45, 25), algo=oracle.DIJKSTRA)
    assert (r.cost, len(r.path), r.n_expand) == (54.04163056034261, 48, 1219)
    r = oracle.plan(c["occ"], (5, 5), (45, 25), algo=oracle.GBFS)
    assert (r.cost, len(r.path), r.n_expand) == (73.254833995939, 61, 197)


def test_fields_exact():
    z = gio.load("fields.npz")
    for name, ndim in (("f3d_26", 3), ("f3d_6", 3), ("f3d_18", 3)):
        occ = gio.unpack_occ(z[name + "_bits"], z[name + "_shape"])
        motions, mg = gio.motions_for(3, int(z[name + "_mset"]))
        r = oracle.plan(occ, z[name + "_src"], None, algo=oracle.DIJKSTRA, motions=motions,
                        motion_g=mg, want_field=True)
        assert np.array_equal(r.g_field, z[name + "_g"]), name
    for name in ("f2d_a", "f2d_b"):
        occ = gio.unpack_occ(z[name + "_bits"], z[name + "_shape"])
        r = oracle.plan(occ, z[name + "_src"], z[name + "_goal"], algo=oracle.DIJKSTRA, want_field=True)
        assert r.cost == float(z[name + "_cost"])
        assert np.array_equal(r.g_field, z[name + "_g"]), name


def test_collision_rule_is_directed_in_3d():
    """with (3,2,2) blocked, (2,2,2)->(3,3,3) collides, the reverse does not."""
    occ = np.zeros((6, 6, 6), dtype=bool)
    occ[3, 2, 2] = True
    assert oracle.is_collision(occ, (2, 2, 2), (3, 3, 3))
    assert not oracle.is_collision(occ, (3, 3, 3), (2, 2, 2))
    occ2 = np.zeros((5, 5), dtype=bool)
    occ2[2, 1] = True
    assert oracle.is_collision(occ2, (1, 1), (2, 2)) and oracle.is_collision(occ2, (2, 2), (1, 1))


def test_batch_matches_single():
    cases = gio.Cases("cases2d.npz", 2)
    occ = cases.grids[2]
    free = np.argwhere(~occ)
    rng = np.random.default_rng(5)
    idx = rng.choice(len(free), size=16, replace=False)
    starts, goals = free[idx[:8]], free[idx[8:]]
    cost, nexp, plen, flags = oracle.plan_batch(occ, starts, goals, algo=oracle.DIJKSTRA, n_threads=4)
    for q in range(8):
        r = oracle.plan(occ, starts[q], goals[q], algo=oracle.DIJKSTRA)
        assert r.cost == cost[q] and r.n_expand == nexp[q] and len(r.path) == plen[q]


def test_field64_config5_crop():
    """config 5 at a multi-tile size: the oracle's 6-motion Dijkstra field == the unmodified
    Dijkstra3D's on a 64^3 crop of the obstacle block (tools/make_golden.py field64)"""
    z = gio.load("field64.npz") if hasattr(gio, "load") else np.load(os.path.join(os.path.dirname(__file__), "golden", "field64.npz"))
    occ, src = z["occ"], tuple(int(v) for v in z["src"])
    keep = oracle.motion_costs(oracle.MOTIONS_3D) == 1.0
    r = oracle.plan(occ, src, None, algo=oracle.DIJKSTRA, motions=oracle.MOTIONS_3D[keep],
                    motion_g=oracle.motion_costs(oracle.MOTIONS_3D)[keep], want_field=True, want_expand=False)
    assert np.array_equal(r.g_field, z["g"])
    assert int(np.isfinite(z["g"]).sum()) == 222454
