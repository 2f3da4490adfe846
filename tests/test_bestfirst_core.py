"""CPU check of the sequential core of the pmp_bestfirst kernel (csrc/pmp_bestfirst_core.h).

tests/bestfirst_host.cpp compiles that header with g++ and drives it exactly as k_bestfirst does
(lane 0: pop / push / close, one bf_eval per motion).  Here its outputs are compared, EXACTLY, with
outputs of the unmodified reference committed under tests/golden/ -- A*, Dijkstra, GBFS
(cases2d / cases3d / the reference's own CSV) and Theta*, Lazy Theta* (anyangle2d / anyangle3d /
the CSV's theta_star and lazy_theta_star rows) and JPS / JPS3D (jpscases2d / jpscases3d; jump functions
in csrc/pmp_jps_core.h): cost, path, CLOSED order, g, h, parent.
The kernel itself is checked the same way on the GPU (tests/test_bestfirst_gpu.py).
"""
import ctypes
import os
import subprocess

import numpy as np
import pytest

from oracle import oracle
from tests import golden_io as gio

HERE = os.path.dirname(os.path.abspath(__file__))
SRC = os.path.join(HERE, "bestfirst_host.cpp")
CORE = os.path.join(HERE, "..", "python_motion_planning_b200", "csrc", "pmp_bestfirst_core.h")
SO = os.path.join(HERE, "libbestfirst_host.so")

THETA, LAZY = 3, 4


@pytest.fixture(scope="module")
def host():
    newest = max(os.path.getmtime(f) for f in (SRC, CORE, os.path.join(os.path.dirname(CORE), "pmp_jps_core.h"),
                                             os.path.join(os.path.dirname(CORE), "pmp_incr_core.h")))
    if not os.path.exists(SO) or os.path.getmtime(SO) < newest:
        subprocess.run(["g++", "-O2", "-std=c++17", "-fPIC", "-Wall", "-shared", "-o", SO, SRC], check=True)
    L = ctypes.CDLL(SO)
    vp = ctypes.c_void_p
    L.bf_host_plan.restype = ctypes.c_int
    L.bf_host_plan.argtypes = [ctypes.c_int, vp, vp, ctypes.c_int, vp, vp, ctypes.c_int, ctypes.c_int,
                               vp, vp, vp, vp, vp, ctypes.c_int64, vp, vp, vp, vp, vp, ctypes.c_int64,
                               vp, vp]
    return L


def plan(L, occ, start, goal, algo, heuristic, motions=None, motion_g=None):
    ndim, dims, occ_c, motions, motion_g = oracle._prep(occ, motions, motion_g)
    # what one step adds to g: motion.g in 2D (node.py:41), the offset's length in 3D (a_star3d.py:67)
    w = motion_g if ndim == 2 else oracle.motion_costs(motions)
    cells = int(occ_c.size)
    cost, plen, nexp = ctypes.c_double(0), ctypes.c_int64(0), ctypes.c_int64(0)
    reached = ctypes.c_int32(0)
    path = np.zeros((cells, 3), dtype=np.int32)
    e_cell = np.zeros(cells, dtype=np.int64)
    e_g, e_h = np.zeros(cells), np.zeros(cells)
    e_p = np.zeros(cells, dtype=np.int64)
    p = oracle._p
    los = ctypes.cast(oracle.lib().pmp_oracle_line_of_sight, ctypes.c_void_p)
    rc = L.bf_host_plan(ndim, p(dims), p(occ_c), len(motions), p(motions), p(np.ascontiguousarray(w)),
                        algo, heuristic, p(oracle._c3(start)), p(oracle._c3(goal)), los,
                        ctypes.byref(cost), p(path), cells, ctypes.byref(plen), p(e_cell), p(e_g),
                        p(e_h), p(e_p), cells, ctypes.byref(nexp), ctypes.byref(reached))
    assert rc == 0, rc
    pts = [tuple(int(v) for v in row[:ndim]) for row in path[:plen.value]]   # goal first
    if ndim == 3:
        pts.reverse()                                                           # a_star3d.py:105
    n = nexp.value
    return dict(cost=cost.value, path=pts, reached=bool(reached.value), e_cell=e_cell[:n], e_g=e_g[:n],
                e_h=e_h[:n], e_parent=e_p[:n])


def check_cases(L, cases):
    for i in range(len(cases)):
        c = cases.case(i)
        motions, mg = gio.motions_for(cases.ndim, c["mset"])
        r = plan(L, c["occ"], c["start"], c["goal"], c["algo"], c["heur"], motions, mg)
        tag = (cases.grid_names[c["gi"]], i, c["algo"], c["heur"])
        assert r["reached"] == c["found"], tag
        if c["found"]:
            assert r["cost"] == c["cost"], tag
            assert r["path"] == c["path"], tag
        if cases.ndim == 2 and not c["found"]:
            continue   # the 2D planners return ([], [], []) (a_star.py:83)
        assert np.array_equal(r["e_cell"], c["e_cell"]), tag
        assert np.array_equal(r["e_g"], c["e_g"]), tag
        assert np.array_equal(r["e_h"], c["e_h"]), tag
        assert np.array_equal(r["e_parent"], c["e_parent"]), tag


@pytest.mark.parametrize("fname,ndim", [("cases2d.npz", 2), ("cases3d.npz", 3),
                                        ("anyangle2d.npz", 2), ("anyangle3d.npz", 3),
                                        ("jpscases2d.npz", 2), ("jpscases3d.npz", 3)])
def test_core_equals_reference(host, fname, ndim):
    check_cases(host, gio.Cases(fname, ndim))


@pytest.mark.parametrize("name,algo,fname", [("astar", 0, "csv_replay.npz"), ("gbfs", 2, "csv_replay.npz"),
                                             ("theta_star", THETA, "csv_anyangle.npz"),
                                             ("lazy_theta_star", LAZY, "csv_anyangle.npz")])
def test_core_csv_known_answers(host, name, algo, fname):
    """Distance and Visited Nodes of the reference's 3d_pathfinding_results.csv (every 5th instance)"""
    z = gio.load("csv_replay.npz")
    a = gio.load(fname)
    dims = tuple(int(v) for v in z["dims"])
    for k in range(0, len(z["starts"]), 5):
        occ = gio.unpack_occ(z["occ_bits"][k], dims)
        r = plan(host, occ, z["starts"][k], z["goals"][k], algo, 0)
        assert len(r["e_cell"]) == int(a[f"{name}_csv_vis"][k]), (name, k)
        assert r["cost"] == pytest.approx(float(a[f"{name}_csv_dist"][k]), rel=4e-16, abs=0), (name, k)
        assert r["cost"] == float(a[f"{name}_ref_cost"][k]), (name, k)


def random_cases(seed, n_grids, ndim):
    """seeded random walled maps with a motion subset, a heuristic and a few free start / goal pairs"""
    rng = np.random.default_rng(seed)
    for _ in range(n_grids):
        shape = tuple(int(v) for v in (rng.integers(9, 28, size=2) if ndim == 2 else rng.integers(6, 13, size=3)))
        occ = rng.random(shape) < rng.choice([0.1, 0.2, 0.3])
        for ax in range(ndim):
            idx = [slice(None)] * ndim
            idx[ax] = 0
            occ[tuple(idx)] = True
            idx[ax] = -1
            occ[tuple(idx)] = True
        free = np.argwhere(~occ)
        if len(free) < 4:
            continue
        mset = int(rng.choice([4, 8]) if ndim == 2 else rng.choice([6, 18, 26]))
        heur = int(rng.integers(0, 2))
        pairs = [(tuple(int(v) for v in free[i]), tuple(int(v) for v in free[j]))
                 for i, j in rng.integers(0, len(free), size=(4, 2))]
        yield occ, mset, heur, pairs


@pytest.mark.parametrize("ndim", [2, 3])
def test_core_equals_oracle_on_random_maps(host, ndim):
    """two independent restatements of the reference (the oracle's C, pinned to the reference, and the
    kernel's core) agree exactly on seeded random cases: 4/8- and 6/18/26-connected, both heuristics,
    unreachable goals included"""
    n = 0
    for occ, mset, heur, pairs in random_cases(11 + ndim, 60, ndim):
        motions, mg = gio.motions_for(ndim, mset)
        for algo in (0, 1, 2):
            for s, g in pairs:
                want = oracle.plan(occ, s, g, algo=algo, heuristic=heur, motions=motions, motion_g=mg)
                got = plan(host, occ, s, g, algo, heur, motions, mg)
                tag = (occ.shape, mset, heur, algo, s, g)
                assert got["reached"] == want.reached, tag
                if want.reached:
                    assert got["cost"] == want.cost and got["path"] == want.path, tag
                assert np.array_equal(got["e_cell"], want.expand_cell), tag
                assert np.array_equal(got["e_g"], want.expand_g) and np.array_equal(got["e_h"], want.expand_h), tag
                assert np.array_equal(got["e_parent"], want.expand_parent), tag
                n += 1
    assert n >= 600
