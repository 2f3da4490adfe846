"""CPU-side checks of the drop-in boundary: the C-ABI library loads, exports every symbol
include/pmp_b200.h declares, and the Python surface mirrors the reference's classes.  No
compute call is made (no GPU here)."""
import ctypes
import inspect
import os
import re

import pytest

import python_motion_planning_b200 as pmp
from python_motion_planning_b200 import _lib

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def declared_symbols():
    text = open(os.path.join(ROOT, "include", "pmp_b200.h")).read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return sorted(set(re.findall(r"\b(pmp_[a-z0-9_]+)\s*\(", text)))


def test_library_builds_and_exports_every_declared_symbol():
    _lib.build()
    L = ctypes.CDLL(_lib.LIB_PATH)
    decl = declared_symbols()
    assert len(decl) >= 20
    for name in decl:
        assert hasattr(L, name), f"{name} declared in include/pmp_b200.h but not exported"
    assert sorted(_lib.SYMBOLS) == decl
    L.pmp_abi_version.restype = ctypes.c_int
    assert L.pmp_abi_version() == _lib.ABI_VERSION == 3


def test_no_cpu_fallback():
    import torch
    if torch.cuda.is_available():
        pytest.skip("GPU present")
    env = pmp.Grid(20, 20)
    with pytest.raises(pmp.PmpError):
        pmp.AStar((2, 2), (10, 10), env).plan()
    with pytest.raises(pmp.PmpError):
        pmp.plan_batch(env, [(2, 2)], [(5, 5)])


def test_product_never_imports_oracle():
    pkg = os.path.join(ROOT, "python_motion_planning_b200")
    for dirpath, _, files in os.walk(pkg):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".h")):
                src = open(os.path.join(dirpath, f)).read()
                assert "oracle" not in src.lower().replace("# oracle", ""), f


def test_python_surface_matches_reference_contract():
    # constructor signatures (planner.py:13, graph_search.py:21, dijkstra3d.py:29)
    def positional(fn):   # the reference's parameters; this package only adds keyword-only options
        return [n for n, p in inspect.signature(fn).parameters.items() if p.kind != p.KEYWORD_ONLY]
    assert positional(pmp.AStar.__init__) == ["self", "start", "goal", "env", "heuristic_type"]
    assert positional(pmp.Dijkstra3D.__init__) == ["self", "start", "goal", "env"]
    assert positional(pmp.ThetaStar3D.__init__) == ["self", "start", "goal", "env", "heuristic_type"]
    env = pmp.Grid(51, 31)
    assert len(env.motions) == 8 and len(env.obstacles) == 2 * 51 + 2 * 31 - 4
    assert [m.current for m in env.motions][:3] == [(-1, 0), (-1, 1), (0, 1)]
    assert (0, 0) in env.grid_map and len(env.grid_map) == 51 * 31
    e3 = pmp.Grid3D(6, 5, 4)
    assert len(e3.motions) == 26
    assert {z for _, _, z in e3.obstacles} == {-1, 0, 1, 2, 3}      # env3d.py:86-92 quirk
    assert (1, 1, 1) in e3.grid_map and (6, 0, 0) not in e3.grid_map
    names = {pmp.AStar: "A*", pmp.Dijkstra: "Dijkstra", pmp.GBFS: "Greedy Best First Search(GBFS)"}
    for cls, s in names.items():
        assert str(cls((5, 5), (45, 25), env)) == s
    assert str(pmp.Dijkstra3D((1, 1, 1), (3, 3, 2), e3)) == "Dijkstra (3D)"
    assert str(pmp.GBFS3D((1, 1, 1), (3, 3, 2), e3)) == "Greedy Best First Search (GBFS) 3D"
    p = pmp.SearchFactory()("a_star", start=(5, 5), goal=(45, 25), env=env)
    assert isinstance(p, pmp.AStar) and p.start.current == (5, 5) and p.motions is env.motions
    with pytest.raises(ValueError):
        pmp.SearchFactory()("nope")
    # Node arithmetic / ordering (node.py:39-54)
    n = pmp.Node((1, 0), (2, 3), 1, 2) + pmp.Node((1, 0), None, 2, 8)
    assert n.current == (2, 0) and n.parent == (2, 3) and n.g == 3 and n.h == 2
    assert pmp.Node((0, 0), None, 1, 1) < pmp.Node((1, 1), None, 1, 2)
    # isCollision restated on the host set agrees with the reference rule (directed in 3D)
    e3.obstacles.add((3, 2, 2))
    a = pmp.AStar3D((1, 1, 1), (4, 3, 2), e3)
    assert a.isCollision(pmp.Node3D((2, 2, 2)), pmp.Node3D((3, 3, 3)))
    assert not a.isCollision(pmp.Node3D((3, 3, 3)), pmp.Node3D((2, 2, 2))) or (3, 3, 3) in e3.obstacles


def test_dense_update_and_lazy_obstacles():
    import numpy as np
    env = pmp.Grid(12, 9)
    occ = np.zeros((12, 9), dtype=bool)
    occ[0, :] = occ[-1, :] = True
    occ[:, 0] = occ[:, -1] = True
    occ[5, 4] = True
    env.update(occ)
    assert (5, 4) in env.obstacles and len(env.obstacles) == int(occ.sum())
    assert np.array_equal(env.occupancy(), occ)


def test_reference_import_paths():
    """the reference's import paths for this path's names (utils/__init__.py, global_planner/graph_search/__init__.py)"""
    from python_motion_planning_b200.utils import Grid, Grid3D, Node, SearchFactory, Planner
    from python_motion_planning_b200.global_planner.graph_search import (AStar, AStar3D, Dijkstra3D, GBFS3D, JPS3D,
                                                                         LazyThetaStar3D, ThetaStar3D, LPAStar3D,
                                                                         DStar3D, DStarLite)
    from python_motion_planning_b200.global_planner import Dijkstra
    assert Grid is pmp.Grid and Grid3D is pmp.Grid3D and Node is pmp.Node and SearchFactory is pmp.SearchFactory
    assert AStar is pmp.AStar and AStar3D is pmp.AStar3D and Dijkstra is pmp.Dijkstra and Planner is pmp.Planner
    assert DStarLite is pmp.DStarLite and ThetaStar3D is pmp.ThetaStar3D and LPAStar3D is pmp.LPAStar3D
    assert all(c is not None for c in (Dijkstra3D, GBFS3D, JPS3D, LazyThetaStar3D, DStar3D))
