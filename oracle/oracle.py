"""ctypes front-end of the parity oracle (oracle/pmp_oracle.c).

TEST INFRASTRUCTURE: only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline /
`--impl reference` legs may import this module.  The product package
(python_motion_planning_b200) never does.

Parity status: PINNED (tests/test_oracle_golden.py: the reference's own
3d_pathfinding_results.csv + outputs of the unmodified reference, tests/golden/).
"""
import ctypes
import math
import os
import subprocess
from dataclasses import dataclass

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_LIB_PATH = os.path.join(_HERE, "libpmp_oracle.so")

ASTAR, DIJKSTRA, GBFS = 0, 1, 2
EUCLIDEAN, MANHATTAN = 0, 1
REACHED, ESCAPED = 1, 2

_SQRT2, _SQRT3 = math.sqrt(2), math.sqrt(3)

# env.py:52-55 -- the 8 motions of Grid, in the reference's order
MOTIONS_2D = np.array([(-1, 0, 0), (-1, 1, 0), (0, 1, 0), (1, 1, 0),
                       (1, 0, 0), (1, -1, 0), (0, -1, 0), (-1, -1, 0)], dtype=np.int32)
# env3d.py:56-70 -- the 26 motions of Grid3D, in the reference's order
MOTIONS_3D = np.array([
    (-1, 0, 0), (-1, 1, 0), (0, 1, 0), (1, 1, 0), (1, 0, 0), (1, -1, 0), (0, -1, 0), (-1, -1, 0),
    (0, 0, 1), (0, 0, -1),
    (-1, 0, 1), (-1, 1, 1), (0, 1, 1), (1, 1, 1), (1, 0, 1), (1, -1, 1), (0, -1, 1), (-1, -1, 1),
    (-1, 0, -1), (-1, 1, -1), (0, 1, -1), (1, 1, -1), (1, 0, -1), (1, -1, -1), (0, -1, -1),
    (-1, -1, -1)], dtype=np.int32)


def motion_costs(motions):
    """sqrt(number of non-zero axes): the 1 / sqrt(2) / sqrt(3) table of env.py / env3d.py."""
    lut = {1: 1.0, 2: _SQRT2, 3: _SQRT3}
    return np.array([lut[int(np.count_nonzero(m))] for m in motions], dtype=np.float64)


def build(force=False):
    """Compile oracle/libpmp_oracle.so with the committed Makefile (gcc)."""
    src = os.path.join(_HERE, "pmp_oracle.c")
    if (not force and os.path.exists(_LIB_PATH)
            and os.path.getmtime(_LIB_PATH) >= os.path.getmtime(src)):
        return _LIB_PATH
    subprocess.run(["make", "-C", _HERE, "-B", "libpmp_oracle.so"], check=True,
                   stdout=subprocess.DEVNULL)
    return _LIB_PATH


_lib = None


def lib():
    global _lib
    if _lib is None:
        build()
        L = ctypes.CDLL(_LIB_PATH)
        vp = ctypes.c_void_p
        L.pmp_oracle_plan.restype = ctypes.c_int
        L.pmp_oracle_plan.argtypes = [ctypes.c_int, vp, vp, ctypes.c_int, vp, vp, ctypes.c_int,
                                      ctypes.c_int, vp, vp, vp, vp, ctypes.c_int64, vp,
                                      vp, vp, vp, vp, ctypes.c_int64, vp, vp, vp]
        L.pmp_oracle_plan_batch.restype = ctypes.c_int
        L.pmp_oracle_plan_batch.argtypes = [ctypes.c_int, vp, vp, ctypes.c_int, vp, vp,
                                            ctypes.c_int, ctypes.c_int, ctypes.c_int64, vp, vp,
                                            vp, vp, vp, vp, ctypes.c_int]
        L.pmp_oracle_is_collision.restype = ctypes.c_int
        L.pmp_oracle_is_collision.argtypes = [ctypes.c_int, vp, vp, vp, vp]
        L.pmp_oracle_max_threads.restype = ctypes.c_int
        L.pmp_oracle_line_of_sight.restype = ctypes.c_int
        L.pmp_oracle_line_of_sight.argtypes = [ctypes.c_int, vp, vp, ctypes.c_int64, vp, vp, vp]
        _lib = L
    return _lib


def _p(a):
    return None if a is None else a.ctypes.data_as(ctypes.c_void_p)


def _prep(occ, motions, motion_g):
    occ = np.asarray(occ)
    ndim = occ.ndim
    assert ndim in (2, 3)
    # occ is indexed [x, y(, z)] like the reference's coordinate tuples; C wants z-major
    occ_c = np.ascontiguousarray(occ.T.astype(np.uint8))
    dims = np.array(list(occ.shape) + [1] * (3 - ndim), dtype=np.int32)
    if motions is None:
        motions = MOTIONS_2D if ndim == 2 else MOTIONS_3D
    motions = np.ascontiguousarray(np.asarray(motions, dtype=np.int32).reshape(-1, 3))
    if motion_g is None:
        motion_g = motion_costs(motions)
    motion_g = np.ascontiguousarray(motion_g, dtype=np.float64)
    return ndim, dims, occ_c, motions, motion_g


def _c3(p):
    a = np.zeros(3, dtype=np.int32)
    a[:len(p)] = p
    return a


@dataclass
class OracleResult:
    cost: float
    path: list            # coordinate tuples in the reference family's order
    reached: bool
    escaped: bool
    expand_cell: np.ndarray   # flat index (z*Y + y)*X + x, CLOSED insertion order
    expand_g: np.ndarray
    expand_h: np.ndarray
    expand_parent: np.ndarray
    g_field: np.ndarray       # indexed [x, y(, z)], +inf where not closed (None unless asked)

    @property
    def n_expand(self):
        return len(self.expand_cell)


def plan(occ, start, goal, algo=ASTAR, heuristic=EUCLIDEAN, motions=None, motion_g=None,
         want_field=False, want_expand=True):
    """One reference plan() on a dense occupancy array indexed [x, y(, z)] (True = obstacle)."""
    ndim, dims, occ_c, motions, motion_g = _prep(occ, motions, motion_g)
    cells = int(occ_c.size)
    cost = ctypes.c_double(0.0)
    plen = ctypes.c_int64(0)
    nexp = ctypes.c_int64(0)
    flags = ctypes.c_int32(0)
    path = np.zeros((cells, 3), dtype=np.int32)
    if want_expand:
        e_cell = np.zeros(cells, dtype=np.int64)
        e_g = np.zeros(cells, dtype=np.float64)
        e_h = np.zeros(cells, dtype=np.float64)
        e_p = np.zeros(cells, dtype=np.int64)
    else:
        e_cell = e_g = e_h = e_p = None
    field = np.zeros(cells, dtype=np.float64) if want_field else None
    s = _c3(start)
    g = _c3(goal) if goal is not None else None
    rc = lib().pmp_oracle_plan(ndim, _p(dims), _p(occ_c), len(motions), _p(motions), _p(motion_g),
                               algo, heuristic, _p(s), _p(g), ctypes.byref(cost),
                               _p(path), cells, ctypes.byref(plen),
                               _p(e_cell), _p(e_g), _p(e_h), _p(e_p), cells, ctypes.byref(nexp),
                               _p(field), ctypes.byref(flags))
    if rc != 0:
        raise RuntimeError(f"pmp_oracle_plan failed: {rc}")
    n = nexp.value
    pts = [tuple(int(v) for v in row[:ndim]) for row in path[:plen.value]]
    if field is not None:
        field = field.reshape(occ_c.shape).T
    z = np.zeros(0)
    return OracleResult(cost.value, pts, bool(flags.value & REACHED), bool(flags.value & ESCAPED),
                        e_cell[:n] if want_expand else z.astype(np.int64),
                        e_g[:n] if want_expand else z, e_h[:n] if want_expand else z,
                        e_p[:n] if want_expand else z.astype(np.int64), field)


def plan_batch(occ, starts, goals, algo=DIJKSTRA, heuristic=EUCLIDEAN, motions=None,
               motion_g=None, n_threads=0):
    """Independent plan() calls on one grid, one query per pthread work item."""
    ndim, dims, occ_c, motions, motion_g = _prep(occ, motions, motion_g)
    starts = np.asarray(starts, dtype=np.int32).reshape(-1, ndim)
    goals = np.asarray(goals, dtype=np.int32).reshape(-1, ndim)
    nq = len(starts)
    s3 = np.zeros((nq, 3), dtype=np.int32); s3[:, :ndim] = starts
    g3 = np.zeros((nq, 3), dtype=np.int32); g3[:, :ndim] = goals
    cost = np.zeros(nq, dtype=np.float64)
    nexp = np.zeros(nq, dtype=np.int64)
    plen = np.zeros(nq, dtype=np.int64)
    flags = np.zeros(nq, dtype=np.int32)
    rc = lib().pmp_oracle_plan_batch(ndim, _p(dims), _p(occ_c), len(motions), _p(motions),
                                     _p(motion_g), algo, heuristic, nq, _p(s3), _p(g3),
                                     _p(cost), _p(nexp), _p(plen), _p(flags), n_threads)
    if rc != 0:
        raise RuntimeError(f"pmp_oracle_plan_batch failed: {rc}")
    return cost, nexp, plen, flags


def is_collision(occ, p1, p2):
    ndim, dims, occ_c, _, _ = _prep(occ, None, None)
    return bool(lib().pmp_oracle_is_collision(ndim, _p(dims), _p(occ_c), _p(_c3(p1)), _p(_c3(p2))))


def line_of_sight(occ, a, b):
    """ThetaStar.lineOfSight / ThetaStar3D.lineOfSight for pairs a[i] -> b[i] ((n, ndim) arrays)"""
    ndim, dims, occ_c, _, _ = _prep(occ, None, None)
    a = np.asarray(a, dtype=np.int32).reshape(-1, ndim)
    b = np.asarray(b, dtype=np.int32).reshape(-1, ndim)
    a3 = np.zeros((len(a), 3), dtype=np.int32)
    b3 = np.zeros((len(a), 3), dtype=np.int32)
    a3[:, :ndim] = a
    b3[:, :ndim] = b
    out = np.zeros(len(a), dtype=np.uint8)
    rc = lib().pmp_oracle_line_of_sight(ndim, _p(dims), _p(occ_c), len(a), _p(a3), _p(b3), _p(out))
    if rc != 0:
        raise RuntimeError(f"pmp_oracle_line_of_sight failed: {rc}")
    return out.astype(bool)


def max_threads():
    return lib().pmp_oracle_max_threads()


def cell_to_coord(cell, shape):
    """flat oracle index -> coordinate tuple for an occupancy array of `shape` ([x,y(,z)])."""
    X, Y = shape[0], shape[1]
    c = (int(cell) % X, (int(cell) // X) % Y)
    return c if len(shape) == 2 else c + (int(cell) // (X * Y),)
