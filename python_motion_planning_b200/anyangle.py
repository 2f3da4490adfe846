"""Any-angle planning.

The reference's Theta* family (global_planner/graph_search/theta_star.py:14-171,
lazy_theta_star.py:12-118, theta_star3d.py:24-232, lazy_theta_star3d.py:24-252) is A* whose
updateVertex tries to connect a node to the parent of the node being expanded when `lineOfSight`
allows it (Lazy Theta*: assumes it does and repairs the parent when the node is popped).  Two modes:

* `mode="reference"` -- Theta* / Lazy Theta* proper, in the reference's own pop order
  (`pmp_bestfirst`, csrc/pmp_bestfirst.cu: one warp per query, the line-of-sight tests of a node's
  neighbours run in parallel lanes).  cost, path, len(expand) and every expand record equal the
  reference's, including the Distance / Visited Nodes columns of the theta_star and
  lazy_theta_star rows of its 3d_pathfinding_results.csv.  Default for grids up to 128 Ki padded
  cells (the reference's own sizes and well beyond).
* `mode="shortcut"` -- for large maps: the tiled wavefront A* (`pmp_sssp`) + path extraction, then
  `pmp_shortcut_paths` walks the extracted path and keeps a waypoint only where the line of sight
  from the previous kept waypoint ends -- the same lineOfSight predicate (`pmp_line_of_sight`,
  bit-exact against the reference's on committed vectors), applied after the search instead of
  inside it.  Result: a collision-free any-angle path with cost <= the grid A* cost; not the
  reference's Theta* path (Theta* may leave the grid-optimal corridor), documented divergence of
  this mode.  Default above 128 Ki cells.
"""
import numpy as np

from . import _lib, engine
from .environment import Node, Node3D
from .planner import GraphSearcher, GraphSearcher3D, _plan_bestfirst

REFERENCE_MODE_MAX_CELLS = 1 << 17


def line_of_sight(env, a, b):
    """Batched lineOfSight(a[i], b[i]) on env's occupancy: (n, ndim) integer arrays -> bool (n,)."""
    dg = env.device_grid()
    t = dg.torch
    ndim = dg.ndim
    a = np.asarray(a, dtype=np.int32).reshape(-1, ndim)
    b = np.asarray(b, dtype=np.int32).reshape(-1, ndim)
    n = len(a)
    if n == 0:
        return np.zeros(0, dtype=bool)
    ab = np.zeros((2, n, 3), dtype=np.int32)
    ab[0, :, :ndim] = a
    ab[1, :, :ndim] = b
    dev = t.from_numpy(ab).pin_memory().to(dg.device, non_blocking=True)
    out = t.empty(n, dtype=t.uint8, device=dg.device)
    _lib.check(dg.L.pmp_line_of_sight(dg.handle, n, dev[0].data_ptr(), dev[1].data_ptr(), out.data_ptr(),
                                      dg.stream()), "pmp_line_of_sight")
    return out.cpu().numpy().astype(bool)


def shortcut(ws):
    """pmp_shortcut_paths on the paths in ws.path / ws.path_len; returns (waypoints, len, cost) tensors"""
    dg = ws.dg
    t = dg.torch
    out_path = t.empty_like(ws.path)
    out_len = t.empty(ws.nq, dtype=t.int32, device=dg.device)
    out_cost = t.empty(ws.nq, dtype=t.float64, device=dg.device)
    _lib.check(dg.L.pmp_shortcut_paths(dg.handle, ws.nq, ws.path.data_ptr(), ws.path_cap, ws.path_len.data_ptr(),
                                       out_path.data_ptr(), out_len.data_ptr(), out_cost.data_ptr(), dg.stream()),
               "pmp_shortcut_paths")
    return out_path, out_len, out_cost


class _AnyAngleMixin:
    _algo = _lib.ASTAR
    _bf_algo = _lib.THETA_STAR

    def __init__(self, start, goal, env, heuristic_type: str = "euclidean", *, mode: str = None) -> None:
        super().__init__(start, goal, env, heuristic_type)
        self.mode = mode   # "reference" | "shortcut" | None: by grid size (REFERENCE_MODE_MAX_CELLS)

    def lineOfSight(self, node1, node2) -> bool:
        """theta_star.py:110-171 / theta_star3d.py:139-215 (evaluated on the device bitmask)"""
        return bool(line_of_sight(self.env, [node1.current], [node2.current])[0])

    def plan(self) -> tuple:
        """(cost, path, expand): path in the family's order (2D goal -> start, 3D start -> goal)"""
        dg = self._prepare()
        mode = self.mode or ("reference" if dg.pcells <= REFERENCE_MODE_MAX_CELLS else "shortcut")
        if mode == "reference":
            self.grid_cost = None
            return _plan_bestfirst(self, self._bf_algo)
        if mode != "shortcut":
            raise ValueError("mode must be 'reference', 'shortcut' or None")
        t = dg.torch
        ws = engine.get_workspace(dg, 1, _lib.W_F64)
        engine.upload_queries(ws, [self.start.current], [self.goal.current], self._ndim)
        engine.sssp(ws, self._algo, self._heuristic_id())
        engine.extract_paths(ws, self._path_cap(dg))
        if int(ws.path_len[0].item()) == -1:
            engine.extract_paths(ws, dg.pcells)
        wp, wl, wc = shortcut(ws)
        res = t.stack([ws.gcost[0], wc[0], wl[0].to(t.float64), ws.status[0].to(t.float64), ws.cost[0]]).cpu()
        self.stats = ws.stats.cpu().numpy().copy()
        self.last_field = ws
        gcost, cost, plen, status = float(res[0]), float(res[1]), int(res[2]), int(res[3])
        self.grid_cost = float(res[4])                       # cost of the grid path that was shortcut
        expand = self._expand_view(ws, gcost if np.isfinite(gcost) else None)
        _lib.check_status(status, str(self))
        if status == _lib.ST_BAD_QUERY or plen <= 0:
            return self._no_path(expand)
        pts = engine.cells_to_points(wp[0, :plen].cpu().numpy(), dg)
        if self._ndim == 3:
            pts.reverse()
        return cost, pts, expand


class ThetaStar(_AnyAngleMixin, GraphSearcher):
    """theta_star.py:14-171"""
    _node_cls = Node

    def __str__(self) -> str:
        return "Theta*"

    def run(self):
        cost, path, expand = self.plan()
        self.plot.animation(path, str(self), cost, expand)


class LazyThetaStar(ThetaStar):
    """lazy_theta_star.py:12-118"""
    _bf_algo = _lib.LAZY_THETA_STAR

    def __str__(self) -> str:
        return "Lazy Theta*"


class ThetaStar3D(_AnyAngleMixin, GraphSearcher3D):
    """theta_star3d.py:24-232"""
    _node_cls = Node3D

    def __str__(self) -> str:
        return "Theta* 3D"

    def run(self):
        cost, path, expand = self.plan()
        self.plot.animation(path, str(self), cost, expand)


class LazyThetaStar3D(ThetaStar3D):
    """lazy_theta_star3d.py:24-252"""
    _bf_algo = _lib.LAZY_THETA_STAR

    def __str__(self) -> str:
        return "Lazy Theta* 3D"
