"""Planner classes with the reference's constructor / plan() / run() contract, running on the GPU.

Mirrors /root/reference/src/python_motion_planning:
  utils/planner/planner.py:12-40, planner3d.py:12-50        (Planner, Planner3D)
  utils/planner/search_factory.py:9-50                      (SearchFactory)
  global_planner/graph_search/graph_search.py:11-87         (GraphSearcher: h, cost, isCollision)
  global_planner/graph_search/graph_search_3d.py:11-107     (GraphSearcher3D)
  global_planner/graph_search/{a_star,dijkstra,gbfs}.py and {a_star3d,dijkstra3d,gbfs3d}.py

Return conventions are kept: 2D path goal->start and ([], [], []) when
no path exists; 3D path start->goal and (inf, [], expand) when none exists; `cost` is the sum of
step lengths accumulated goal->start.  Documented differences: among equal-cost paths a
different one may be returned; `expand` is the set of cells the wavefront settled (for A* the
cells with g + h <= cost), not the reference's pop order; A* with an inadmissible heuristic
(manhattan on 8/26-connected grids) returns the optimal cost, which can be lower than the
reference's.

`reference_order=True` (constructor keyword, `planner.reference_order`, the module default
`set_reference_order()`, or PMP_B200_REFERENCE_ORDER=1) removes those differences: plan() then runs
the reference's own heap discipline on one warp (`pmp_bestfirst`, csrc/pmp_bestfirst.cu) and cost,
path, len(expand) and every expand record (current, parent, g, h, in CLOSED order) equal the
reference's.  It is the mode for the reference's own grid sizes; the wavefront is the mode for
large grids and batches.
"""
import math
import os
from collections.abc import Sequence

import numpy as np

from . import _lib, engine
from .environment import Env, Env3D, Grid, Grid3D, Node, Node3D, motion_table


_REFERENCE_ORDER = os.environ.get("PMP_B200_REFERENCE_ORDER", "0") not in ("", "0")


def set_reference_order(on: bool) -> None:
    """module default of `reference_order` for planners constructed afterwards"""
    global _REFERENCE_ORDER
    _REFERENCE_ORDER = bool(on)


class HeadlessPlot:
    """Stand-in for utils/plot/plot.py / plot3d.py (matplotlib / pyvista windows): there is no
    graphics stack on the target machines, so animation() records its arguments instead."""

    def __init__(self, start, goal, env):
        self.start, self.goal, self.env = start, goal, env
        self.last = None

    def animation(self, path, name, cost=None, expand=None, *args, **kwargs):
        self.last = dict(path=path, name=name, cost=cost, expand=expand, args=args, kwargs=kwargs)

    def __getattr__(self, item):            # plotPath, plotExpand, ... : accepted and ignored
        if item.startswith("_"):
            raise AttributeError(item)
        return lambda *a, **k: None


class ExpandView(Sequence):
    """`expand` (third value of plan()): the settled cells as Node / Node3D records, built on
    demand from arrays so that 10^8-cell searches do not allocate 10^8 Python objects up front.
    Supports len(), indexing, slicing, iteration and list(expand)."""

    def __init__(self, node_cls, coords, g, h, parents):
        self._cls = node_cls
        self._coords, self._g, self._h, self._parents = coords, g, h, parents

    def __len__(self):
        return len(self._coords)

    def __getitem__(self, i):
        if isinstance(i, slice):
            return ExpandView(self._cls, self._coords[i], self._g[i], self._h[i], self._parents[i])
        c = tuple(int(v) for v in self._coords[i])
        p = tuple(int(v) for v in self._parents[i])
        return self._cls(c, p, float(self._g[i]), float(self._h[i]))

    def __eq__(self, other):
        if isinstance(other, (list, tuple)) and len(other) == 0:
            return len(self) == 0
        return NotImplemented if not isinstance(other, ExpandView) else self is other


class Planner:
    """planner.py:12-40"""
    _node_cls = Node
    _ndim = 2

    def __init__(self, start: tuple, goal: tuple, env) -> None:
        self.start = self._node_cls(start, start, 0, 0)
        self.goal = self._node_cls(goal, goal, 0, 0)
        self.env = env
        self.plot = HeadlessPlot(start, goal, env)

    def dist(self, node1, node2) -> float:
        return math.hypot(node2.x - node1.x, node2.y - node1.y)

    def angle(self, node1, node2) -> float:
        return math.atan2(node2.y - node1.y, node2.x - node1.x)

    def plan(self):
        raise NotImplementedError

    def run(self):
        raise NotImplementedError


class Planner3D(Planner):
    """planner3d.py:12-50"""
    _node_cls = Node3D
    _ndim = 3

    def dist(self, node1, node2) -> float:
        return math.sqrt((node2.x - node1.x) ** 2 + (node2.y - node1.y) ** 2 + (node2.z - node1.z) ** 2)

    def angle(self, node1, node2):
        dx, dy, dz = node2.x - node1.x, node2.y - node1.y, node2.z - node1.z
        return math.atan2(dy, dx), math.atan2(dz, math.hypot(dx, dy))


class _SearchMixin:
    """What GraphSearcher / GraphSearcher3D add: heuristic, cost, isCollision, and the GPU plan."""
    _algo = _lib.ASTAR

    def _init_search(self, heuristic_type, reference_order=None):
        self.heuristic_type = heuristic_type
        self.reference_order = _REFERENCE_ORDER if reference_order is None else bool(reference_order)
        self.motions = self.env.motions        # alias, as graph_search.py:26
        self._stats = None                     # device counters of the last plan(): rounds, tiles, cells, edges
        self._stats_dev = None
        self.last_field = None                 # (Workspace) device cost field of the last plan()

    @property
    def stats(self):
        """counters of the last plan() (rounds, tile activations, cells, edges, ...), fetched on demand"""
        if self._stats is None and self._stats_dev is not None:
            self._stats = self._stats_dev.cpu().numpy().copy()
        return self._stats

    @stats.setter
    def stats(self, value):
        self._stats = value
        self._stats_dev = None

    @property
    def obstacles(self):
        """graph_search.py:28 aliases env.obstacles at construction; here it is read through
        the environment so that dense (array-backed) grids never build a tuple set unasked."""
        return self.env.obstacles

    # ---- reference helpers, unchanged semantics ----
    def h(self, node, goal) -> float:
        d = [abs(a - b) for a, b in zip(goal.current, node.current)]
        if self.heuristic_type == "manhattan":
            return sum(d)
        elif self.heuristic_type == "euclidean":
            return math.hypot(*d) if self._ndim == 2 else math.sqrt(sum(v ** 2 for v in d))

    def cost(self, node1, node2) -> float:
        if self.isCollision(node1, node2):
            return float("inf")
        return self.dist(node1, node2)

    def isCollision(self, node1, node2) -> bool:
        """graph_search.py:61-87 / graph_search_3d.py:66-107 on the host set (for user code;
        the kernels evaluate the same rule on the bitmask)."""
        obs = self.obstacles
        p1, p2 = tuple(node1.current), tuple(node2.current)
        if p1 in obs or p2 in obs:
            return True
        d = [b - a for a, b in zip(p1, p2)]
        if self._ndim == 3 and max(abs(v) for v in d) > 1:
            return False
        moving = [i for i, v in enumerate(d) if v != 0]
        if len(moving) <= 1:
            return False
        if self._ndim == 2:
            x1, y1 = p1
            x2, y2 = p2
            if x2 - x1 == y1 - y2:
                s1, s2 = (min(x1, x2), min(y1, y2)), (max(x1, x2), max(y1, y2))
            else:
                s1, s2 = (min(x1, x2), max(y1, y2)), (max(x1, x2), min(y1, y2))
            return s1 in obs or s2 in obs
        for i in moving:   # face neighbours of the source along each moving axis
            c = list(p1)
            c[i] += d[i]
            if tuple(c) in obs:
                return True
        return False

    def getNeighbor(self, node):
        return [node + m for m in self.motions if not self.isCollision(node, node + m)]

    # ---- GPU plan ----
    def _heuristic_id(self):
        return _lib.H_MANHATTAN if self.heuristic_type == "manhattan" else _lib.H_EUCLIDEAN

    def _install_motions(self, dg):
        """this planner's motion table and edge rule on the (shared) device grid; returns the
        installed (motion key, corner-cutting flag)"""
        offsets, costs = motion_table(self.env if self.motions is self.env.motions else self, self._ndim)
        dg.set_motions(offsets, costs, cut=getattr(self, "_cut", False))
        return dg._motion_key, getattr(dg, "_cut", False)

    def _prepare(self):
        dg = self.env.device_grid()
        self._install_motions(dg)
        return dg

    def _no_path(self, expand):
        if self._ndim == 2:
            return [], [], []                  # a_star.py:83
        return float("inf"), [], expand        # a_star3d.py:77-78

    def _plan_sssp(self):
        dg = self._prepare()
        t = dg.torch
        ws = engine.get_workspace(dg, 1, _lib.W_F64)
        engine.upload_queries(ws, [self.start.current], [self.goal.current], self._ndim)
        engine.sssp(ws, self._algo, self._heuristic_id())
        engine.extract_paths(ws, self._path_cap(dg))
        # one synchronisation for everything plan() returns: the scalars and the head of the path
        # land in pinned host buffers, the counters are cloned on the device and read on demand
        ws.host_res.copy_(t.stack([ws.gcost[0], ws.cost[0], ws.path_len[0].to(t.float64),
                                   ws.status[0].to(t.float64)]), non_blocking=True)
        nhead = min(ws.path_cap, ws.host_path.numel())
        ws.host_path[:nhead].copy_(ws.path[0, :nhead], non_blocking=True)
        self._stats_dev = ws.stats.clone()
        self._stats = None
        t.cuda.current_stream().synchronize()
        self.last_field = ws
        res = ws.host_res
        gcost, cost, plen, status = float(res[0]), float(res[1]), int(res[2]), int(res[3])
        _lib.check_status(status, str(self))
        if status == _lib.ST_BAD_QUERY:
            return self._no_path(self._expand_view(ws, None))
        if plen == -1:                         # path longer than the first guess: retry full size
            engine.extract_paths(ws, dg.pcells)
            plen = int(ws.path_len[0].item())
            cost = float(ws.cost[0].item())
            nhead = 0
        expand = self._expand_view(ws, gcost if math.isfinite(gcost) else None)
        if plen <= 0:
            return self._no_path(expand)
        cells = ws.host_path[:plen].numpy().copy() if plen <= nhead else ws.path[0, :plen].cpu().numpy()
        pts = engine.cells_to_points(cells, dg)
        if self._ndim == 3:
            pts.reverse()                      # a_star3d.py:105: start -> goal
        return cost, pts, expand

    def _path_cap(self, dg):
        return int(min(dg.pcells, 4 * sum(dg.dims) + 4096))

    def _expand_view(self, ws, bound):
        """settled cells (finite g; with a bound: g + h <= bound) as an ExpandView"""
        planner = self

        class _Lazy(ExpandView):
            def __init__(self):
                self._built = False
                self._cls = planner._node_cls

            def _build(self):
                if not self._built:
                    c, g, h, p = planner._expand_arrays(ws, bound, getattr(self, "_snap", None))
                    self._snap = None
                    self._coords, self._g, self._h, self._parents = c, g, h, p
                    self._built = True

            def __len__(self):
                self._build()
                return len(self._coords)

            def __getitem__(self, i):
                self._build()
                return ExpandView.__getitem__(self, i)

        view = _Lazy()
        if ws.dg.pcells <= (1 << 22):          # small grids: keep a device copy of the field (the
            view._snap = ws.g[0].clone()       # workspace is reused by the next plan()); records
        return view                            # are still built on first use only

    def _expand_arrays(self, ws, bound, snap=None):
        dg = ws.dg
        f = engine.field_to_numpy(ws, 0, snap)
        finite = np.isfinite(f)
        coords = np.argwhere(finite)
        g = f[finite]
        goal = np.array(self.goal.current)
        d = np.abs(coords - goal)
        if self._algo == _lib.DIJKSTRA:
            h = np.zeros(len(coords))
        elif self.heuristic_type == "manhattan":
            h = d.sum(axis=1).astype(np.float64)
        else:
            h = np.sqrt((d.astype(np.float64) ** 2).sum(axis=1))
        if bound is not None:
            # the region the device closed: g + h_adm <= cost, h_adm the admissible heuristic
            # the kernels order by (max of scaled euclid / manhattan; csrc/pmp_grid.cu)
            h_adm = 0.0
            if self._algo == _lib.ASTAR:
                offs, costs = motion_table(self.env if self.motions is self.env.motions else self, self._ndim)
                he = min(c / math.sqrt(sum(v * v for v in o)) for o, c in zip(offs, costs))
                hm = min(c / sum(1 for v in o if v) for o, c in zip(offs, costs))
                h_adm = np.maximum(he * np.sqrt((d.astype(np.float64) ** 2).sum(axis=1)), hm * d.sum(axis=1))
            keep = g + h_adm <= bound * (1 + 1e-9)
            coords, g, h = coords[keep], g[keep], h[keep]
        order = np.lexsort((g, g + h))         # by f, then g; the start cell always first
        is_start = np.all(coords[order] == np.array(self.start.current), axis=1)
        order = np.concatenate([order[is_start], order[~is_start]])
        coords, g, h = coords[order], g[order], h[order]
        parents = _parents_from_field(f, self.env.occupancy(), coords, g, motion_table(
            self.env if self.motions is self.env.motions else self, self._ndim))
        return coords, g, h, parents


def _parents_from_field(f, occ, coords, g, table):
    """parent of every listed cell: the legal in-neighbour u minimising g(u) + w (lowest motion
    index on ties) -- the same rule the path-extraction kernel applies; vectorised numpy,
    used only to fill `expand` records."""
    offsets, costs = table
    ndim = f.ndim
    n = len(coords)
    parents = coords.copy()
    if n == 0:
        return parents
    pad = 2
    fp = np.pad(f, pad, constant_values=np.inf)
    op = np.pad(occ, pad, constant_values=True)
    best = np.full(n, np.inf)
    c = coords + pad
    for off, w in zip(offsets, costs):
        off = np.array(off[:ndim])
        u = c - off
        ui = tuple(u.T)
        legal = ~op[ui]
        moving = [i for i in range(ndim) if off[i] != 0]
        if len(moving) >= 2:                   # face neighbours of the source u
            for i in moving:
                fc = u.copy()
                fc[:, i] += off[i]
                legal &= ~op[tuple(fc.T)]
        cand = np.where(legal, fp[ui] + w, np.inf)
        better = cand < best
        best = np.where(better, cand, best)
        parents[better] = (u - pad)[better]
    return parents


class GraphSearcher(_SearchMixin, Planner):
    """graph_search.py:11-87"""

    def __init__(self, start: tuple, goal: tuple, env: Grid, heuristic_type: str = "euclidean", *,
                 reference_order: bool = None) -> None:
        Planner.__init__(self, start, goal, env)
        self._init_search(heuristic_type, reference_order)


class GraphSearcher3D(_SearchMixin, Planner3D):
    """graph_search_3d.py:11-107"""

    def __init__(self, start: tuple, goal: tuple, env: Grid3D, heuristic_type: str = "euclidean", *,
                 reference_order: bool = None) -> None:
        Planner3D.__init__(self, start, goal, env)
        self._init_search(heuristic_type, reference_order)


class AStar(GraphSearcher):
    """a_star.py:10-124.  plan() -> (cost, path goal->start, expand)."""
    _algo = _lib.ASTAR

    def __str__(self) -> str:
        return "A*"

    def plan(self) -> tuple:
        if self.reference_order:
            return _plan_bestfirst(self, self._algo)
        return self._plan_sssp()

    def extractPath(self, closed_list: dict) -> tuple:
        """a_star.py:98-117 on a {coord: Node} dict (kept for user code)."""
        cost = 0
        node = closed_list[self.goal.current]
        path = [node.current]
        while node != self.start:
            node_parent = closed_list[node.parent]
            cost += self.dist(node, node_parent)
            node = node_parent
            path.append(node.current)
        return cost, path

    def run(self):
        cost, path, expand = self.plan()
        self.plot.animation(path, str(self), cost, expand)


class Dijkstra(AStar):
    """dijkstra.py:10-85"""
    _algo = _lib.DIJKSTRA

    def __str__(self) -> str:
        return "Dijkstra"


class GBFS(AStar):
    """gbfs.py:10-86: greedy best-first, run as the reference's sequential loop on one warp."""
    _algo = _lib.GBFS

    def __str__(self) -> str:
        return "Greedy Best First Search(GBFS)"

    def plan(self) -> tuple:
        if self.reference_order:               # also CLOSED order and the g / h of every record
            return _plan_bestfirst(self, _lib.GBFS)
        return _plan_gbfs(self)


class AStar3D(GraphSearcher3D):
    """a_star3d.py:18-111.  plan() -> (cost, path start->goal, expand)."""
    _algo = _lib.ASTAR

    def __init__(self, start: tuple, goal: tuple, env: Grid3D, heuristic_type: str = "euclidean", *,
                 reference_order: bool = None) -> None:
        super().__init__(start, goal, env, heuristic_type, reference_order=reference_order)
        self.start.g = 0.0
        self.start.h = self.h(self.start, self.goal)

    def __str__(self) -> str:
        return "A*"

    def plan(self) -> tuple:
        if self.reference_order:
            return _plan_bestfirst(self, self._algo)
        return self._plan_sssp()

    def extractPath(self, closed_list: dict) -> tuple:
        cost = 0.0
        node = closed_list[self.goal.current]
        path = [node.current]
        while node != self.start:
            parent = closed_list[node.parent]
            cost += self.dist(node, parent)
            node = parent
            path.append(node.current)
        path.reverse()
        return cost, path

    def run(self):
        cost, path, expand = self.plan()
        self.plot.animation(path, str(self), cost, expand)


class Dijkstra3D(AStar3D):
    """dijkstra3d.py:19-147: the constructor takes no heuristic argument (dijkstra3d.py:29-31)."""
    _algo = _lib.DIJKSTRA

    def __init__(self, start: tuple, goal: tuple, env: Grid3D, *, reference_order: bool = None) -> None:
        GraphSearcher3D.__init__(self, start, goal, env, None, reference_order=reference_order)
        self.start.g = 0.0
        self.start.h = 0.0

    def __str__(self) -> str:
        return "Dijkstra (3D)"


class GBFS3D(AStar3D):
    """gbfs3d.py:20-114"""
    _algo = _lib.GBFS

    def __str__(self) -> str:
        return "Greedy Best First Search (GBFS) 3D"

    def plan(self) -> tuple:
        if self.reference_order:               # also CLOSED order and the g / h of every record
            return _plan_bestfirst(self, _lib.GBFS)
        return _plan_gbfs(self)


def _plan_gbfs(planner):
    dg = planner._prepare()
    t = dg.torch
    ndim = planner._ndim
    ws = engine.get_workspace(dg, 1, _lib.W_F64)
    engine.upload_queries(ws, [planner.start.current], [planner.goal.current], ndim)
    heap_cap = int(min(len(planner.motions) * dg.pcells + 1, 1 << 26))
    cache = dg.__dict__.setdefault("_gbfs_cache", {})
    if cache.get("cap", 0) < heap_cap:
        cache["parent"] = t.empty(dg.pcells, dtype=t.int32, device=dg.device)
        cache["heap"] = t.empty(heap_cap * 16, dtype=t.uint8, device=dg.device)
        cache["nexp"] = t.empty(1, dtype=t.int32, device=dg.device)
        cache["cap"] = heap_cap
    ws.ensure_path(dg.pcells if dg.pcells <= (1 << 22) else planner._path_cap(dg))
    _lib.check(dg.L.pmp_gbfs(dg.handle, planner._heuristic_id(), 1, ws.starts.data_ptr(),
                             ws.goals.data_ptr(), cache["parent"].data_ptr(),
                             cache["heap"].data_ptr(), heap_cap, ws.path.data_ptr(), ws.path_cap,
                             ws.path_len.data_ptr(), ws.cost.data_ptr(), cache["nexp"].data_ptr(),
                             dg.stream()), "pmp_gbfs")
    plen = int(ws.path_len[0].item())
    if plen < 0:
        raise _lib.PmpError("pmp_gbfs: path or heap capacity exceeded (%d)" % plen)
    cost = float(ws.cost[0].item())
    # expand: the CLOSED set with parents, from the parent array
    par = cache["parent"].cpu().numpy().astype(np.int64) & 0xFFFFFFFF
    closed = np.nonzero(par != 0xFFFFFFFF)[0]
    coords = engine.cells_to_coords(closed, dg)
    parents = engine.cells_to_coords(par[closed], dg)
    goal = np.array(planner.goal.current)
    d = np.abs(coords - goal)
    if planner.heuristic_type == "manhattan":
        h = d.sum(axis=1).astype(np.float64)
    else:
        h = np.sqrt((d.astype(np.float64) ** 2).sum(axis=1))
    expand = ExpandView(planner._node_cls, coords, np.zeros(len(coords)), h, parents)
    if plen == 0:
        return planner._no_path(expand)
    pts = engine.cells_to_points(ws.path[0, :plen].cpu().numpy(), dg)
    if ndim == 3:
        pts.reverse()
    return cost, pts, expand


def _plan_bestfirst(planner, algo):
    """plan() in the reference's pop order (pmp_bestfirst): (cost, path, expand) equal to the
    reference's, expand records included.  algo: _lib.ASTAR / DIJKSTRA / GBFS / THETA_STAR /
    LAZY_THETA_STAR."""
    dg = planner._prepare()
    t = dg.torch
    ndim = planner._ndim
    ws = engine.get_workspace(dg, 1, _lib.W_F64)
    engine.upload_queries(ws, [planner.start.current], [planner.goal.current], ndim)
    heap_cap = int(min(len(planner.motions) * dg.pcells + 1, 1 << 23))
    cache = dg.__dict__.setdefault("_bestfirst_cache", {})
    if cache.get("cap", 0) < heap_cap:
        cache["heap"] = t.empty(heap_cap * _lib.BESTFIRST_HEAP_REC_BYTES, dtype=t.uint8, device=dg.device)
        cache["g"] = t.empty(dg.pcells, dtype=t.float64, device=dg.device)
        cache["h"] = t.empty(dg.pcells, dtype=t.float64, device=dg.device)
        cache["parent"] = t.empty(dg.pcells, dtype=t.int32, device=dg.device)
        cache["order"] = t.empty(dg.pcells, dtype=t.int32, device=dg.device)
        cache["nexp"] = t.empty(1, dtype=t.int32, device=dg.device)
        cache["cap"] = heap_cap
    ws.ensure_path(dg.pcells if dg.pcells <= (1 << 22) else planner._path_cap(dg))
    heur = planner._heuristic_id() if planner.heuristic_type is not None else _lib.H_EUCLIDEAN
    _lib.check(dg.L.pmp_bestfirst(dg.handle, algo, heur, 1, ws.starts.data_ptr(), ws.goals.data_ptr(),
                                  cache["heap"].data_ptr(), heap_cap, cache["g"].data_ptr(),
                                  cache["h"].data_ptr(), cache["parent"].data_ptr(),
                                  cache["order"].data_ptr(), dg.pcells, ws.path.data_ptr(), ws.path_cap,
                                  ws.path_len.data_ptr(), ws.cost.data_ptr(), cache["nexp"].data_ptr(),
                                  dg.stream()), "pmp_bestfirst")
    res = t.stack([ws.cost[0], ws.path_len[0].to(t.float64), cache["nexp"][0].to(t.float64)]).cpu()
    cost, plen, n = float(res[0]), int(res[1]), int(res[2])
    if plen == -2:
        raise _lib.PmpError("pmp_bestfirst: OPEN / CLOSED capacity exceeded; use the wavefront mode "
                            "(reference_order=False) for grids of this size")
    if plen == -1:
        raise _lib.PmpError("pmp_bestfirst: path longer than the path buffer")
    # expand = list(CLOSED.values()): insertion order, latest values
    order = cache["order"][:n].to(t.int64)
    e_g = cache["g"][order].cpu().numpy()
    e_h = cache["h"][order].cpu().numpy()
    e_par = cache["parent"][order].cpu().numpy()
    coords = engine.cells_to_coords(order.cpu().numpy(), dg)
    parents = engine.cells_to_coords(e_par, dg)
    expand = ExpandView(planner._node_cls, coords, e_g, e_h, parents)
    if plen == 0:
        return planner._no_path(expand)
    pts = engine.cells_to_points(ws.path[0, :plen].cpu().numpy(), dg)
    if ndim == 3:
        pts.reverse()                          # a_star3d.py:105: start -> goal
    return cost, pts, expand


class SearchFactory(object):
    """search_factory.py:9-50 -- the graph-search entries this package implements."""

    def __call__(self, planner_name, **config):
        if planner_name == "a_star":
            return AStar(**config)
        elif planner_name == "dijkstra":
            return Dijkstra(**config)
        elif planner_name == "gbfs":
            return GBFS(**config)
        elif planner_name == "lpa_star":
            from .replan import LPAStar
            return LPAStar(**config)
        elif planner_name == "jps":
            from .jps import JPS
            return JPS(**config)
        elif planner_name == "d_star":
            from .replan import DStar
            return DStar(**config)
        elif planner_name == "d_star_lite":
            from .replan import DStarLite
            return DStarLite(**config)
        elif planner_name == "theta_star":
            from .anyangle import ThetaStar
            return ThetaStar(**config)
        elif planner_name == "lazy_theta_star":
            from .anyangle import LazyThetaStar
            return LazyThetaStar(**config)
        raise ValueError("The `planner_name` must be set correctly.")
