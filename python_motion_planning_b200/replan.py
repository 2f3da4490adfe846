"""Incremental replanning on a resident device field.

Mirrors the reference's incremental planners for this path -- LPAStar / LPAStar3D
(global_planner/graph_search/lpa_star.py:38-230, lpa_star3d.py:46-225) and the dynamic-obstacle
entry of DStar3D (d_star3d.py:115-151): plan() once, then apply_change() / apply_changes() /
apply_dynamic_obstacles() after the map changed.  The reference keeps (g, rhs) per cell and
repairs with a priority queue in Python; here the cost field of the previous search stays in HBM
and `pmp_replan_update` (csrc/pmp_replan.cuh) flips the occupancy bits, resets the values that
lost their support, re-arms the cells around the change and resumes the ordinary search kernel.
The repaired field equals the field of a search from scratch on the new map (tests check this
bit for bit), so the returned cost is the optimum; the reference's LPA* extracts its path by
greedy descent on g alone (lpa_star.py:209-229) and returns a longer path in ~17 % of the rows of
its own 3d_pathfinding_results.csv -- documented divergence: cost <= the reference's.
"""
import ctypes
import math

import numpy as np

from . import _lib, engine
from .environment import Node, Node3D
from .planner import GraphSearcher, GraphSearcher3D


class ReplanSession:
    """One repairable single-source search: private field + workspace on the device."""

    def __init__(self, dg, algo, heuristic=_lib.H_EUCLIDEAN, weight=_lib.W_F64, full_field=False,
                 scratch_cells=1 << 20, window=0.0):
        t = dg.torch
        self.dg = dg
        self.ws = engine.Workspace(dg, 1, weight)          # private: the field must stay resident
        self.P = _lib.SsspParams(algo=algo, heuristic=heuristic, weight=weight,
                                 full_field=1 if full_field else 0, window=window, max_rounds=0, reserved=0)
        self.full_field = bool(full_field)
        self.scratch = t.zeros(16 + 2 * int(scratch_cells), dtype=t.int32, device=dg.device)
        self.started = False
        self.info = {"reset_cells": 0, "waves": 0, "fallback": False, "changes": 0}

    def _goal_ptr(self):
        return None if self.full_field else self.ws.goals.data_ptr()

    def begin(self, start, goal=None):
        ws, dg = self.ws, self.dg
        engine.upload_queries(ws, [start], None if goal is None else [goal], dg.ndim)
        self._begin()

    def _begin(self):
        ws, dg = self.ws, self.dg
        _lib.check(dg.L.pmp_replan_begin(dg.handle, ctypes.byref(self.P), ws.starts.data_ptr(), self._goal_ptr(),
                                         ws.g.data_ptr(), ws.ws.data_ptr(), ws.ws_bytes, ws.gcost.data_ptr(),
                                         ws.status.data_ptr(), ws.stats.data_ptr(), dg.stream()),
                   "pmp_replan_begin")
        self.started = True

    def update(self, cells, blocked):
        """cells: (n, ndim) box coordinates, blocked: (n,) new state (True = obstacle).  Returns the
        info dict; falls back to a search from scratch when the invalidated region overflows."""
        if not self.started:
            raise _lib.PmpError("ReplanSession.update before begin")
        ws, dg = self.ws, self.dg
        t = dg.torch
        cells = np.asarray(cells, dtype=np.int32).reshape(-1, dg.ndim)
        n = len(cells)
        c3 = np.zeros((n, 3), dtype=np.int32)
        c3[:, :dg.ndim] = cells
        b = np.ascontiguousarray(np.broadcast_to(np.asarray(blocked, dtype=np.uint8), (n,)))
        if n:
            host = t.from_numpy(np.concatenate([c3.view(np.uint8).reshape(-1), b])).pin_memory()
            dev = host.to(dg.device, non_blocking=True)
            cells_ptr, blocked_ptr = dev.data_ptr(), dev.data_ptr() + c3.nbytes
        else:
            dev, cells_ptr, blocked_ptr = None, None, None
        _lib.check(dg.L.pmp_replan_update(dg.handle, ctypes.byref(self.P), ws.starts.data_ptr(), self._goal_ptr(),
                                          ws.g.data_ptr(), ws.ws.data_ptr(), ws.ws_bytes, cells_ptr, blocked_ptr,
                                          n, self.scratch.data_ptr(), self.scratch.numel(), ws.gcost.data_ptr(),
                                          ws.status.data_ptr(), ws.stats.data_ptr(), dg.stream()),
                   "pmp_replan_update")
        ctl = self.scratch[:8].cpu().numpy()               # synchronises: `dev` may be released now
        self.info = {"reset_cells": int(ctl[4]), "waves": int(ctl[5]), "fallback": bool(ctl[3]), "changes": n}
        if ctl[3]:
            self._begin()
        return self.info


class _ReplanMixin:
    """plan() / apply_change() on top of a ReplanSession; shared by the 2D and 3D classes."""
    _algo = _lib.ASTAR
    _scratch_cells = 1 << 20

    def _init_replan(self):
        self._session = None
        self._result = None
        self.EXPAND = []
        self.info = None

    # ---- reference API ----
    def plan(self) -> tuple:
        """lpa_star.py:78-89 / lpa_star3d.py:88-92: (cost, path start -> goal, expand)."""
        if self._session is None:
            dg = self._prepare()
            self._session = ReplanSession(dg, self._algo, self._heuristic_id(), scratch_cells=self._scratch_cells)
            self._session.begin(self.start.current, self.goal.current)
            self._result = None
        if self._result is None:
            self._result = self._collect()
        return self._result

    def apply_change(self, coord: tuple, blocked=None) -> tuple:
        """lpa_star3d.py:98-133 (2D: the body of OnPress, lpa_star.py:112-127).  blocked = None
        toggles, True adds an obstacle, False removes one; returns plan()."""
        coord = tuple(int(v) for v in coord)
        if blocked is None:
            blocked = coord not in self.obstacles
        return self.apply_changes([coord], [bool(blocked)])

    def apply_changes(self, coords, blocked) -> tuple:
        """several cells at once: one upload, one invalidation, one resumed search"""
        coords = [tuple(int(v) for v in c) for c in coords]
        flags = [bool(b) for b in (blocked if hasattr(blocked, "__len__") else [blocked] * len(coords))]
        if self._session is None:
            self.plan()
        self._edit_host_map(coords, flags)
        self.info = self._session.update(np.array(coords, dtype=np.int32).reshape(-1, self._ndim), flags)
        self._result = None
        return self.plan()

    def apply_dynamic_obstacles(self, newly_blocked) -> tuple:
        """d_star3d.py:115-151: add obstacles, repair, return (cost, path)."""
        cells = [tuple(c) for c in newly_blocked]
        cost, path, _ = self.apply_changes(cells, [True] * len(cells)) if cells else self.plan()
        return cost, path

    def OnPress(self, event) -> None:
        """lpa_star.py:102-133 without the plot: toggle the clicked cell and replan"""
        xy = (int(event.xdata), int(event.ydata)) + ((int(event.zdata),) if self._ndim == 3 else ())
        if any(v < 0 or v > r - 1 for v, r in zip(xy, self.env._dims())):
            print("Please choose right area!")
            return
        print("Change position: " + ", ".join(f"{n} = {v}" for n, v in zip("xyz", xy)))
        self.apply_change(xy)

    # ---- internals ----
    def _edit_host_map(self, coords, flags):
        """keep env.obstacles (the set the reference planners alias) in step with the device"""
        env = self.env
        if env._dense is not None:
            d = env._dense
            for c, b in zip(coords, flags):
                d[c] = 1 if b else 0
            env._obstacles = None
            env._tree = None
        else:
            obs = env._obstacles
            for c, b in zip(coords, flags):
                if b:
                    obs.add(c)
                else:
                    obs.discard(c)
            env.update(obs)                                # env.py:78-80 (invalidates the k-d tree)
        env._device_version = env._version                 # the bitmask was patched on the device

    def _collect(self):
        ws = self._session.ws
        dg = ws.dg
        t = dg.torch
        engine.extract_paths(ws, self._path_cap(dg))
        res = t.stack([ws.gcost[0], ws.cost[0], ws.path_len[0].to(t.float64),
                       ws.status[0].to(t.float64)]).cpu()
        self.stats = ws.stats.cpu().numpy().copy()
        self.last_field = ws
        gcost, cost, plen, status = float(res[0]), float(res[1]), int(res[2]), int(res[3])
        if plen == -1:
            engine.extract_paths(ws, dg.pcells)
            plen = int(ws.path_len[0].item())
            cost = float(ws.cost[0].item())
        expand = None
        if self._ndim == 3:                                # the 2D planner returns None (lpa_star.py:89)
            expand = self._expand_view(ws, gcost if math.isfinite(gcost) else None)
            self.EXPAND = expand
        if status == _lib.ST_BAD_QUERY or plen <= 0:
            return float("inf"), [], expand
        cells = ws.path[0, :plen].cpu().numpy()
        pts = engine.cells_to_points(cells, dg)
        pts.reverse()                                      # start -> goal (lpa_star.py:229, lpa_star3d.py:224)
        return cost, pts, expand


class LPAStar(_ReplanMixin, GraphSearcher):
    """lpa_star.py:38-230"""
    _node_cls = Node

    def __init__(self, start: tuple, goal: tuple, env, heuristic_type: str = "euclidean") -> None:
        GraphSearcher.__init__(self, start, goal, env, heuristic_type)
        self._init_replan()

    def __str__(self) -> str:
        return "Lifelong Planning A*"

    def run(self) -> None:
        cost, path, _ = self.plan()
        self.plot.animation(path, str(self), cost=cost)


class LPAStar3D(_ReplanMixin, GraphSearcher3D):
    """lpa_star3d.py:46-225"""
    _node_cls = Node3D

    def __init__(self, start: tuple, goal: tuple, env, heuristic_type: str = "euclidean") -> None:
        GraphSearcher3D.__init__(self, start, goal, env, heuristic_type)
        self._init_replan()

    def __str__(self) -> str:
        return "Lifelong Planning A* 3D"

    def run(self) -> None:
        cost, path, _ = self.plan()
        if not path:
            return
        self.plot.animation(path, str(self), cost, self.EXPAND)


class DStarLite(LPAStar):
    """d_star_lite.py:14-187: same (g, rhs) repair as LPA*, searched from the goal; plan() and
    the toggling OnPress have the LPA* signatures, the path is start -> goal"""

    def __str__(self) -> str:
        return "D* Lite"


class DStar(_ReplanMixin, GraphSearcher):
    """d_star.py:38-291: plan() -> (cost, path start -> goal, None); OnPress only ADDS an obstacle
    (d_star.py:102-134) and repairs"""
    _node_cls = Node

    def __init__(self, start: tuple, goal: tuple, env) -> None:
        GraphSearcher.__init__(self, start, goal, env, None)
        self._init_replan()

    def __str__(self) -> str:
        return "Dynamic A*(D*)"

    def OnPress(self, event) -> None:
        xy = (int(event.xdata), int(event.ydata))
        if any(v < 0 or v > r - 1 for v, r in zip(xy, self.env._dims())):
            print("Please choose right area!")
        elif xy not in self.obstacles:
            print("Add obstacle at: ({}, {})".format(*xy))
            self.apply_change(xy, True)

    def run(self) -> None:
        cost, path, _ = self.plan()
        self.plot.animation(path, str(self), cost=cost)


class DStar3D(_ReplanMixin, GraphSearcher3D):
    """d_star3d.py:60-280: plan() -> (cost, path start -> goal, expand);
    apply_dynamic_obstacles(newly_blocked) -> (cost, path) (d_star3d.py:115-151)"""
    _node_cls = Node3D

    def __init__(self, start: tuple, goal: tuple, env) -> None:
        GraphSearcher3D.__init__(self, start, goal, env, None)
        self._init_replan()

    def __str__(self) -> str:
        return "Dynamic A* (D*) 3D"

    def run(self) -> None:
        cost, path, expand = self.plan()
        self.plot.animation(path, str(self), cost, expand)


class DynamicField:
    """A full cost-to-go field that follows map edits: cost_field() once, then update()."""

    def __init__(self, env, source, weight="f64", scratch_cells=1 << 22):
        from .environment import motion_table
        dg = env.device_grid()
        offsets, costs = motion_table(env, env._ndim)
        dg.set_motions(offsets, costs)
        self.env, self.dg = env, dg
        self.source = tuple(int(v) for v in source)
        w = _lib.W_F64 if weight == "f64" else _lib.W_U32
        self.session = ReplanSession(dg, _lib.DIJKSTRA, weight=w, full_field=True, scratch_cells=scratch_cells)
        self.session.begin(self.source)

    def update(self, coords, blocked):
        coords = [tuple(int(v) for v in c) for c in coords]
        flags = [bool(b) for b in (blocked if hasattr(blocked, "__len__") else [blocked] * len(coords))]
        _ReplanMixin._edit_host_map(self, coords, flags)
        return self.session.update(np.array(coords, dtype=np.int32).reshape(-1, self.dg.ndim), flags)

    def numpy(self):
        return engine.field_to_numpy(self.session.ws, 0)

    def verify(self):
        """(violations, reached cells, pending flags) of the shortest-path certificate"""
        ws, dg = self.session.ws, self.dg
        t = dg.torch
        out = t.zeros(3, dtype=t.int64, device=dg.device)
        _lib.check(dg.L.pmp_verify_field(dg.handle, ws.weight, ws.g.data_ptr(), ws.starts.data_ptr(), 0,
                                         dg.padded[2], out.data_ptr(), dg.stream()), "pmp_verify_field")
        return tuple(int(v) for v in out.cpu())
