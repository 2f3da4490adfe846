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

`reference_order=True` removes that divergence for LPAStar / LPAStar3D / DStar / DStar3D: plan() and
the change entry points then run the reference's own containers and processing order on the device
(`pmp_lpa_star` / `pmp_d_star`, csrc/pmp_incr_core.h) and return the reference's cost, path and
EXPAND -- greedy-descent detours, D*'s inf costs over directed corner edges and all.  Sequential; for
the reference's own grid sizes.
"""
import ctypes
import math

import numpy as np

from . import _lib, engine
from .environment import Node, Node3D
from .planner import ExpandView, GraphSearcher, GraphSearcher3D


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


class ReferenceIncrSession:
    """Device state of pmp_lpa_star / pmp_d_star: the reference's own LPA* / D* containers (per-cell
    g, rhs, key or h, k, parent, tag; the OPEN list; EXPAND), kept between calls so that
    apply_change / apply_dynamic_obstacles continue where plan() stopped, as the reference's planner
    objects do (csrc/pmp_incr_core.h)."""

    def __init__(self, dg, kind, heuristic, start, goal, variant=0):
        t = dg.torch
        n = dg.pcells
        self.dg, self.kind, self.heuristic = dg, kind, heuristic      # kind 0: LPA* (variant 1: D* Lite), 1: D*
        self.variant = variant
        dev = dg.device
        self.a = t.empty(n, dtype=t.float64, device=dev)               # g   | h
        self.b = t.empty(n, dtype=t.float64, device=dev)               # rhs | k
        self.key = t.empty(2 * n, dtype=t.float64, device=dev) if kind == 0 else None
        self.parent = t.empty(n, dtype=t.int32, device=dev) if kind == 1 else None
        self.tag = t.empty(n, dtype=t.uint8, device=dev) if kind == 1 else None
        self.list_cap = 8 * n + 64
        self.list = t.empty(self.list_cap, dtype=t.int32, device=dev)
        self.ctl = t.zeros(4, dtype=t.int64, device=dev)
        self.expand_cap = int(min(16 * n + 1024, 1 << 24))
        self.expand = t.empty(self.expand_cap, dtype=t.int32, device=dev)
        self.path_cap = int(max(n, 4096))
        self.path = t.empty(self.path_cap, dtype=t.int32, device=dev)
        self.out_i = t.zeros(2, dtype=t.int32, device=dev)             # path_len, n_expand
        self.out_cost = t.zeros(1, dtype=t.float64, device=dev)
        sg = np.zeros((2, 3), dtype=np.int32)
        sg[0, :dg.ndim] = start
        sg[1, :dg.ndim] = goal
        self.sg = t.from_numpy(sg).to(dev)
        self.change = t.zeros(4, dtype=t.int32, device=dev)

    def step(self, op, change=None):
        """op 0: plan from scratch; op 1: continue after an occupancy edit (LPA*: change = (coord, freed)).
        Returns (cost, path cells in kernel order, EXPAND cells)."""
        dg = self.dg
        t = dg.torch
        if change is not None:
            c = np.zeros(4, dtype=np.int32)
            c[:dg.ndim] = change[0]
            c[3] = 1 if change[1] else 0
            self.change.copy_(t.from_numpy(c))
        if self.kind == 0:
            rc = dg.L.pmp_lpa_star(dg.handle, self.variant, op, self.heuristic, self.sg.data_ptr(), self.change.data_ptr(),
                                   self.a.data_ptr(), self.b.data_ptr(), self.key.data_ptr(), self.list.data_ptr(),
                                   self.list_cap, self.ctl.data_ptr(), self.expand.data_ptr(), self.expand_cap,
                                   self.path.data_ptr(), self.path_cap, self.out_i.data_ptr(),
                                   self.out_cost.data_ptr(), self.out_i.data_ptr() + 4, dg.stream())
            _lib.check(rc, "pmp_lpa_star")
        else:
            rc = dg.L.pmp_d_star(dg.handle, op, self.sg.data_ptr(), self.a.data_ptr(), self.b.data_ptr(),
                                 self.parent.data_ptr(), self.tag.data_ptr(), self.list.data_ptr(), self.list_cap,
                                 self.ctl.data_ptr(), self.expand.data_ptr(), self.expand_cap, self.path.data_ptr(),
                                 self.path_cap, self.out_i.data_ptr(), self.out_cost.data_ptr(),
                                 self.out_i.data_ptr() + 4, dg.stream())
            _lib.check(rc, "pmp_d_star")
        oi = self.out_i.cpu().numpy()
        plen, nexp = int(oi[0]), int(oi[1])
        if plen == -2:
            raise _lib.PmpError("reference-order LPA*/D*: a list capacity was exceeded; use reference_order=False")
        if plen == -1:
            raise _lib.PmpError("reference-order LPA*/D*: path longer than the path buffer")
        cost = float(self.out_cost.cpu()[0])
        cells = self.path[:plen].cpu().numpy()
        exp = self.expand[:nexp].cpu().numpy()
        return cost, cells, exp

    def edit(self, coords, value):
        """set the occupancy of the listed cells on the device bitmask (pmp_grid_edit_cells)"""
        dg = self.dg
        t = dg.torch
        c3 = np.zeros((len(coords), 3), dtype=np.int32)
        c3[:, :dg.ndim] = np.asarray(coords, dtype=np.int32).reshape(-1, dg.ndim)
        dev = t.from_numpy(c3).to(dg.device)
        _lib.check(dg.L.pmp_grid_edit_cells(dg.handle, dev.data_ptr(), len(coords), 1 if value else 0, dg.stream()),
                   "pmp_grid_edit_cells")
        t.cuda.current_stream().synchronize()


class _ReplanMixin:
    """plan() / apply_change() on top of a ReplanSession; shared by the 2D and 3D classes."""
    _algo = _lib.ASTAR
    _scratch_cells = 1 << 20

    _ref_kind = 0          # reference-order mode: 0 = the LPA* containers, 1 = D*'s
    _ref_variant = 0       # 1: D* Lite (LPA* from the goal)

    def _init_replan(self):
        self._session = None
        self._result = None
        self.EXPAND = []
        self.info = None
        self._ref = None       # ReferenceIncrSession (reference_order=True)

    # ---- reference-order mode (pmp_lpa_star / pmp_d_star): the reference's own processing order ----
    def _ref_result(self, cost, cells, exp):
        dg = self._ref.dg
        pts = engine.cells_to_points(cells, dg)
        if self._ref_kind == 0 and self._ref_variant == 0:
            pts.reverse()                                  # start -> goal (lpa_star.py:229, lpa_star3d.py:224)
        coords = engine.cells_to_coords(exp, dg)
        t = dg.torch
        vals = self._ref.a[t.from_numpy(exp.astype(np.int64) & 0xFFFFFFFF).to(dg.device)].cpu().numpy() if len(exp) else np.zeros(0)
        self.EXPAND = ExpandView(self._node_cls, coords, vals, np.zeros(len(coords)), coords)
        expand = self.EXPAND if self._ndim == 3 else None  # the 2D classes return None (lpa_star.py:89, d_star.py:88)
        return cost, pts, expand

    def _ref_plan(self):
        if self._ref is None:
            dg = self._prepare()
            self._ref = ReferenceIncrSession(dg, self._ref_kind, self._heuristic_id(), self.start.current,
                                             self.goal.current, self._ref_variant)
            self._result = self._ref_result(*self._ref.step(0))
        return self._result

    def _ref_changes(self, coords, flags):
        self._ref_plan()
        if self._ref_variant == 1:                         # D* Lite: OnPress walks and toggles the cell itself
            for c, b in zip(coords, flags):
                if b != (c not in self.obstacles):
                    raise _lib.PmpError("reference-order D* Lite toggles the cell (d_star_lite.py:82-86)")
                self._result = self._ref_result(*self._ref.step(1, (c, False)))
                self._edit_host_map([c], [b])              # the kernel has flipped the device bit already
        elif self._ref_kind == 0:                          # LPA*: one updateVertex round + plan() per cell
            for c, b in zip(coords, flags):
                was = c in self.obstacles
                self._edit_host_map([c], [b])
                self._ref.edit([c], b)
                self._result = self._ref_result(*self._ref.step(1, (c, was and not b)))
        else:                                              # D*: obstacles are only ever added
            add = [c for c, b in zip(coords, flags) if b]
            if len(add) != len(coords):
                raise _lib.PmpError("reference-order D* only adds obstacles (d_star.py:102-127, d_star3d.py:115-151)")
            self._edit_host_map(add, [True] * len(add))
            self._ref.edit(add, True)
            self._result = self._ref_result(*self._ref.step(1))
        return self._result

    # ---- reference API ----
    def plan(self) -> tuple:
        """lpa_star.py:78-89 / lpa_star3d.py:88-92: (cost, path start -> goal, expand)."""
        if self.reference_order:
            return self._ref_plan()
        if self._session is not None and self._reinstall() != self._session_key:
            self._session = None       # the motion table changed under the resident field: search again
        if self._session is None:
            dg = self._prepare()
            self._session_key = (dg._motion_key, getattr(dg, "_cut", False))
            self._session = ReplanSession(dg, self._algo, self._heuristic_id(), scratch_cells=self._scratch_cells)
            self._session.begin(self.start.current, self.goal.current)
            self._result = None
        if self._result is None:
            self._result = self._collect()
        return self._result

    def _reinstall(self):
        """The pmp_grid handle is shared by every planner of the environment, and the repair and
        the path walk read its motion table: put this planner's table and edge rule back before
        any of them runs (another planner -- JPS with corner cutting, a filtered motion set --
        may have planned on the same env in between)."""
        return self._install_motions(self._session.ws.dg)

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
        if self.reference_order:
            return self._ref_changes(coords, flags)
        if self._session is None or self._reinstall() != self._session_key:
            self._session = None
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
        self._reinstall()
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
        _lib.check_status(status, str(self))
        if status == _lib.ST_BAD_QUERY or plen <= 0:
            return float("inf"), [], expand
        cells = ws.path[0, :plen].cpu().numpy()
        pts = engine.cells_to_points(cells, dg)
        pts.reverse()                                      # start -> goal (lpa_star.py:229, lpa_star3d.py:224)
        return cost, pts, expand


class LPAStar(_ReplanMixin, GraphSearcher):
    """lpa_star.py:38-230"""
    _node_cls = Node

    def __init__(self, start: tuple, goal: tuple, env, heuristic_type: str = "euclidean", *,
                 reference_order: bool = None) -> None:
        GraphSearcher.__init__(self, start, goal, env, heuristic_type, reference_order=reference_order)
        self._init_replan()

    def __str__(self) -> str:
        return "Lifelong Planning A*"

    def run(self) -> None:
        cost, path, _ = self.plan()
        self.plot.animation(path, str(self), cost=cost)


class LPAStar3D(_ReplanMixin, GraphSearcher3D):
    """lpa_star3d.py:46-225"""
    _node_cls = Node3D

    def __init__(self, start: tuple, goal: tuple, env, heuristic_type: str = "euclidean", *,
                 reference_order: bool = None) -> None:
        GraphSearcher3D.__init__(self, start, goal, env, heuristic_type, reference_order=reference_order)
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
    the toggling OnPress have the LPA* signatures, the path is start -> goal.  With
    reference_order=True apply_change(coord) is the walk of the reference's OnPress: it returns the
    cells walked from the start (one step on the old field, then on the repaired one) and their cost."""

    _ref_variant = 1

    def __str__(self) -> str:
        return "D* Lite"


class DStar(_ReplanMixin, GraphSearcher):
    """d_star.py:38-291: plan() -> (cost, path start -> goal, None); OnPress only ADDS an obstacle
    (d_star.py:102-134) and repairs"""
    _node_cls = Node

    _ref_kind = 1

    def __init__(self, start: tuple, goal: tuple, env, *, reference_order: bool = None) -> None:
        GraphSearcher.__init__(self, start, goal, env, None, reference_order=reference_order)
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

    _ref_kind = 1

    def __init__(self, start: tuple, goal: tuple, env, *, reference_order: bool = None) -> None:
        GraphSearcher3D.__init__(self, start, goal, env, None, reference_order=reference_order)
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
