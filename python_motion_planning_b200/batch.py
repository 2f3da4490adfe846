"""Batched queries and full cost-to-go fields: the data-parallel entry points that have no
counterpart in the reference's one-query-at-a-time API but run the same planners' arithmetic
(BASELINE.json configs 3-5).  Independent queries share one occupancy bitmask and are
processed concurrently by one persistent kernel."""
import math

import numpy as np

from . import _lib, engine
from .environment import motion_table

_ALGO = {"a_star": _lib.ASTAR, "astar": _lib.ASTAR, "dijkstra": _lib.DIJKSTRA}


class BatchResult:
    """costs: (nq,) float64 (+inf = no path); paths: list of (n_i, ndim) int arrays in
    goal->start order (None when not requested); stats: kernel counters; timing in ms."""

    def __init__(self):
        self.costs = None
        self.gcosts = None
        self.status = None
        self.paths = None
        self.path_len = None
        self.stats = np.zeros(16, dtype=np.int64)
        self.settled = None
        self.h2d_bytes = 0
        self.d2h_bytes = 0


def _wave_size(dg, nq, weight, max_bytes):
    esz = 8 if weight == _lib.W_F64 else 4
    per_q = dg.pcells * esz + dg.ntiles * (8 + 128 + 8)
    w = max(1, int(max_bytes // per_q))
    w = min(w, nq, max(1, (2 ** 32 - 2) // max(dg.ntiles, 1) - 1))
    return w


def plan_batch(env, starts, goals, algo="dijkstra", heuristic_type="euclidean", weight="f64",
               want_paths=True, want_settled=False, window=0.0, max_concurrent=None,
               max_field_bytes=48 << 30, path_cap=None):
    """Plan nq independent (start, goal) queries on one environment.

    starts, goals: (nq, ndim) integer arrays.  Returns a BatchResult.  Costs follow the
    reference's convention (sum of step lengths accumulated goal->start along the returned
    path); paths are goal->start coordinate arrays."""
    ndim = env._ndim
    starts = np.asarray(starts, dtype=np.int32).reshape(-1, ndim)
    goals = np.asarray(goals, dtype=np.int32).reshape(-1, ndim)
    nq = len(starts)
    dg = env.device_grid()
    dg.set_motions(*motion_table(env, ndim))
    t = dg.torch
    wt = _lib.W_U32 if weight == "u32" else _lib.W_F64
    wave = _wave_size(dg, nq, wt, max_field_bytes)
    if max_concurrent:
        wave = min(wave, int(max_concurrent))
    res = BatchResult()
    res.costs = np.full(nq, np.inf)
    res.gcosts = np.full(nq, np.inf)
    res.status = np.zeros(nq, dtype=np.int32)
    res.path_len = np.zeros(nq, dtype=np.int32)
    if want_paths:
        res.paths = [None] * nq
    if want_settled:
        res.settled = np.zeros(nq, dtype=np.int64)
    hid = _lib.H_MANHATTAN if heuristic_type == "manhattan" else _lib.H_EUCLIDEAN
    cap = path_cap or int(min(dg.pcells, 6 * sum(dg.dims) + 4096))
    for q0 in range(0, nq, wave):
        n = min(wave, nq - q0)
        ws = engine.get_workspace(dg, n, wt)
        res.h2d_bytes += engine.upload_queries(ws, starts[q0:q0 + n], goals[q0:q0 + n], ndim)
        engine.sssp(ws, _ALGO[algo], hid, window=window)
        engine.extract_paths(ws, cap)
        plen = ws.path_len.cpu().numpy()
        if (plen == -1).any():                      # rare: a path longer than the first guess
            cap = int(dg.pcells)
            engine.extract_paths(ws, cap)
            plen = ws.path_len.cpu().numpy()
        res.costs[q0:q0 + n] = ws.cost.cpu().numpy()
        res.gcosts[q0:q0 + n] = ws.gcost.cpu().numpy()
        res.status[q0:q0 + n] = ws.status.cpu().numpy()
        _lib.check_status(res.status[q0:q0 + n], "plan_batch")
        res.path_len[q0:q0 + n] = plen
        res.stats += ws.stats.cpu().numpy()
        res.d2h_bytes += n * (8 + 8 + 4 + 4) + 64
        if want_settled:
            res.settled[q0:q0 + n] = engine.count_settled(ws, bounded=True, use_h=_ALGO[algo] == _lib.ASTAR).cpu().numpy()
        if want_paths:
            # one compacted D2H of all path cells of the wave
            lens = np.maximum(plen, 0)
            mask = t.arange(ws.path_cap, device=dg.device)[None, :] < t.from_numpy(lens).to(dg.device)[:, None]
            flat = ws.path[mask].cpu().numpy()
            res.d2h_bytes += flat.nbytes
            coords = engine.cells_to_coords(flat, dg)
            off = np.concatenate([[0], np.cumsum(lens)])
            for i in range(n):
                res.paths[q0 + i] = coords[off[i]:off[i + 1]]
    res.costs[res.path_len <= 0] = np.inf
    return res


def cost_field(env, source, weight="f64", window=0.0, to_host=True):
    """Full single-source cost-to-go field (Dijkstra run to exhaustion, as Dijkstra3D.plan()
    does when the goal is unreachable, dijkstra3d.py:86-87).  Returns a float64 numpy array
    indexed [x, y(, z)] with +inf for unreachable cells, or (Workspace, stats) with
    to_host=False (the field stays on the device: Workspace.g[0], padded layout)."""
    ndim = env._ndim
    dg = env.device_grid(sync=False)
    wt = _lib.W_U32 if weight == "u32" else _lib.W_F64
    ws = engine.get_workspace(dg, 1, wt)
    # the query goes up first: a streamed occupancy upload (update_packed with pinned rows) keeps the H2D copy engine
    # busy for milliseconds, and the search is meant to start behind its first chunk, not behind its last
    engine.upload_queries(ws, [tuple(source)], None, ndim)
    env.device_grid()
    dg.set_motions(*motion_table(env, ndim))
    engine.sssp(ws, _lib.DIJKSTRA, 0, full_field=True, window=window, have_goals=False)
    if not to_host:
        return ws, ws.stats        # asynchronous: the caller checks ws.status (engine.check_field_status)
    _lib.check_status(ws.status.cpu().numpy(), "cost_field")
    return engine.field_to_numpy(ws, 0)


def shard_range(nq, rank, world):
    """contiguous slice [q0, q1) of the queries planned by `rank` (sizes differ by at most one)"""
    base, rem = divmod(nq, world)
    q0 = rank * base + min(rank, rem)
    return q0, q0 + base + (1 if rank < rem else 0)


def plan_batch_sharded(env, starts, goals, dist=None, gather=True, _plan=None, **kw):
    """plan_batch over several GPUs: the occupancy bitmask is
    replicated, the queries are split 1/P per rank, no communication on the data path.  With
    gather=True the per-query scalars (costs, gcosts, status, path_len) are all-gathered so
    every rank returns the complete arrays; paths stay on the rank that planned them
    (`res.paths[q]` is None elsewhere).  `dist`: torch.distributed (initialised), or None."""
    plan = _plan or plan_batch
    ndim = env._ndim
    starts = np.asarray(starts, dtype=np.int32).reshape(-1, ndim)
    goals = np.asarray(goals, dtype=np.int32).reshape(-1, ndim)
    nq = len(starts)
    world = dist.get_world_size() if dist is not None else 1
    rank = dist.get_rank() if dist is not None else 0
    q0, q1 = shard_range(nq, rank, world)
    local = plan(env, starts[q0:q1], goals[q0:q1], **kw) if q1 > q0 else None
    res = BatchResult()
    res.costs = np.full(nq, np.inf)
    res.gcosts = np.full(nq, np.inf)
    res.status = np.zeros(nq, dtype=np.int32)
    res.path_len = np.zeros(nq, dtype=np.int32)
    res.shard = (q0, q1)
    if local is not None:
        res.costs[q0:q1], res.gcosts[q0:q1] = local.costs, local.gcosts
        res.status[q0:q1], res.path_len[q0:q1] = local.status, local.path_len
        res.stats = local.stats
        res.h2d_bytes, res.d2h_bytes = local.h2d_bytes, local.d2h_bytes
        if local.paths is not None:
            res.paths = [None] * nq
            res.paths[q0:q1] = local.paths
        if local.settled is not None:
            res.settled = np.zeros(nq, dtype=np.int64)
            res.settled[q0:q1] = local.settled
    if gather and world > 1:
        import torch
        dev = "cuda" if dist.get_backend() == "nccl" else "cpu"
        # +inf / zeros outside the own shard: min / max reductions assemble the full arrays
        c = torch.from_numpy(np.stack([res.costs, res.gcosts])).to(dev)
        dist.all_reduce(c, op=dist.ReduceOp.MIN)
        s = torch.from_numpy(np.stack([res.status, res.path_len]).astype(np.int64)).to(dev)
        dist.all_reduce(s, op=dist.ReduceOp.MAX)
        c, s = c.cpu().numpy(), s.cpu().numpy()
        res.costs, res.gcosts = c[0], c[1]
        res.status, res.path_len = s[0].astype(np.int32), s[1].astype(np.int32)
    return res


_REF_ALGO = {"a_star": _lib.ASTAR, "astar": _lib.ASTAR, "dijkstra": _lib.DIJKSTRA, "gbfs": _lib.GBFS,
             "theta_star": _lib.THETA_STAR, "lazy_theta_star": _lib.LAZY_THETA_STAR, "jps": _lib.JPS}


def plan_batch_reference(env, starts, goals, algo="a_star", heuristic_type="euclidean", want_paths=True,
                         max_bytes=8 << 30):
    """nq independent queries on one environment in the REFERENCE'S pop order (pmp_bestfirst, one warp
    per query, all queries of a wave concurrently): costs, paths and `n_expand` (len(CLOSED), the
    benchmark's "Visited Nodes") equal what the reference's planner returns for each query.
    algo: a_star | dijkstra | gbfs | theta_star | lazy_theta_star | jps.  Paths are goal -> start
    coordinate arrays as in plan_batch.  Usable as `_plan` of plan_batch_sharded."""
    ndim = env._ndim
    starts = np.asarray(starts, dtype=np.int32).reshape(-1, ndim)
    goals = np.asarray(goals, dtype=np.int32).reshape(-1, ndim)
    nq = len(starts)
    dg = env.device_grid()
    offsets, costs = motion_table(env, ndim)
    dg.set_motions(offsets, costs)
    t = dg.torch
    n = dg.pcells
    heap_cap = int(min(len(offsets) * n + 1, 1 << 23))
    per_q = heap_cap * _lib.BESTFIRST_HEAP_REC_BYTES + 32 * n
    wave = int(max(1, min(nq, max_bytes // per_q)))
    hid = _lib.H_MANHATTAN if heuristic_type == "manhattan" else _lib.H_EUCLIDEAN
    res = BatchResult()
    res.costs = np.full(nq, np.inf)
    res.gcosts = res.costs
    res.status = np.zeros(nq, dtype=np.int32)
    res.path_len = np.zeros(nq, dtype=np.int32)
    res.n_expand = np.zeros(nq, dtype=np.int64)
    if want_paths:
        res.paths = [None] * nq
    dev = dg.device
    heap = t.empty(wave * heap_cap * _lib.BESTFIRST_HEAP_REC_BYTES, dtype=t.uint8, device=dev)
    cg = t.empty((wave, n), dtype=t.float64, device=dev)
    ch = t.empty((wave, n), dtype=t.float64, device=dev)
    cpar = t.empty((wave, n), dtype=t.int32, device=dev)
    order = t.empty((wave, n), dtype=t.int32, device=dev)
    path = t.empty((wave, n), dtype=t.int32, device=dev)
    plen_d = t.empty(wave, dtype=t.int32, device=dev)
    cost_d = t.empty(wave, dtype=t.float64, device=dev)
    nexp_d = t.empty(wave, dtype=t.int32, device=dev)
    for q0 in range(0, nq, wave):
        m = min(wave, nq - q0)
        sg = np.zeros((2, m, 3), dtype=np.int32)
        sg[0, :, :ndim] = starts[q0:q0 + m]
        sg[1, :, :ndim] = goals[q0:q0 + m]
        sg_d = t.from_numpy(sg).to(dev)
        res.h2d_bytes += sg.nbytes
        _lib.check(dg.L.pmp_bestfirst(dg.handle, _REF_ALGO[algo], hid, m, sg_d[0].data_ptr(), sg_d[1].data_ptr(),
                                      heap.data_ptr(), heap_cap, cg.data_ptr(), ch.data_ptr(), cpar.data_ptr(),
                                      order.data_ptr(), n, path.data_ptr(), n, plen_d.data_ptr(),
                                      cost_d.data_ptr(), nexp_d.data_ptr(), dg.stream()), "pmp_bestfirst")
        plen = plen_d[:m].cpu().numpy()
        if (plen < 0).any():
            raise _lib.PmpError("pmp_bestfirst: OPEN / CLOSED capacity exceeded for a query of the batch")
        res.costs[q0:q0 + m] = cost_d[:m].cpu().numpy()
        res.path_len[q0:q0 + m] = plen
        res.n_expand[q0:q0 + m] = nexp_d[:m].cpu().numpy()
        res.status[q0:q0 + m] = np.where(plen > 0, _lib.ST_REACHED, _lib.ST_EXHAUSTED)
        res.d2h_bytes += m * (8 + 4 + 4)
        if want_paths:
            mask = t.arange(n, device=dev)[None, :] < plen_d[:m, None]
            flat = path[:m][mask].cpu().numpy()
            res.d2h_bytes += flat.nbytes
            coords = engine.cells_to_coords(flat, dg)
            off = np.concatenate([[0], np.cumsum(plen)])
            for i in range(m):
                res.paths[q0 + i] = coords[off[i]:off[i + 1]]
    return res
