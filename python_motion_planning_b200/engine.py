"""Host-side driver of the CUDA search kernels: buffers (torch tensors as carriers), the
calls into libpmp_b200.so, and conversion of results.  No arithmetic of the search happens
here."""
import ctypes
import os

import numpy as np

from . import _lib
from ._lib import ASTAR, DIJKSTRA, W_F64, W_U32, SsspParams  # noqa: F401


class Workspace:
    """Reusable device buffers for nq concurrent searches on one DeviceGrid."""

    def __init__(self, dg, nq, weight):
        t = dg.torch
        self.dg, self.nq, self.weight = dg, nq, weight
        gb, wb = ctypes.c_int64(0), ctypes.c_int64(0)
        _lib.check(dg.L.pmp_sssp_sizes(dg.handle, nq, weight, ctypes.byref(gb), ctypes.byref(wb)),
                   "pmp_sssp_sizes")
        self.g_bytes_per_query, self.ws_bytes = gb.value, wb.value
        dt = t.float64 if weight == W_F64 else t.int32
        self.g = t.empty((nq, dg.pcells), dtype=dt, device=dg.device)
        self.ws = t.empty(self.ws_bytes, dtype=t.uint8, device=dg.device)
        self.gcost = t.empty(nq, dtype=t.float64, device=dg.device)
        self.status = t.empty(nq, dtype=t.int32, device=dg.device)
        self.stats = t.zeros(16, dtype=t.int64, device=dg.device)
        self.starts = t.empty((nq, 3), dtype=t.int32, device=dg.device)
        self.goals = t.empty((nq, 3), dtype=t.int32, device=dg.device)
        self.starts_host = t.empty((nq, 3), dtype=t.int32).pin_memory()
        self.goals_host = t.empty((nq, 3), dtype=t.int32).pin_memory()
        self.path = None
        self.path_cap = 0
        self.path_len = t.empty(nq, dtype=t.int32, device=dg.device)
        self.cost = t.empty(nq, dtype=t.float64, device=dg.device)
        # pinned landing buffers of the single-query result: scalars and the head of the path in one sync
        self.host_res = t.empty(4, dtype=t.float64).pin_memory()
        self.host_path = t.empty(16384, dtype=t.int32).pin_memory()

    def ensure_path(self, cap):
        if self.path is None or self.path_cap < cap:
            t = self.dg.torch
            self.path = t.empty((self.nq, cap), dtype=t.int32, device=self.dg.device)
            self.path_cap = cap


_small_pool = {}   # (device, padded dims, nq, weight) -> Workspace, for grids whose field is snapshotted


def get_workspace(dg, nq, weight):
    """Buffers for nq searches on dg, reused across calls.  Small grids (<= 4 Mi cells, where
    plan() hands out a private copy of the field for `expand`) share one pool across
    environments of the same shape, so that planning on a freshly built Grid does not pay for
    device and pinned-host allocations again; large grids keep their buffers per DeviceGrid."""
    if dg.pcells <= (1 << 22):
        key = (str(dg.device), dg.ndim, dg.padded, nq, weight)
        ws = _small_pool.get(key)
        if ws is None:
            if len(_small_pool) >= 8:
                _small_pool.clear()
            ws = _small_pool[key] = Workspace(dg, nq, weight)
        ws.dg = dg
        return ws
    cache = dg.__dict__.setdefault("_ws_cache", {})
    key = (nq, weight)
    ws = cache.get(key)
    if ws is None:
        if len(cache) >= 2:      # keep the footprint bounded
            cache.clear()
        ws = cache[key] = Workspace(dg, nq, weight)
    return ws


def _coords3(a, nq, ndim):
    out = np.zeros((nq, 3), dtype=np.int32)
    out[:, :ndim] = np.asarray(a, dtype=np.int32).reshape(nq, ndim)
    return out


def upload_queries(ws, starts, goals, ndim):
    """pinned host -> device copy of the start / goal coordinates; returns bytes copied"""
    t = ws.dg.torch
    nq = ws.nq
    ws.starts_host.copy_(t.from_numpy(_coords3(starts, nq, ndim)))
    ws.starts.copy_(ws.starts_host, non_blocking=True)
    n = ws.starts_host.numel() * 4
    if goals is not None:
        ws.goals_host.copy_(t.from_numpy(_coords3(goals, nq, ndim)))
        ws.goals.copy_(ws.goals_host, non_blocking=True)
        n += ws.goals_host.numel() * 4
    return n


def default_window(ndim):
    v = os.environ.get("PMP_B200_WINDOW")
    return float(v) if v else 0.0


def sssp(ws, algo, heuristic=0, full_field=False, window=0.0, max_rounds=0, have_goals=True):
    """Launch the search for the queries already uploaded into ws.starts / ws.goals."""
    dg = ws.dg
    P = SsspParams(algo=algo, heuristic=heuristic, weight=ws.weight, full_field=1 if full_field else 0,
                   window=window or default_window(dg.ndim), max_rounds=max_rounds,
                   reserved=int(os.environ.get("PMP_B200_SUB", "0") or 0))
    goals_ptr = ws.goals.data_ptr() if have_goals else None
    _lib.check(dg.L.pmp_sssp(dg.handle, ctypes.byref(P), ws.nq, ws.starts.data_ptr(), goals_ptr,
                             ws.g.data_ptr(), ws.ws.data_ptr(), ws.ws_bytes, ws.gcost.data_ptr(),
                             ws.status.data_ptr(), ws.stats.data_ptr(), dg.stream()), "pmp_sssp")


def check_field_status(ws):
    """synchronising check of the status words of the last search in ws: raises PmpError when the
    kernel aborted (PMP_ST_ERROR) -- for callers that keep the field on the device"""
    _lib.check_status(ws.status.cpu().numpy(), "search")


def extract_paths(ws, path_cap):
    dg = ws.dg
    ws.ensure_path(path_cap)
    _lib.check(dg.L.pmp_extract_paths(dg.handle, ws.weight, ws.nq, ws.starts.data_ptr(),
                                      ws.goals.data_ptr(), ws.g.data_ptr(), ws.path.data_ptr(),
                                      ws.path_cap, ws.path_len.data_ptr(), ws.cost.data_ptr(),
                                      dg.stream()), "pmp_extract_paths")


def count_settled(ws, bounded, use_h=True):
    dg = ws.dg
    t = dg.torch
    counts = t.zeros(ws.nq, dtype=t.int64, device=dg.device)
    out = []
    for q0 in range(0, ws.nq, 65535):     # grid.y limit
        n = min(65535, ws.nq - q0)
        esz = 8 if ws.weight == W_F64 else 4
        _lib.check(dg.L.pmp_count_settled(
            dg.handle, ws.weight, n, ws.g.data_ptr() + q0 * dg.pcells * esz,
            (ws.goals.data_ptr() + q0 * 12) if bounded else None,
            (ws.gcost.data_ptr() + q0 * 8) if bounded else None, 0 if use_h else -1,
            counts.data_ptr() + q0 * 8, dg.stream()), "pmp_count_settled")
    return counts


def cells_to_coords(cells, dg):
    """padded cell indices (uint32 carried as int32) -> (n, ndim) coordinates"""
    c = np.asarray(cells).astype(np.int64) & 0xFFFFFFFF
    PX, PY, _ = dg.padded
    x = c % PX
    y = (c // PX) % PY
    if dg.ndim == 2:
        return np.stack([x, y], axis=1)
    return np.stack([x, y, c // (PX * PY)], axis=1)


def cells_to_points(cells, dg):
    """padded cell indices -> list of coordinate tuples of Python ints (the reference's path form)"""
    return list(map(tuple, cells_to_coords(cells, dg).tolist()))


def field_to_numpy(ws, q=0, raw=None):
    """cost field of query q as a float64 array indexed [x, y(, z)] (+inf = unreached);
    `raw`: a private device copy of the field words to decode instead of ws.g[q]"""
    dg = ws.dg
    t = dg.torch
    PX, PY, PZ = dg.padded
    raw = ws.g[q].clone() if raw is None else raw
    _lib.check(dg.L.pmp_field_finalize(raw.data_ptr(), ws.weight, raw.numel(), dg.stream()),
               "pmp_field_finalize")
    if ws.weight == W_F64:
        f = raw
    else:
        f = raw.to(t.float64)
        f = t.where(raw == -1, t.full_like(f, float("inf")), f)
    f = f.view(PZ, PY, PX)
    X, Y = dg.dims[0], dg.dims[1]
    if dg.ndim == 2:
        return f[0, :Y, :X].T.contiguous().cpu().numpy()
    return f[:dg.dims[2], :Y, :X].permute(2, 1, 0).contiguous().cpu().numpy()
