"""z-slab decomposition of one large 3D cost-to-go field over several GPUs (BASELINE.json
config 5).

The reference computes a full field with `Dijkstra3D(start, unreachable_goal, env).plan()`
(global_planner/graph_search/dijkstra3d.py:39-87) in one Python process.  Here rank p owns the
planes z in [z0, z1) of the grid and keeps one ghost plane per interior side.  All ranks repeat

    run      relax every pending key below the common bound to local convergence (k_sssp);
    exchange send the ghost plane(s) -- this rank's candidate values for the neighbour's boundary
             plane -- to rank p-1 / p+1 (NCCL send/recv over NVLink; gloo in the CPU tests) and
             min-merge the plane(s) received into the own boundary plane(s), improved cells
             becoming pending;
    agree    all-reduce(min) of the smallest pending key -> next bound, or stop when no rank
             has anything pending.

The host logic is independent of where the slab lives: `SlabDriver` talks to an *engine* (begin /
run / plane / merge / refresh_plane).  `CudaSlabEngine` is the product engine (C ABI
`pmp_slab_*`); tests drive the same protocol on CPU tensors with a stand-in engine over gloo.
"""
import ctypes
import math

import numpy as np

from . import _lib, engine as _engine
from .environment import motion_table

INF = float("inf")


def slab_range(Z, rank, world):
    """owned planes [z0, z1) of `rank`: contiguous, sizes differ by at most one plane"""
    base, rem = divmod(Z, world)
    z0 = rank * base + min(rank, rem)
    return z0, z0 + base + (1 if rank < rem else 0)


def ghost_range(Z, rank, world):
    """planes held locally: the owned range plus one ghost plane per interior side"""
    z0, z1 = slab_range(Z, rank, world)
    return max(z0 - 1, 0), min(z1 + 1, Z)


class CudaSlabEngine:
    """One slab (+ ghost planes) on the current CUDA device, behind the C ABI.

    env: Grid3D of the LOCAL box (x, y, local z incl. ghosts) with its occupancy already set."""

    def __init__(self, env, weight="u32", window=0.0):
        self.env = env
        self.dg = env.device_grid()
        self.dg.set_motions(*motion_table(env, 3))
        self.weight = _lib.W_U32 if weight == "u32" else _lib.W_F64
        self.ws = _engine.Workspace(self.dg, 1, self.weight)   # private: the slab field lives across calls
        t = self.dg.torch
        self.torch = t
        self.P = _lib.SsspParams(algo=_lib.DIJKSTRA, heuristic=0, weight=self.weight, full_field=1,
                                 window=window, max_rounds=0, reserved=0)
        self.out = t.zeros(8, dtype=t.int64, device=self.dg.device)
        self.plane_cells = self.dg.padded[0] * self.dg.padded[1]
        self.launches = 0

    # -- protocol ---------------------------------------------------------------------------
    def begin(self, start_local):
        dg, ws = self.dg, self.ws
        sp = None
        if start_local is not None:
            _engine.upload_queries(ws, [tuple(start_local)], None, 3)
            sp = ws.starts.data_ptr()
        _lib.check(dg.L.pmp_slab_begin(dg.handle, ctypes.byref(self.P), sp, ws.g.data_ptr(),
                                       ws.ws.data_ptr(), ws.ws_bytes, dg.stream()), "pmp_slab_begin")
        self.launches += 1

    def _key(self, raw):
        raw = int(raw) & 0xFFFFFFFFFFFFFFFF
        if raw == 0xFFFFFFFFFFFFFFFF:
            return INF
        if self.weight == _lib.W_U32:
            return float(raw)
        return float(np.array([raw], dtype=np.uint64).view(np.float64)[0])

    def run(self, bound):
        dg, ws = self.dg, self.ws
        self.out.zero_()
        _lib.check(dg.L.pmp_slab_run(dg.handle, ctypes.byref(self.P), ws.g.data_ptr(), ws.ws.data_ptr(),
                                     ws.ws_bytes, float(bound), self.out.data_ptr(), dg.stream()),
                   "pmp_slab_run")
        self.launches += 3

    def plane(self, z_local):
        n = self.plane_cells
        return self.ws.g[0, z_local * n:(z_local + 1) * n]

    def new_plane(self):
        return self.torch.empty_like(self.plane(0))

    def merge(self, z_local, plane, mark=True):
        dg, ws = self.dg, self.ws
        _lib.check(dg.L.pmp_slab_merge(dg.handle, self.weight, ws.g.data_ptr(), ws.ws.data_ptr(),
                                       ws.ws_bytes, int(z_local), 1 if mark else 0, plane.data_ptr(),
                                       self.out.data_ptr(), dg.stream()), "pmp_slab_merge")
        self.launches += 1

    def pending(self):
        """(smallest pending key incl. merged cells, cells improved by merges) -- one D2H"""
        o = self.out.cpu()
        if int(o[5].item()):
            raise _lib.PmpError("pmp_slab_run: the search kernel aborted (queue wait timed out)")
        return self._key(o[0].item()), int(o[4].item())

    def stats(self):
        o = self.out.cpu().numpy()
        return {"rounds": int(o[1]), "tile_activations": int(o[2]), "cells_expanded": int(o[3])}

    def verify(self, source_local, z_lo, z_hi):
        dg, ws = self.dg, self.ws
        t = self.torch
        out = t.zeros(3, dtype=t.int64, device=dg.device)
        sp = None
        if source_local is not None:
            src = t.tensor(list(source_local), dtype=t.int32, device=dg.device)
            sp = src.data_ptr()
        _lib.check(dg.L.pmp_verify_field(dg.handle, self.weight, ws.g.data_ptr(), sp, int(z_lo), int(z_hi),
                                         out.data_ptr(), dg.stream()), "pmp_verify_field")
        v = out.cpu().numpy()
        return int(v[0]), int(v[1]), int(v[2])

    def field(self, z_lo, z_hi):
        """float64 numpy array [x, y, z_lo:z_hi] of the local planes (+inf = unreached)"""
        f = _engine.field_to_numpy(self.ws, 0)
        return f[:, :, z_lo:z_hi]


class SlabDriver:
    """The exchange protocol of one rank.  `dist` is torch.distributed (or None / world 1)."""

    def __init__(self, engine, Z, rank=0, world=1, dist=None, outer_window=512.0):
        self.e, self.Z, self.rank, self.world, self.dist = engine, Z, rank, world, dist
        self.z0, self.z1 = slab_range(Z, rank, world)
        self.h0, self.h1 = ghost_range(Z, rank, world)
        self.has_lo = self.h0 < self.z0          # ghost plane below / neighbour rank - 1
        self.has_hi = self.h1 > self.z1
        self.nzl = self.h1 - self.h0             # local planes incl. ghosts
        self.outer_window = outer_window
        self.iterations = 0
        self.bytes_sent = 0
        if world > 1:
            self.rx_lo = engine.new_plane() if self.has_lo else None
            self.rx_hi = engine.new_plane() if self.has_hi else None

    def local_z(self, z_global):
        return z_global - self.h0

    def owns(self, z_global):
        return self.z0 <= z_global < self.z1

    def _exchange(self, lo_plane, hi_plane):
        """send lo_plane to rank-1 and hi_plane to rank+1; receive into rx_lo / rx_hi"""
        d = self.dist
        ops = []
        if self.has_lo:
            ops.append(d.P2POp(d.isend, lo_plane, self.rank - 1))
            ops.append(d.P2POp(d.irecv, self.rx_lo, self.rank - 1))
            self.bytes_sent += lo_plane.numel() * lo_plane.element_size()
        if self.has_hi:
            ops.append(d.P2POp(d.isend, hi_plane, self.rank + 1))
            ops.append(d.P2POp(d.irecv, self.rx_hi, self.rank + 1))
            self.bytes_sent += hi_plane.numel() * hi_plane.element_size()
        if ops:
            for w in d.batch_isend_irecv(ops):
                w.wait()

    def _allreduce_min(self, key, improved):
        if self.world == 1:
            return key, improved
        import torch
        d = self.dist
        dev = self.e.plane(0).device
        t = torch.tensor([key if math.isfinite(key) else 1e300, -float(improved)], dtype=torch.float64, device=dev)
        d.all_reduce(t, op=d.ReduceOp.MIN)
        k = float(t[0].item())
        return (INF if k >= 1e300 else k), int(-t[1].item())

    def solve(self, source_global, max_iterations=1_000_000):
        """Runs to global convergence.  Returns the number of run/exchange iterations."""
        e = self.e
        sx, sy, sz = source_global
        e.begin((sx, sy, self.local_z(sz)) if self.h0 <= sz < self.h1 and self.owns(sz) else None)
        key = 0.0 if self.owns(sz) else INF
        key, _ = self._allreduce_min(key, 0)
        it = 0
        while math.isfinite(key) and it < max_iterations:
            bound = key + self.outer_window if self.world > 1 else INF
            e.run(bound)
            if self.world > 1:
                # ghost planes = this rank's candidates for the neighbours' boundary planes
                self._exchange(e.plane(0) if self.has_lo else None,
                               e.plane(self.nzl - 1) if self.has_hi else None)
                if self.has_lo:
                    e.merge(1, self.rx_lo, mark=True)              # own lowest plane
                if self.has_hi:
                    e.merge(self.nzl - 2, self.rx_hi, mark=True)   # own highest plane
            k, improved = e.pending()
            key, _ = self._allreduce_min(k, improved)
            it += 1
        self.iterations = it
        return it

    def refresh_ghosts(self):
        """make the ghost planes exact copies of the neighbours' boundary planes (values only);
        needed before the certificate is evaluated on the owned boundary planes"""
        if self.world == 1:
            return
        e = self.e
        self._exchange(e.plane(1) if self.has_lo else None,
                       e.plane(self.nzl - 2) if self.has_hi else None)
        if self.has_lo:
            e.merge(0, self.rx_lo, mark=False)
        if self.has_hi:
            e.merge(self.nzl - 1, self.rx_hi, mark=False)

    def verify(self, source_global):
        """global (violations, reached cells, pending flags) of the certificate over owned planes"""
        self.refresh_ghosts()
        sx, sy, sz = source_global
        src = (sx, sy, self.local_z(sz)) if self.owns(sz) else None
        bad, reached, pend = self.e.verify(src, self.local_z(self.z0), self.local_z(self.z1))
        if self.world > 1:
            import torch
            d = self.dist
            t = torch.tensor([bad, reached, pend], dtype=torch.int64, device=self.e.plane(0).device)
            d.all_reduce(t, op=d.ReduceOp.SUM)
            bad, reached, pend = (int(v) for v in t.tolist())
        return bad, reached, pend


def build_local_env(occ_local, motions="6", packed_shape=None):
    """Grid3D of the local box from a dense occupancy array [x, y, z_local] -- or, with
    packed_shape = (X, Y, Z_local), from bit rows [z][y][X/32] (Grid3D.update_packed); `motions`:
    "6" keeps the unit moves only (config 5), "26" the full table."""
    from .environment import Grid3D
    env = Grid3D(*(packed_shape or occ_local.shape), walls=False)
    if motions == "6":
        env.motions = [m for m in env.motions if m.g == 1]
    elif motions == "18":
        env.motions = [m for m in env.motions if m.g < 1.5]
    if packed_shape:
        env.update_packed(occ_local)
    else:
        env.update(occ_local)
    return env


def solve_in_process(engines, Z, source_global, outer_window=512.0, max_iterations=1_000_000):
    """The same protocol with every slab held by ONE process (slabs of a grid too large for one
    field buffer processed in turn on one GPU, or the single-GPU check of the multi-GPU path):
    `engines[p]` holds slab p of `len(engines)`; planes are handed over by device copies instead
    of send/recv.  Launches are strictly sequential -- no kernel ever waits for another.
    Returns (iterations, drivers)."""
    world = len(engines)
    drv = [SlabDriver(e, Z, p, world, None, outer_window) for p, e in enumerate(engines)]
    sx, sy, sz = source_global
    for d in drv:
        d.e.begin((sx, sy, d.local_z(sz)) if d.owns(sz) else None)
    key, it = 0.0, 0
    while math.isfinite(key) and it < max_iterations:
        bound = key + outer_window if world > 1 else INF
        for d in drv:
            d.e.run(bound)
        for p in range(world - 1):      # boundary between slab p and p + 1
            lo, hi = drv[p], drv[p + 1]
            up = lo.e.plane(lo.nzl - 1).clone()   # lo's candidates for hi's lowest owned plane
            dn = hi.e.plane(0).clone()            # hi's candidates for lo's highest owned plane
            hi.e.merge(1, up, mark=True)
            lo.e.merge(lo.nzl - 2, dn, mark=True)
        key = min(d.e.pending()[0] for d in drv)
        it += 1
    for p in range(world - 1):          # exact ghosts for the certificate
        lo, hi = drv[p], drv[p + 1]
        hi.e.merge(0, lo.e.plane(lo.nzl - 2).clone(), mark=False)
        lo.e.merge(lo.nzl - 1, hi.e.plane(1).clone(), mark=False)
    return it, drv


# ------------------------------------------------------------------------------------------------
# Round 2: z-slabs with DEVICE-SIDE halo exchange (pmp_bfs_slab_*, csrc/pmp_bfs.cu).
#
# For the unit-weight 6-connected field (config 5) the host-driven loop above is replaced by one
# kernel launch per rank: boundary tiles read their z halo out of the neighbour GPU's field over
# NVLink (CUDA IPC mapping of the peer's arena) and queue the neighbour's tiles with remote
# atomics.  torch.distributed only carries the 128-byte (IPC handle, layout) records once and the
# two barriers around a search.
# ------------------------------------------------------------------------------------------------
def slab_range8(Z, rank, world):
    """owned planes [z0, z1) of `rank` when slab boundaries fall on tile rows (multiples of 8)"""
    ntz = (Z + 7) // 8
    t0, t1 = ntz * rank // world, ntz * (rank + 1) // world
    return t0 * 8, min(Z, t1 * 8)


class _DevMem:
    """a device allocation of the library, viewed by torch without a copy"""

    def __init__(self, ptr, n_words, owner):
        self.__cuda_array_interface__ = {"shape": (int(n_words),), "typestr": "<i4", "data": (int(ptr), False),
                                         "version": 2}
        self._owner = owner


class BfsSlab:
    """One z-slab (owned planes only, no ghosts) of a 6-connected unit-weight grid on the current
    CUDA device.  occ_local: dense occupancy [x, y, z_local]."""

    def __init__(self, occ_local, rank=0, world=1, packed_shape=None):
        self.env = build_local_env(occ_local, "6", packed_shape)
        self.dg = self.env.device_grid()
        self.dg.set_motions(*motion_table(self.env, 3))
        self.rank, self.world = rank, world
        t = self.torch = self.dg.torch
        L = self.L = self.dg.L
        h = ctypes.c_void_p()
        _lib.check(L.pmp_bfs_slab_create(self.dg.handle, rank, world, ctypes.byref(h)), "pmp_bfs_slab_create")
        self.handle = h
        fp, sp, cp, nb = ctypes.c_void_p(), ctypes.c_void_p(), ctypes.c_void_p(), ctypes.c_int64(0)
        _lib.check(L.pmp_bfs_slab_buffers(h, ctypes.byref(fp), ctypes.byref(nb), ctypes.byref(sp), ctypes.byref(cp)),
                   "pmp_bfs_slab_buffers")
        self.field = t.as_tensor(_DevMem(fp.value, nb.value // 4, self), device=self.dg.device)   # raw words
        self._stats_arena = t.as_tensor(_DevMem(sp.value, 64, self), device=self.dg.device).view(t.int64)   # 32 counters
        self.stats = t.zeros(16, dtype=t.int64, device=self.dg.device)
        self.status = t.zeros(1, dtype=t.int32, device=self.dg.device)
        self._start = t.zeros(3, dtype=t.int32, device=self.dg.device)
        self.launches = 0

    def close(self):
        if self.handle is not None:
            self.field = None
            self._stats_arena = None
            self.L.pmp_bfs_slab_destroy(self.handle)
            self.handle = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def export(self):
        hb = (ctypes.c_uint8 * 64)()
        info = (ctypes.c_int64 * 8)()
        _lib.check(self.L.pmp_bfs_slab_export(self.handle, hb, info), "pmp_bfs_slab_export")
        return bytes(hb), [int(v) for v in info]

    def connect(self, handles, infos):
        """handles: list of 64-byte strings by rank; infos: list of 8-int lists by rank"""
        hb = (ctypes.c_uint8 * (64 * self.world)).from_buffer_copy(b"".join(handles))
        ib = (ctypes.c_int64 * (8 * self.world))(*[v for inf in infos for v in inf])
        _lib.check(self.L.pmp_bfs_slab_connect(self.handle, hb, ib), "pmp_bfs_slab_connect")

    def begin(self, start_local):
        sp = None
        if start_local is not None:
            self._start.copy_(self.torch.tensor(list(start_local), dtype=self.torch.int32))
            sp = self._start.data_ptr()
        self.status.zero_()
        _lib.check(self.L.pmp_bfs_slab_begin(self.handle, sp, self.status.data_ptr(), self.dg.stream()),
                   "pmp_bfs_slab_begin")
        self.launches += 1

    def run(self):
        _lib.check(self.L.pmp_bfs_slab_run(self.handle, self.status.data_ptr(), self.stats.data_ptr(),
                                           self.dg.stream()), "pmp_bfs_slab_run")
        self.launches += 3

    def check(self):
        _lib.check_status(self.status.cpu().numpy(), "pmp_bfs_slab_run")

    def stats_host(self):
        """the 16 counters of the last search, read from the arena (layout of pmp_sssp's stats)"""
        v = self._stats_arena.cpu().numpy()
        if int(v[15]):
            raise _lib.PmpError("pmp_bfs_slab: the search kernel aborted (queue wait timed out)")
        return v

    def plane(self, z_local):
        n = self.dg.padded[0] * self.dg.padded[1]
        return self.field[z_local * n:(z_local + 1) * n]

    def field_numpy(self):
        """float64 [x, y, z_local] (+inf = unreached)"""
        t = self.torch
        PX, PY, PZ = self.dg.padded
        raw = self.field.clone()
        f = (raw >> 1).to(t.float64)
        f = t.where(raw == -1, t.full_like(f, float("inf")), f).view(PZ, PY, PX)
        X, Y, Z = self.dg.dims
        return f[:Z, :Y, :X].permute(2, 1, 0).contiguous().cpu().numpy()


def solve_group(slabs, Z, source_global):
    """several BfsSlab objects of ONE process / one GPU (slab p = rank p of len(slabs)): one kernel
    launch serves all of them, with the peer protocol running between them."""
    n = len(slabs)
    L = slabs[0].L
    arr = (ctypes.c_void_p * n)(*[s.handle for s in slabs])
    _lib.check(L.pmp_bfs_slab_connect_local(arr, n), "pmp_bfs_slab_connect_local")
    sx, sy, sz = source_global
    for p, s in enumerate(slabs):
        z0, z1 = slab_range8(Z, p, n)
        s.begin((sx, sy, sz - z0) if z0 <= sz < z1 else None)
    _lib.check(L.pmp_bfs_slab_run_group(arr, n, slabs[0].dg.stream()), "pmp_bfs_slab_run_group")


def gather_slab_records(dist, handle64, info8, device="cpu"):
    """all-gather the (64-byte IPC handle, 8 x int64 layout) records of every rank's slab arena, ordered by rank --
    the only data the ranks exchange through torch.distributed for the device-side halo protocol"""
    import torch
    rec = np.zeros(128, dtype=np.uint8)
    rec[:64] = np.frombuffer(handle64, dtype=np.uint8)
    rec[64:] = np.array(info8, dtype=np.int64).view(np.uint8)
    mine = torch.from_numpy(rec).to(device)
    allr = [torch.empty_like(mine) for _ in range(dist.get_world_size())]
    dist.all_gather(allr, mine)
    recs = [a.cpu().numpy() for a in allr]
    return [r[:64].tobytes() for r in recs], [r[64:].view(np.int64).tolist() for r in recs]


class BfsSlabSolver:
    """The unit-weight field of one large grid over `world` GPUs, one process per GPU.

    occ_planes(z0, z1) -> dense occupancy [x, y, z0:z1] of the GLOBAL grid of `shape` (packed=True:
    bit rows [z1 - z0][Y][X/32], Grid3D.update_packed);
    dist: torch.distributed (initialised; NCCL on the GPU box) or None for world 1."""

    def __init__(self, occ_planes, shape, rank=0, world=1, dist=None, packed=False):
        self.shape, self.rank, self.world, self.dist = tuple(shape), rank, world, dist
        self.occ_planes, self.packed = occ_planes, packed
        self.z0, self.z1 = slab_range8(shape[2], rank, world)
        self.slab = BfsSlab(occ_planes(self.z0, self.z1), rank, world,
                            (shape[0], shape[1], self.z1 - self.z0) if packed else None)
        self.torch = self.slab.torch
        if world > 1:
            h, info = self.slab.export()
            dev = self.slab.dg.device if dist.get_backend() == "nccl" else "cpu"
            handles, infos = gather_slab_records(dist, h, info, dev)
            self.slab.connect(handles, infos)
        self.bytes_halo = 0

    def _barrier(self):
        self.torch.cuda.synchronize()
        if self.world > 1:
            self.dist.barrier()

    def solve(self, source_global):
        """one search to global termination; returns nothing (the field stays in self.slab.field)"""
        sx, sy, sz = source_global
        self.slab.begin((sx, sy, sz - self.z0) if self.z0 <= sz < self.z1 else None)
        self._barrier()          # every rank has reset its workspace before any rank searches
        self.slab.run()
        self._barrier()          # ... and has finished before any rank resets again

    def check(self):
        self.slab.check()

    def verify(self, source_global):
        """global (violations, reached cells, pending flags): the local-optimality certificate
        (pmp_verify_field) over the owned planes, evaluated on a copy with one ghost plane per
        interior side fetched from the neighbour ranks (send / recv; test / accounting path only)."""
        t, s = self.torch, self.slab
        Z = self.shape[2]
        h0, h1 = max(self.z0 - 1, 0), min(self.z1 + 1, Z)
        env_g = build_local_env(self.occ_planes(h0, h1), "6", (self.shape[0], self.shape[1], h1 - h0) if self.packed else None)
        dg = env_g.device_grid()
        dg.set_motions(*motion_table(env_g, 3))
        PX, PY, PZg = dg.padded
        n = PX * PY
        f = t.full((PZg * n,), -1, dtype=t.int32, device=dg.device)
        off = self.z0 - h0
        nz = self.z1 - self.z0
        f[off * n:(off + nz) * n] = s.field[:nz * n]
        if self.world > 1:
            d = self.dist
            ops = []
            lo_rx = hi_rx = None
            if self.rank > 0:
                lo_rx = t.empty(n, dtype=t.int32, device=dg.device)
                ops += [d.P2POp(d.isend, s.plane(0).contiguous(), self.rank - 1), d.P2POp(d.irecv, lo_rx, self.rank - 1)]
            if self.rank < self.world - 1:
                hi_rx = t.empty(n, dtype=t.int32, device=dg.device)
                ops += [d.P2POp(d.isend, s.plane(nz - 1).contiguous(), self.rank + 1), d.P2POp(d.irecv, hi_rx, self.rank + 1)]
            for w in d.batch_isend_irecv(ops):
                w.wait()
            if lo_rx is not None:
                f[0:n] = lo_rx
            if hi_rx is not None:
                f[(off + nz) * n:(off + nz + 1) * n] = hi_rx
        out = t.zeros(3, dtype=t.int64, device=dg.device)
        sx, sy, sz = source_global
        sp = None
        if self.z0 <= sz < self.z1:
            src = t.tensor([sx, sy, sz - h0], dtype=t.int32, device=dg.device)
            sp = src.data_ptr()
        _lib.check(dg.L.pmp_verify_field(dg.handle, _lib.W_U32, f.data_ptr(), sp, off, off + nz, out.data_ptr(),
                                         dg.stream()), "pmp_verify_field")
        if self.world > 1:
            self.dist.all_reduce(out, op=self.dist.ReduceOp.SUM)
        bad, reached, pend = (int(v) for v in out.cpu().tolist())
        return bad, reached, pend
