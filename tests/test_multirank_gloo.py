"""world_size-2/3 CPU tests (gloo) of the multi-GPU host logic: the z-slab exchange protocol
(python_motion_planning_b200/slab.py) and the query sharding of plan_batch_sharded.

No CUDA here: the slab protocol is driven with a stand-in engine that keeps its slab in numpy
(a plain heap Dijkstra restricted to the local box), the batch sharding with a stand-in planner
that answers from the oracle.  What is under test is the partitioning, the plane exchange, the
merge / termination agreement and the gather -- the parts that do not run on one GPU.
"""
import heapq
import os
import socket
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


class NumpySlabEngine:
    """stand-in for CudaSlabEngine: 6-connected unit-weight Dijkstra on the local box"""

    def __init__(self, occ_local):
        import torch
        self.torch = torch
        self.occ = np.ascontiguousarray(occ_local.transpose(2, 1, 0))      # [z, y, x]
        self.g = np.full(self.occ.shape, np.inf)
        self.heap = []
        self.improved = 0
        self.runs = 0

    def begin(self, start_local):
        self.g[:] = np.inf
        self.heap = []
        if start_local is not None:
            x, y, z = start_local
            self.g[z, y, x] = 0.0
            heapq.heappush(self.heap, (0.0, z, y, x))

    def run(self, bound):
        self.improved = 0
        self.runs += 1
        Z, Y, X = self.g.shape
        while self.heap and self.heap[0][0] < bound:
            k, z, y, x = heapq.heappop(self.heap)
            if k > self.g[z, y, x]:
                continue
            for dz, dy, dx in ((0, 0, 1), (0, 0, -1), (0, 1, 0), (0, -1, 0), (1, 0, 0), (-1, 0, 0)):
                a, b, c = z + dz, y + dy, x + dx
                if 0 <= a < Z and 0 <= b < Y and 0 <= c < X and not self.occ[a, b, c] and k + 1 < self.g[a, b, c]:
                    self.g[a, b, c] = k + 1
                    heapq.heappush(self.heap, (k + 1, a, b, c))

    def plane(self, z_local):
        return self.torch.from_numpy(self.g[z_local].reshape(-1))

    def new_plane(self):
        return self.torch.empty(self.g.shape[1] * self.g.shape[2], dtype=self.torch.float64)

    def merge(self, z_local, plane, mark=True):
        p = plane.numpy().reshape(self.g.shape[1:])
        better = (p < self.g[z_local]) & ~self.occ[z_local]
        self.g[z_local][better] = p[better]
        if mark:
            self.improved += int(better.sum())
            for y, x in np.argwhere(better):
                heapq.heappush(self.heap, (float(p[y, x]), z_local, int(y), int(x)))

    def pending(self):
        while self.heap and self.heap[0][0] > self.g[self.heap[0][1], self.heap[0][2], self.heap[0][3]]:
            heapq.heappop(self.heap)
        return (self.heap[0][0] if self.heap else float("inf")), self.improved

    def verify(self, source_local, z_lo, z_hi):
        return 0, int(np.isfinite(self.g[z_lo:z_hi]).sum()), 0


def _make_grid():
    rng = np.random.default_rng(4)
    occ = np.zeros((14, 12, 23), dtype=bool)
    for ax in range(3):
        sl = [slice(None)] * 3
        sl[ax] = 0
        occ[tuple(sl)] = True
        sl[ax] = -1
        occ[tuple(sl)] = True
    for _ in range(30):
        c = [int(rng.integers(1, occ.shape[i] - 2)) for i in range(3)]
        e = rng.integers(1, 5, size=3)
        occ[c[0]:c[0] + e[0], c[1]:c[1] + e[1], c[2]:c[2] + e[2]] = True
    occ[1:3, 1:3, 1:3] = False
    # a wall with one hole forces paths to cross the slab boundaries back and forth
    occ[1:-1, 6, 1:-1] = True
    occ[2, 6, 20] = False
    occ[11, 6, 2] = False
    for ax in range(3):
        sl = [slice(None)] * 3
        sl[ax] = 0
        occ[tuple(sl)] = True
        sl[ax] = -1
        occ[tuple(sl)] = True
    return occ


def _slab_worker(rank, world, port, outer_window, q):
    sys.path.insert(0, ROOT)
    import torch.distributed as dist
    from python_motion_planning_b200 import slab
    dist.init_process_group("gloo", init_method=f"tcp://127.0.0.1:{port}", rank=rank, world_size=world)
    try:
        occ = _make_grid()
        Z = occ.shape[2]
        h0, h1 = slab.ghost_range(Z, rank, world)
        eng = NumpySlabEngine(occ[:, :, h0:h1])
        drv = slab.SlabDriver(eng, Z, rank, world, dist, outer_window=outer_window)
        its = drv.solve((1, 1, 1))
        bad, reached, pend = drv.verify((1, 1, 1))
        z0, z1 = drv.z0, drv.z1
        own = eng.g[drv.local_z(z0):drv.local_z(z1)].transpose(2, 1, 0).copy()   # [x, y, z]
        ghosts_exact = True
        q.put((rank, z0, z1, own, its, reached, eng.runs, drv.bytes_sent))
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("world,outer_window", [(2, 4.0), (3, 1e9), (2, 1.0)])
def test_slab_protocol_matches_oracle(world, outer_window):
    import torch.multiprocessing as mp
    from oracle import oracle
    occ = _make_grid()
    motions = oracle.MOTIONS_3D[oracle.motion_costs(oracle.MOTIONS_3D) == 1.0]
    ref = oracle.plan(occ, (1, 1, 1), None, algo=oracle.DIJKSTRA, motions=motions,
                      motion_g=np.ones(len(motions)), want_field=True, want_expand=False).g_field
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_slab_worker, args=(r, world, port, outer_window, q)) for r in range(world)]
    for p in procs:
        p.start()
    got = [q.get(timeout=120) for _ in range(world)]
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    field = np.full(occ.shape, np.inf)
    reached = 0
    its = set()
    for rank, z0, z1, own, it, r, runs, sent in got:
        field[:, :, z0:z1] = own
        reached = r
        its.add(it)
        assert runs == it
    assert len(its) == 1                     # every rank did the same number of exchanges
    assert np.array_equal(field, ref)        # bit-exact (integer weights)
    assert reached == int(np.isfinite(ref).sum())
    assert np.isfinite(ref[:, 7:, :]).any()  # the far side of the wall was reached through the holes


def test_slab_ranges_cover_grid():
    from python_motion_planning_b200 import slab
    for Z in (5, 16, 23, 1536):
        for world in (1, 2, 3, 8):
            if world > Z:
                continue
            nxt = 0
            for r in range(world):
                z0, z1 = slab.slab_range(Z, r, world)
                h0, h1 = slab.ghost_range(Z, r, world)
                assert z0 == nxt and z1 > z0
                assert h0 == max(z0 - 1, 0) and h1 == min(z1 + 1, Z)
                nxt = z1
            assert nxt == Z


def _batch_worker(rank, world, port, q):
    sys.path.insert(0, ROOT)
    import torch.distributed as dist
    import python_motion_planning_b200 as pmp
    from python_motion_planning_b200.batch import BatchResult
    from oracle import oracle
    dist.init_process_group("gloo", init_method=f"tcp://127.0.0.1:{port}", rank=rank, world_size=world)
    try:
        rng = np.random.default_rng(0)
        occ = rng.random((40, 40)) < 0.25
        occ[0, :] = occ[-1, :] = True
        occ[:, 0] = occ[:, -1] = True
        free = np.argwhere(~occ)
        idx = np.random.default_rng(1).integers(0, len(free), size=(11, 2))
        starts, goals = free[idx[:, 0]], free[idx[:, 1]]

        def fake_plan(env, s, g, **kw):          # stands in for the CUDA plan_batch of one rank
            cost, nexp, plen, flags = oracle.plan_batch(occ, s, g, algo=oracle.DIJKSTRA)
            r = BatchResult()
            r.costs, r.gcosts = cost, cost.copy()
            r.status = np.where(np.isfinite(cost), 2, 1).astype(np.int32)
            r.path_len = plen.astype(np.int32)
            r.paths = [np.zeros((int(n), 2), dtype=np.int64) for n in plen]
            return r

        env = pmp.Grid(40, 40)
        env.update(occ)
        res = pmp.plan_batch_sharded(env, starts, goals, dist=dist, _plan=fake_plan)
        q.put((rank, res.shard, res.costs, res.status, res.path_len,
               [i for i, p in enumerate(res.paths) if p is not None]))
    finally:
        dist.destroy_process_group()


def test_batch_sharding_gathers_every_query():
    import torch.multiprocessing as mp
    from oracle import oracle
    world = 2
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_batch_worker, args=(r, world, port, q)) for r in range(world)]
    for p in procs:
        p.start()
    got = sorted([q.get(timeout=120) for _ in range(world)], key=lambda t: t[0])
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    rng = np.random.default_rng(0)
    occ = rng.random((40, 40)) < 0.25
    occ[0, :] = occ[-1, :] = True
    occ[:, 0] = occ[:, -1] = True
    free = np.argwhere(~occ)
    idx = np.random.default_rng(1).integers(0, len(free), size=(11, 2))
    ref, _, ref_len, _ = oracle.plan_batch(occ, free[idx[:, 0]], free[idx[:, 1]], algo=oracle.DIJKSTRA)
    assert got[0][1] == (0, 6) and got[1][1] == (6, 11)
    for rank, shard, costs, status, plen, have_paths in got:
        assert np.array_equal(costs, ref)                    # every rank holds all answers
        assert np.array_equal(plen, ref_len.astype(np.int32))
        assert have_paths == list(range(*shard))             # paths only for the own shard


# ---- round 2: host logic of the device-side halo protocol (slab.BfsSlabSolver) ----
def _records_worker(rank, world, port, q):
    sys.path.insert(0, ROOT)
    import torch.distributed as dist
    from python_motion_planning_b200 import slab
    dist.init_process_group("gloo", init_method=f"tcp://127.0.0.1:{port}", rank=rank, world_size=world)
    try:
        handle = bytes([(rank * 7 + i) % 251 for i in range(64)])          # stands in for a cudaIpcMemHandle_t
        z0, z1 = slab.slab_range8(200, rank, world)
        info = [48, 192, (z1 - z0 + 7) // 8, z1 - z0, 1000 + rank, 2000 + rank, 1536, 1536]
        handles, infos = slab.gather_slab_records(dist, handle, info, "cpu")
        q.put((rank, handles, infos))
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("world", [2, 3])
def test_slab_record_exchange_is_ordered_by_rank(world):
    """every rank ends up with every rank's (IPC handle, layout) record, ordered by rank: what
    pmp_bfs_slab_connect needs; and the 8-aligned slab ranges tile the grid"""
    import torch.multiprocessing as mp
    from python_motion_planning_b200 import slab
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_records_worker, args=(r, world, port, q)) for r in range(world)]
    for p in procs:
        p.start()
    got = [q.get(timeout=120) for _ in range(world)]
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    for rank, handles, infos in got:
        assert len(handles) == world and len(infos) == world
        for r in range(world):
            assert handles[r] == bytes([(r * 7 + i) % 251 for i in range(64)])
            z0, z1 = slab.slab_range8(200, r, world)
            assert infos[r][3] == z1 - z0 and infos[r][4] == 1000 + r
    for Z in (8, 64, 200, 1536, 1537):
        for w in (1, 2, 3, 5, 8):
            if (Z + 7) // 8 < w:
                continue
            edges = [slab.slab_range8(Z, r, w) for r in range(w)]
            assert edges[0][0] == 0 and edges[-1][1] == Z
            for (a0, a1), (b0, b1) in zip(edges[:-1], edges[1:]):
                assert a1 == b0 and a1 % 8 == 0 and a1 > a0
