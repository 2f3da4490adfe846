"""GPU parity of the warp-per-tile unit-weight field solver (csrc/pmp_bfs.cu) through the public API
and the C ABI: full fields bit-exact against the oracle (the pinned restatement of Dijkstra3D run
to exhaustion on the six unit moves, dijkstra3d.py:39-126 / env3d.py:56-70), the BASELINE config 5
block at 128^3, status words, the z-slab protocol with device-side halo pulls (several slabs on
one GPU in one launch; real multi-GPU run under torchrun when >= 2 devices are visible)."""
import json
import os
import subprocess
import sys

import numpy as np
import pytest

import python_motion_planning_b200 as pmp
from python_motion_planning_b200 import _lib, engine, slab, workloads
from oracle import oracle

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def six():
    keep = oracle.motion_costs(oracle.MOTIONS_3D) == 1.0
    return oracle.MOTIONS_3D[keep], oracle.motion_costs(oracle.MOTIONS_3D)[keep]


def env6(occ):
    env = pmp.Grid3D(*occ.shape, walls=False)
    env.motions = [m for m in env.motions if m.g == 1]
    env.update(occ)
    return env


def oracle_field(occ, src):
    motions, mg = six()
    return oracle.plan(occ, src, None, algo=oracle.DIJKSTRA, motions=motions, motion_g=mg, want_field=True,
                       want_expand=False).g_field


@pytest.mark.parametrize("shape,dens,seed", [((40, 24, 20), 0.10, 0), ((48, 17, 9), 0.30, 1), ((70, 33, 41), 0.35, 2),
                                             ((33, 9, 9), 0.05, 3), ((130, 70, 50), 0.20, 4), ((16, 8, 8), 0.0, 5),
                                             ((200, 40, 24), 0.40, 6)])
def test_field_bit_exact_vs_oracle(shape, dens, seed):
    rng = np.random.default_rng(seed)
    occ = rng.random(shape) < dens
    free = np.argwhere(~occ)
    src = tuple(int(v) for v in free[rng.integers(0, len(free))])
    ref = oracle_field(occ, src)
    env = env6(occ)
    f = pmp.cost_field(env, src, weight="u32")
    assert np.array_equal(f, ref)
    # the tiled kernel of round 1 computes the same field (two independent CUDA paths agree)
    os.environ["PMP_B200_OLD_BFS"] = "1"
    try:
        assert np.array_equal(pmp.cost_field(env, src, weight="u32"), ref)
    finally:
        del os.environ["PMP_B200_OLD_BFS"]


def test_field_with_shell_walls_and_update():
    """Grid3D's own shell walls (env3d.py:72-103, with the reference's z - 1 quirk) as initialised, then two maps set through
    env.update() on the same grid: a later field must not see anything of an earlier map (lazy tile initialisation,
    workspace reuse)"""
    rng = np.random.default_rng(77)
    env = pmp.Grid3D(50, 30, 20)               # walls=True
    env.motions = [m for m in env.motions if m.g == 1]
    occ = np.asarray(env.occupancy()).astype(bool)
    assert occ[0].all() and occ[:, :, 0].all() and not occ[5, 5, 5]
    f = pmp.cost_field(env, (5, 5, 5), weight="u32")
    assert np.array_equal(f, oracle_field(occ, (5, 5, 5)))
    for k in range(2):
        occ = rng.random((50, 30, 20)) < (0.15 + 0.1 * k)
        occ[0, :, :] = occ[-1, :, :] = True
        occ[:, 0, :] = occ[:, -1, :] = True
        occ[:, :, 0] = occ[:, :, -1] = True
        env.update(occ)
        free = np.argwhere(~occ)
        src = tuple(int(v) for v in free[rng.integers(0, len(free))])
        f = pmp.cost_field(env, src, weight="u32")
        assert np.array_equal(f, oracle_field(occ, src)), k


def test_c5_block_128_bit_exact():
    """BASELINE.json config 5's parity statement: the full field bit-exact on the 128^3 copy"""
    blk, _ = workloads.replicated_boxes3d(base=128, reps=1)
    occ = workloads.replicated_slab(blk, (128, 128, 128), 0, 128)
    ref = oracle_field(occ, (1, 1, 1))
    f = pmp.cost_field(env6(occ), (1, 1, 1), weight="u32")
    assert np.array_equal(f, ref)
    assert np.isfinite(ref).sum() > 1_500_000


def test_golden_field_from_reference_64():
    """reference (Dijkstra3D, unmodified) -> fixture -> kernel on a 64^3 crop of the config 5 block"""
    p = os.path.join(ROOT, "tests", "golden", "field64.npz")
    if not os.path.exists(p):
        pytest.skip("tests/golden/field64.npz not generated")
    z = np.load(p)
    occ, src, g = z["occ"], tuple(int(v) for v in z["src"]), z["g"]
    f = pmp.cost_field(env6(occ), src, weight="u32")
    assert np.array_equal(f, g)


def test_status_words():
    """exhausted searches report PMP_ST_EXHAUSTED (full field, unreachable goal, batch)"""
    occ = np.zeros((40, 24, 20), dtype=bool)
    occ[20, :, :] = True                       # a wall: the far half is unreachable
    env = env6(occ)
    ws, _ = pmp.cost_field(env, (1, 1, 1), weight="u32", to_host=False)
    assert ws.status.cpu().tolist() == [_lib.ST_EXHAUSTED]
    ws, _ = pmp.cost_field(env, (1, 1, 1), weight="f64", to_host=False)
    assert ws.status.cpu().tolist() == [_lib.ST_EXHAUSTED]
    res = pmp.plan_batch(env, [(1, 1, 1), (1, 1, 1)], [(30, 5, 5), (10, 5, 5)], algo="dijkstra")
    assert res.status.tolist() == [_lib.ST_EXHAUSTED, _lib.ST_REACHED]
    occ[3, 3, 3] = True
    env = env6(occ)
    ws, _ = pmp.cost_field(env, (3, 3, 3), weight="u32", to_host=False)
    assert ws.status.cpu().tolist() == [_lib.ST_BAD_QUERY]
    assert not np.isfinite(pmp.cost_field(env, (3, 3, 3), weight="u32")).any()


@pytest.mark.parametrize("nslab", [2, 3, 5])
def test_slab_group_one_gpu(nslab):
    """several z-slabs on ONE GPU served by one launch: the peer protocol (halo pulls from the
    neighbour slab's field, remote arming, distributed termination) against the oracle"""
    rng = np.random.default_rng(nslab)
    shape = (70, 40, 88)
    occ = rng.random(shape) < 0.25
    free = np.argwhere(~occ)
    src = tuple(int(v) for v in free[rng.integers(0, len(free))])
    ref = oracle_field(occ, src)
    slabs = []
    for p in range(nslab):
        z0, z1 = slab.slab_range8(shape[2], p, nslab)
        slabs.append(slab.BfsSlab(occ[:, :, z0:z1], p, nslab))
    slab.solve_group(slabs, shape[2], src)
    got = np.concatenate([s.field_numpy() for s in slabs], axis=2)
    assert np.array_equal(got, ref)
    assert sum(int(s.stats_host()[7]) for s in slabs) > 0      # remote pushes happened
    for s in slabs:
        s.close()


def test_slab_multi_gpu_torchrun():
    """the real thing: one process per GPU, IPC-mapped arenas, NVLink halo pulls; certificate 0
    violations and the reached-cell count of the single-GPU run"""
    import torch
    ng = torch.cuda.device_count()
    if ng < 2:
        pytest.skip("needs >= 2 GPUs")
    n = 256
    blk, _ = workloads.replicated_boxes3d(base=128, reps=n // 128)
    occ = workloads.replicated_slab(blk, (n, n, n), 0, n)
    ws, _ = pmp.cost_field(env6(occ), (1, 1, 1), weight="u32", to_host=False)
    reached = int(engine.count_settled(ws, bounded=False).sum().item())
    world = 2 if ng < 4 else 4
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={world}",
           "--master-addr", "127.0.0.1", "--master-port", "29541", os.path.join(ROOT, "tools", "run_bfs_slab.py"),
           "--size", str(n), "--steps", "2", "--verify", "--reached", str(reached)]
    r = subprocess.run(cmd, stdout=subprocess.PIPE, stderr=subprocess.PIPE, text=True, timeout=600, cwd=ROOT)
    assert r.returncode == 0, r.stdout[-1500:] + r.stderr[-3000:]
    line = json.loads([ln for ln in r.stdout.splitlines() if ln.startswith("{")][-1])
    assert line["ok"] and line["violations"] == 0 and line["pending"] == 0 and line["reached"] == reached


def test_streamed_packed_upload_overlaps_search():
    """Grid3D.update_packed(pinned host rows) -> pmp_grid_set_packed_host_streamed: the field search starts while the
    occupancy is still arriving (k_bfs waits per tile for the chunks it needs); results equal the oracle's, twice in a
    row with different maps on the same grid, and entry points that cannot wait chunk-wise see the complete map."""
    import torch
    rng = np.random.default_rng(21)
    shape = (96, 40, 160)                    # X % 32 == 0, Y % 8 == 0, Z % 8 == 0: the rows are the grid's own layout
    env = pmp.Grid3D(*shape, walls=False)
    env.motions = [m for m in env.motions if m.g == 1]
    for k in range(2):
        occ = rng.random(shape) < (0.2 + 0.1 * k)
        free = np.argwhere(~occ)
        src = tuple(int(v) for v in free[rng.integers(0, len(free))])
        rows = torch.from_numpy(workloads.pack_rows(occ).view(np.int32)).pin_memory()
        env.update_packed(rows)
        f = pmp.cost_field(env, src, weight="u32")
        assert np.array_equal(f, oracle_field(occ, src)), k
        assert np.array_equal(env.device_grid().read_bits(), occ)          # the whole map landed
    # a planner that cannot wait chunk by chunk, right after a streamed upload
    occ = rng.random(shape) < 0.15
    env.update_packed(torch.from_numpy(workloads.pack_rows(occ).view(np.int32)).pin_memory())
    f64 = pmp.cost_field(env, src if not occ[src] else tuple(int(v) for v in np.argwhere(~occ)[0]))   # fp64 path: k_sssp
    s2 = src if not occ[src] else tuple(int(v) for v in np.argwhere(~occ)[0])
    assert np.array_equal(f64, oracle_field(occ, s2))
