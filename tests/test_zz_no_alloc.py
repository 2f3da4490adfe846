"""pmp_cost_field_u32_static -- the entry point that allocates nothing (workspace in a static device
array inside the library; include/pmp_b200.h).  Called through the C ABI and compared, bit for bit
(integer work: exact), with the existing entry point pmp_sssp(full_field, PMP_W_U32) on the same
inputs, on two calls in a row (the array keeps the previous call's contents).  Collected last on
purpose."""
import ctypes

import numpy as np
import pytest

import python_motion_planning_b200 as pmp
from python_motion_planning_b200 import _lib, engine
from python_motion_planning_b200.environment import motion_table

pytestmark = pytest.mark.gpu


def env6(occ):
    env = pmp.Grid3D(*occ.shape, walls=False)
    env.motions = [m for m in env.motions if m.g == 1]
    env.update(occ)
    return env


def test_static_workspace_field_equals_pmp_sssp_twice():
    import torch
    rng = np.random.default_rng(7)
    occ = rng.random((150, 70, 50)) < 0.25
    free = np.argwhere(~occ)
    env = env6(occ)
    dg = env.device_grid()
    dg.set_motions(*motion_table(env, 3))
    L = dg.L
    out = torch.empty(dg.pcells, dtype=torch.int32, device=dg.device)          # caller-owned output
    start = torch.zeros(3, dtype=torch.int32, device=dg.device)
    status = torch.zeros(1, dtype=torch.int32, device=dg.device)
    stats = torch.zeros(16, dtype=torch.int64, device=dg.device)
    ws = engine.get_workspace(dg, 1, _lib.W_U32)
    for k in (0, 1):                                                           # two calls in a row
        src = tuple(int(v) for v in free[rng.integers(0, len(free))])
        start.copy_(torch.tensor(src, dtype=torch.int32))
        out.fill_(12345)                                                       # garbage: every word must be written
        _lib.check(L.pmp_cost_field_u32_static(dg.handle, start.data_ptr(), out.data_ptr(), status.data_ptr(),
                                               stats.data_ptr(), dg.stream()), "pmp_cost_field_u32_static")
        # the existing entry point on the same inputs
        engine.upload_queries(ws, [src], None, 3)
        engine.sssp(ws, _lib.DIJKSTRA, 0, full_field=True, have_goals=False)
        torch.cuda.synchronize()
        assert status.cpu().tolist() == [_lib.ST_EXHAUSTED]
        assert int(stats[15].item()) == 0
        assert torch.equal(out, ws.g[0]), f"call {k}: field differs from pmp_sssp's"
        assert int((out != -1).sum().item()) > 100_000


def test_static_workspace_rejects_large_grids_and_bad_motions():
    env = pmp.Grid3D(1024, 512, 520, walls=False)              # 133 120 tiles: beyond the 1 MiB array
    env.motions = [m for m in env.motions if m.g == 1]
    env.update(set())
    dg = env.device_grid()
    dg.set_motions(*motion_table(env, 3))
    import torch
    start = torch.ones(3, dtype=torch.int32, device=dg.device)
    dummy = torch.empty(16, dtype=torch.int32, device=dg.device)
    rc = dg.L.pmp_cost_field_u32_static(dg.handle, start.data_ptr(), dummy.data_ptr(), None, None, dg.stream())
    assert rc == -1 and b"static array holds" in dg.L.pmp_last_error()
    e26 = pmp.Grid3D(40, 24, 16, walls=False)                  # 26 motions: not this kernel's graph
    e26.update(set())
    d26 = e26.device_grid()
    d26.set_motions(*motion_table(e26, 3))
    rc = d26.L.pmp_cost_field_u32_static(d26.handle, start.data_ptr(), dummy.data_ptr(), None, None, d26.stream())
    assert rc == -1 and b"six unit moves" in d26.L.pmp_last_error()
