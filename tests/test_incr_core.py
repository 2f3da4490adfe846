"""CPU check of csrc/pmp_incr_core.h (LPA* / D* / D* Lite in the reference's processing order; the sequential
code the pmp_lpa_star / pmp_d_star kernels run on one thread), compiled with g++ by
tests/bestfirst_host.cpp.  Compared EXACTLY with runs of the unmodified reference committed in
tests/golden/incr.npz (tools/make_golden_incr.py): after plan() and after every occupancy change the
cost, the path and the EXPAND list; and with the lpastar / dstar rows of the reference's own CSV.
"""
import ctypes

import numpy as np
import pytest

from oracle import oracle
from tests import golden_io as gio
from tests.test_bestfirst_core import host  # noqa: F401  (fixture: builds and loads the harness)


class Session:
    def __init__(self, L, kind, occ, start, goal, heuristic=0, motions=None):
        self.L = L
        self.ndim = occ.ndim
        ndim, dims, occ_c, motions, _ = oracle._prep(occ, motions, None)
        self.dims, self.occ_c, self.motions = dims, occ_c, motions          # kept alive: the harness references them
        self.shape = occ.shape
        vp = ctypes.c_void_p
        L.incr_host_create.restype = vp
        L.incr_host_create.argtypes = [ctypes.c_int, ctypes.c_int, vp, vp, ctypes.c_int, vp, ctypes.c_int, vp, vp]
        L.incr_host_step.restype = ctypes.c_int
        L.incr_host_step.argtypes = [vp, ctypes.c_int, vp, vp, vp, ctypes.c_int64, vp, vp, ctypes.c_int64, vp]
        L.incr_host_destroy.argtypes = [vp]
        p = oracle._p
        self.h = L.incr_host_create(kind, ndim, p(dims), p(occ_c), len(motions), p(motions), heuristic,
                                    p(oracle._c3(start)), p(oracle._c3(goal)))

    def set_cell(self, c, blocked):
        idx = tuple(c)[::-1]            # occ_c is z-major
        self.occ_c[idx] = 1 if blocked else 0

    def step(self, op, change=None):
        cells = int(self.occ_c.size)
        cap = 200000
        cost, plen, nexp = ctypes.c_double(0), ctypes.c_int64(0), ctypes.c_int64(0)
        path = np.zeros((cap, 3), dtype=np.int32)
        ecap = 64 * cells
        e_cell = np.zeros(ecap, dtype=np.int64)
        ch = np.zeros(4, dtype=np.int32)
        if change is not None:
            ch[:self.ndim] = change[0]
            ch[3] = 1 if change[1] else 0
        p = oracle._p
        rc = self.L.incr_host_step(self.h, op, p(ch), ctypes.byref(cost), p(path), cap, ctypes.byref(plen),
                                   p(e_cell), ecap, ctypes.byref(nexp))
        assert rc == 0, rc
        pts = [tuple(int(v) for v in row[:self.ndim]) for row in path[:plen.value]]
        return cost.value, pts, e_cell[:nexp.value]

    def close(self):
        self.L.incr_host_destroy(self.h)


def load_cases():
    z = gio.load("incr.npz")
    for k in range(int(z["n_cases"])):
        shape = tuple(int(v) for v in z[f"c{k}_shape"])
        yield dict(kind=int(z[f"c{k}_kind"]), shape=shape, occ=gio.unpack_occ(z[f"c{k}_bits"], shape),
                   start=tuple(int(v) for v in z[f"c{k}_start"]), goal=tuple(int(v) for v in z[f"c{k}_goal"]),
                   changes=z[f"c{k}_changes"], blocked=z[f"c{k}_blocked"], cost=z[f"c{k}_cost"],
                   path=gio.Ragged(z, f"c{k}_path"), expand=gio.Ragged(z, f"c{k}_expand"))


def check_step(c, i, got, kind):
    cost, pts, exp = got
    ndim = len(c["shape"])
    want_path = [tuple(int(v) for v in p[:ndim]) for p in c["path"][i]]
    if kind == 0:
        pts = pts[::-1]                 # the core writes goal first; lpa_star.py:229 returns start -> goal
    tag = (kind, c["shape"], i)
    assert pts == want_path, tag
    assert cost == float(c["cost"][i]) or (np.isinf(cost) and np.isinf(c["cost"][i])), tag
    assert np.array_equal(exp, c["expand"][i][:, 0] if len(c["expand"][i]) else np.zeros(0, dtype=np.int64)), tag


def test_incremental_cores_equal_reference(host):  # noqa: F811
    n = 0
    for c in load_cases():
        s = Session(host, c["kind"], c["occ"], c["start"], c["goal"])
        check_step(c, 0, s.step(0), c["kind"])
        for i, (cell, b) in enumerate(zip(c["changes"], c["blocked"])):
            cell = tuple(int(v) for v in cell)
            if c["kind"] != 2:          # D* Lite toggles the cell itself, one step into its walk (d_star_lite.py:82-90)
                s.set_cell(cell, bool(b))
            check_step(c, i + 1, s.step(1, (cell, not bool(b))), c["kind"])
            assert bool(s.occ_c[cell[::-1]]) == bool(b)
            n += 1
        s.close()
    assert n >= 80


@pytest.mark.parametrize("name,kind", [("lpastar", 0), ("dstar", 1)])
def test_incremental_cores_csv_known_answers(host, name, kind):  # noqa: F811
    """Distance and Visited Nodes of the lpastar / dstar rows of the reference's CSV (every 5th instance)"""
    z = gio.load("csv_replay.npz")
    a = gio.load("csv_incr.npz")
    dims = tuple(int(v) for v in z["dims"])
    for k in range(0, len(z["starts"]), 5):
        occ = gio.unpack_occ(z["occ_bits"][k], dims)
        s = Session(host, kind, occ, z["starts"][k], z["goals"][k])
        cost, pts, exp = s.step(0)
        s.close()
        assert len(exp) == int(a[f"{name}_csv_vis"][k]), (name, k)
        want = float(a[f"{name}_csv_dist"][k])
        assert (np.isinf(cost) and np.isinf(want)) or cost == pytest.approx(want, rel=4e-16, abs=0), (name, k)
        assert len(pts) == int(a[f"{name}_ref_plen"][k]), (name, k)
