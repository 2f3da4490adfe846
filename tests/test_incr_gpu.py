"""GPU parity of the reference-order incremental planners (pmp_lpa_star / pmp_d_star) through the
planner classes (LPAStar, LPAStar3D, DStar, DStar3D, DStarLite with reference_order=True): EXACT equality with
runs of the unmodified reference (tests/golden/incr.npz) after plan() and after every occupancy
change -- cost, path, EXPAND -- and with the Distance / Visited Nodes columns of the lpastar and dstar
rows of the reference's own 3d_pathfinding_results.csv (tests/golden/csv_incr.npz).
"""
import math

import numpy as np
import pytest

import python_motion_planning_b200 as pmp
from tests import golden_io as gio
from tests.test_gpu_parity import make_env
from tests.test_incr_core import load_cases

pytestmark = pytest.mark.gpu

CLS = {(0, 2): pmp.LPAStar, (0, 3): pmp.LPAStar3D, (1, 2): pmp.DStar, (1, 3): pmp.DStar3D, (2, 2): pmp.DStarLite}


def flat(coords, shape):
    c = np.asarray(coords, dtype=np.int64).reshape(-1, len(shape))
    if len(shape) == 2:
        return c[:, 1] * shape[0] + c[:, 0]
    return (c[:, 2] * shape[1] + c[:, 1]) * shape[0] + c[:, 0]


def check(c, i, planner, cost, path):
    ndim = len(c["shape"])
    tag = (c["kind"], c["shape"], i)
    want_path = [tuple(int(v) for v in p[:ndim]) for p in c["path"][i]]
    assert path == want_path, tag
    want = float(c["cost"][i])
    assert cost == want or (math.isinf(cost) and math.isinf(want)), tag
    want_exp = c["expand"][i][:, 0] if len(c["expand"][i]) else np.zeros(0, dtype=np.int64)
    got = flat([n.current for n in planner.EXPAND], c["shape"]) if len(planner.EXPAND) else np.zeros(0, dtype=np.int64)
    assert np.array_equal(got, want_exp), tag


def test_lpa_and_dstar_sequences_equal_reference():
    steps = 0
    for c in load_cases():
        ndim = len(c["shape"])
        env = make_env(c["occ"])
        planner = CLS[(c["kind"], ndim)](c["start"], c["goal"], env, reference_order=True)
        cost, path, expand = planner.plan()
        assert (expand is None) == (ndim == 2)                 # lpa_star.py:89, d_star.py:88
        check(c, 0, planner, cost, path)
        for i, (cell, b) in enumerate(zip(c["changes"], c["blocked"])):
            cell = tuple(int(v) for v in cell)
            if c["kind"] in (0, 2):
                cost, path, _ = planner.apply_change(cell)     # toggles, like the reference run
            elif ndim == 3:
                cost, path = planner.apply_dynamic_obstacles([cell])
            else:
                cost, path, _ = planner.apply_change(cell, True)
            assert (cell in planner.obstacles) == bool(b)
            check(c, i + 1, planner, cost, path)
            steps += 1
    assert steps >= 80


@pytest.mark.parametrize("name,cls", [("lpastar", pmp.LPAStar3D), ("dstar", pmp.DStar3D)])
def test_csv_known_answers_lpa_dstar(name, cls):
    z = gio.load("csv_replay.npz")
    a = gio.load("csv_incr.npz")
    dims = tuple(int(v) for v in z["dims"])
    n = 0
    for k in range(0, len(z["starts"]), 2):
        occ = gio.unpack_occ(z["occ_bits"][k], dims)
        s, g = tuple(int(v) for v in z["starts"][k]), tuple(int(v) for v in z["goals"][k])
        cost, path, expand = cls(s, g, make_env(occ), reference_order=True).plan()
        assert len(expand) == int(a[f"{name}_csv_vis"][k]), (name, k)
        want = float(a[f"{name}_csv_dist"][k])
        assert (math.isinf(cost) and math.isinf(want)) or cost == pytest.approx(want, rel=4e-16, abs=0), (name, k)
        assert len(path) == int(a[f"{name}_ref_plen"][k]), (name, k)
        n += 1
    assert n == 250
