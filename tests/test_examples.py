"""examples/benchmark.py: same CSV schema as the reference's
benchmark writer (benchmark.py:39), and the rows it writes for the replayed instances agree with
the reference's own 3d_pathfinding_results.csv (values committed in tests/golden/csv_replay.npz)."""
import csv
import os
import subprocess
import sys

import numpy as np
import pytest

from tests import golden_io as gio

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
HEADER = ["Scenario", "Algorithm", "Runtime (s)", "Distance", "Visited Nodes", "Start", "Goal", "Seed"]


@pytest.mark.gpu
def test_benchmark_writer_schema_and_known_answers(tmp_path):
    out = tmp_path / "results.csv"
    subprocess.run([sys.executable, os.path.join(ROOT, "examples", "benchmark.py"), "--out", str(out),
                    "--iterations", "4", "--repeats", "2"], check=True, cwd=str(tmp_path), timeout=600)
    rows = list(csv.reader(open(out)))
    assert rows[0] == HEADER
    body = [dict(zip(HEADER, r)) for r in rows[1:]]
    assert len(body) == 5 * 7 * 4 * 2
    z = gio.load("csv_replay.npz")
    az = gio.load("csv_anyangle.npz")
    iz = gio.load("csv_incr.npz")
    names = [str(s) for s in z["scen_names"]]
    index = {(names[int(z["scen_idx"][k])], int(z["seeds"][k])): k for k in range(len(z["seeds"]))}
    for r in body:
        k = index[(r["Scenario"], int(r["Seed"]))]
        assert eval(r["Start"]) == tuple(int(v) for v in z["starts"][k])
        assert eval(r["Goal"]) == tuple(int(v) for v in z["goals"][k])
        d, a = float(r["Distance"]), r["Algorithm"]
        assert float(r["Runtime (s)"]) >= 0.0 and int(r["Visited Nodes"]) >= 1
        if a in ("astar", "dijkstra", "gbfs"):                     # reference order: the CSV's own columns
            assert abs(d - float(z[f"{a}_csv_dist"][k])) <= 4e-16 * max(d, 1.0)
            assert int(r["Visited Nodes"]) == int(z[f"{a}_csv_vis"][k])
        if a in ("lpastar", "dstar"):                              # incl. greedy detours and D*'s inf distances
            want = float(iz[f"{a}_csv_dist"][k])
            assert (d == want) or abs(d - want) <= 4e-16 * max(d, 1.0)
            assert int(r["Visited Nodes"]) == int(iz[f"{a}_csv_vis"][k])
        if a in ("theta_star", "lazy_theta_star"):
            assert abs(d - float(az[f"{a}_csv_dist"][k])) <= 4e-16 * max(d, 1.0)
            assert int(r["Visited Nodes"]) == int(az[f"{a}_csv_vis"][k])


@pytest.mark.gpu
@pytest.mark.parametrize("flag", [[], ["--reference-order"]])
def test_global_examples_script(flag):
    """examples/global_examples.py (the reference's script with the package swapped): every 2D planner name of
    SearchFactory plans and 'animates' headlessly; README costs as measured on the reference"""
    out = subprocess.run([sys.executable, os.path.join(ROOT, "examples", "global_examples.py")] + flag, check=True,
                         capture_output=True, text=True, timeout=300).stdout
    rows = {line.split()[0]: line for line in out.strip().splitlines()}
    assert set(rows) == {"a_star", "dijkstra", "gbfs", "theta_star", "lazy_theta_star", "jps", "d_star", "lpa_star",
                         "d_star_lite"}
    assert "cost 54.041631" in rows["a_star"] and "cost 54.041631" in rows["dijkstra"]
    assert "cost 73.254834" in rows["gbfs"]
    if flag:                                   # reference order: the reference's own expansion counts
        assert rows["a_star"].rstrip().endswith("visited 579") and rows["dijkstra"].rstrip().endswith("visited 1219")
        assert rows["gbfs"].rstrip().endswith("visited 197")
        assert "cost 51.707233" in rows["theta_star"] and "cost 51.978792" in rows["lazy_theta_star"]
        assert "cost 53.455844" in rows["jps"]
