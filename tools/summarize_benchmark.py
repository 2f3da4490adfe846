"""Runs examples/benchmark.py and prints one JSON object: median plan() time and mean Distance /
Visited Nodes per algorithm, next to the medians of the reference's own 3d_pathfinding_results.csv
(copied from profiles/r01_examples_benchmark.json, which was written where that file is readable).

    python tools/summarize_benchmark.py [--order reference|wavefront] [--iterations 20] [--repeats 2]
"""
import argparse
import csv
import json
import os
import statistics
import subprocess
import sys
import tempfile

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))

ap = argparse.ArgumentParser()
ap.add_argument("--order", default="reference")
ap.add_argument("--iterations", type=int, default=20)
ap.add_argument("--repeats", type=int, default=2)
args = ap.parse_args()
out = os.path.join(tempfile.mkdtemp(), "rows.csv")
subprocess.run([sys.executable, os.path.join(ROOT, "examples", "benchmark.py"), "--out", out, "--order", args.order,
                "--iterations", str(args.iterations), "--repeats", str(args.repeats)], check=True,
               stdout=subprocess.DEVNULL)
rows = list(csv.DictReader(open(out)))
algos = list(dict.fromkeys(r["Algorithm"] for r in rows))
prev = json.load(open(os.path.join(ROOT, "profiles", "r01_examples_benchmark.json")))
res = {"order": args.order, "rows": len(rows), "iterations": args.iterations, "repeats": args.repeats,
       "ours_median_ms": {}, "ours_mean_distance": {}, "ours_mean_visited": {},
       "reference_csv_median_ms": prev["reference_csv_median_ms"]}
for a in algos:
    rr = [r for r in rows if r["Algorithm"] == a]
    res["ours_median_ms"][a] = round(1e3 * statistics.median(float(r["Runtime (s)"]) for r in rr), 3)
    res["ours_mean_distance"][a] = round(statistics.fmean(float(r["Distance"]) for r in rr), 4)
    res["ours_mean_visited"][a] = round(statistics.fmean(int(r["Visited Nodes"]) for r in rr), 2)
print(json.dumps(res))
