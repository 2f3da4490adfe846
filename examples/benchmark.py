"""Drop-in for the reference's examples/benchmark.py (benchmark.py:37-89): the same loop over
scenarios x algorithms x seeds, the same CSV schema

    Scenario, Algorithm, Runtime (s), Distance, Visited Nodes, Start, Goal, Seed

(benchmark.py:39), with the planners of python_motion_planning_b200.

The reference builds its 21 x 15 x 11 scenario maps with examples/scenarios.py, seeds `random`
with the iteration index and draws start / goal (benchmark.py:46-75).  Those 500 instances
(5 scenarios x 100 seeds: occupancy after carve_safety_bubble, start, goal) are committed in
tests/golden/csv_replay.npz, written from the unmodified reference by tools/make_golden.py, so this
script replays exactly the instances behind the reference's 3d_pathfinding_results.csv and its
output can be compared with that file row by row: with --order reference (default) Distance and
Visited Nodes of all seven algorithms' rows (astar, dstar, dijkstra, gbfs, lazy_theta_star, theta_star,
lpastar) are equal
(tests/test_examples.py); with --order wavefront Distance of astar / dijkstra.
With --scenarios-module the maps are generated live instead, from a module that offers the
reference's interface (`scenarios` dict of env -> obstacle set, `carve_safety_bubble`).

    python examples/benchmark.py [--out 3d_pathfinding_results.csv] [--repeats 10] [--algorithms astar,dijkstra,...]
"""
import argparse
import csv
import importlib
import os
import random
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

import python_motion_planning_b200 as pmp  # noqa: E402

algorithms = {                     # examples/algorithms.py, the entries this package implements
    "astar": pmp.AStar3D,
    "dstar": pmp.DStar3D,
    "dijkstra": pmp.Dijkstra3D,
    "gbfs": pmp.GBFS3D,
    "lazy_theta_star": pmp.LazyThetaStar3D,
    "theta_star": pmp.ThetaStar3D,
    "lpastar": pmp.LPAStar3D,
    "jps": pmp.JPS3D,
}


def replay_instances():
    """(scenario, seed, occupancy [x,y,z], start, goal) of the reference benchmark, from the fixture"""
    z = np.load(os.path.join(ROOT, "tests", "golden", "csv_replay.npz"), allow_pickle=False)
    dims = tuple(int(v) for v in z["dims"])
    names = [str(s) for s in z["scen_names"]]
    n = int(np.prod(dims))
    for k in range(len(z["seeds"])):
        flat = np.unpackbits(z["occ_bits"][k], bitorder="little")[:n].astype(bool)
        occ = flat.reshape(dims[::-1]).T.copy()
        yield (names[int(z["scen_idx"][k])], int(z["seeds"][k]), occ,
               tuple(int(v) for v in z["starts"][k]), tuple(int(v) for v in z["goals"][k]))


def live_instances(module, iterations, width=21, height=15, depth=11):
    """benchmark.py:41-77 with a user-supplied scenarios module"""
    mod = importlib.import_module(module)
    for scenario, scenario_func in mod.scenarios.items():
        for i in range(iterations):
            env = pmp.Grid3D(width, height, depth)
            XR, YR, ZR = env.x_range, env.y_range, env.z_range
            random.seed(i)
            draw = lambda: (random.randint(1, XR - 2), random.randint(1, YR - 2), random.randint(1, ZR - 2))  # noqa: E731
            start = draw()
            goal = draw()
            while goal == start:
                goal = draw()
            obstacles = scenario_func(env)
            mod.carve_safety_bubble(obstacles, start, radius=2)
            mod.carve_safety_bubble(obstacles, goal, radius=2)
            env.update(obstacles)
            yield scenario, i, env.occupancy(), start, goal


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--out", default="3d_pathfinding_results.csv")
    ap.add_argument("--repeats", type=int, default=10, help="rows per (scenario, algorithm, seed); the reference writes 10")
    ap.add_argument("--iterations", type=int, default=100)
    ap.add_argument("--algorithms", default="astar,dstar,dijkstra,gbfs,lazy_theta_star,theta_star,lpastar")
    ap.add_argument("--scenarios-module", default=None)
    ap.add_argument("--order", choices=("reference", "wavefront"), default="reference",
                    help="reference: plan() in the reference's pop order (pmp_bestfirst), so Distance AND Visited "
                         "Nodes equal the reference's rows; wavefront: the tiled relaxation kernels (equal Distance)")
    args = ap.parse_args()
    names = [a for a in args.algorithms.split(",") if a]
    inst = list(live_instances(args.scenarios_module, args.iterations) if args.scenarios_module
                else (t for t in replay_instances() if t[1] < args.iterations))
    scen_order = list(dict.fromkeys(t[0] for t in inst))
    with open(args.out, mode="w", newline="") as f:
        writer = csv.writer(f)
        writer.writerow(["Scenario", "Algorithm", "Runtime (s)", "Distance", "Visited Nodes", "Start", "Goal", "Seed"])
        for scenario in scen_order:
            for name in names:
                cls = algorithms[name]
                for sc, seed, occ, start, goal in inst:
                    if sc != scenario:
                        continue
                    for _ in range(args.repeats):
                        env = pmp.Grid3D(*occ.shape, walls=False)
                        env.update(set(map(tuple, np.argwhere(occ).tolist())))
                        planner = cls(start=start, goal=goal, env=env)
                        if args.order == "reference":
                            planner.reference_order = True         # astar / dijkstra / gbfs / lpastar / dstar
                        elif hasattr(planner, "mode"):             # theta_star / lazy_theta_star; jps
                            planner.mode = "wavefront" if name == "jps" else "shortcut"
                        t0 = time.time()
                        cost, path, expand = planner.plan()
                        t1 = time.time()
                        writer.writerow([scenario, name, t1 - t0, cost, len(expand), start, goal, seed])
    print(f"Results saved to {args.out}")


if __name__ == "__main__":
    main()
