"""The graph-search part of the reference's examples/global_examples.py (lines 11-49) with the package
swapped: every planner name SearchFactory knows here, on the README map, headless (run() records
its animation arguments instead of opening a window).

    python examples/global_examples.py [--reference-order]
"""
import os
import sys

sys.path.append(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from python_motion_planning_b200.utils import Grid, SearchFactory  # noqa: E402
import python_motion_planning_b200 as pmp  # noqa: E402

if __name__ == '__main__':
    if "--reference-order" in sys.argv:
        pmp.set_reference_order(True)
    search_factory = SearchFactory()
    start = (5, 5)
    goal = (45, 25)
    env = Grid(51, 31)
    obstacles = env.obstacles
    for i in range(10, 21):
        obstacles.add((i, 15))
    for i in range(15):
        obstacles.add((20, i))
    for i in range(15, 30):
        obstacles.add((30, i))
    for i in range(16):
        obstacles.add((40, i))
    env.update(obstacles)
    for name in ("a_star", "dijkstra", "gbfs", "theta_star", "lazy_theta_star", "jps", "d_star", "lpa_star",
                 "d_star_lite"):
        planner = search_factory(name, start=start, goal=goal, env=env)
        cost, path, expand = planner.plan()
        planner.run()
        visited = len(expand) if expand is not None else len(getattr(planner, "EXPAND", []))
        print(f"{name:16s} {str(planner):34s} cost {cost:.6f}  path {len(path):3d} points  visited {visited}")
