"""A handful of reference-order plans (smoke run of every reference-order entry point):
    python tools/reference_order_smoke.py
Every algorithm of pmp_bestfirst, pmp_lpa_star (LPA*, D* Lite) and pmp_d_star once, on small maps."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np

import python_motion_planning_b200 as pmp

rng = np.random.default_rng(5)


def env2d():
    env = pmp.Grid(31, 21)
    obs = env.obstacles
    for x in range(1, 30):
        for y in range(1, 20):
            if rng.random() < 0.15 and (x, y) not in ((2, 2), (28, 18)):
                obs.add((x, y))
    env.update(obs)
    return env


def env3d():
    env = pmp.Grid3D(14, 12, 10)
    obs = env.obstacles
    for x in range(1, 13):
        for y in range(1, 11):
            for z in range(1, 9):
                if rng.random() < 0.12 and (x, y, z) not in ((1, 1, 1), (12, 10, 8)):
                    obs.add((x, y, z))
    env.update(obs)
    return env


out = []
e2, e3 = env2d(), env3d()
for cls in (pmp.AStar, pmp.Dijkstra, pmp.GBFS):
    out.append((cls.__name__, pmp_cost := cls((2, 2), (28, 18), e2, reference_order=True).plan()[0]))
for cls in (pmp.ThetaStar, pmp.LazyThetaStar, pmp.JPS):
    out.append((cls.__name__, cls((2, 2), (28, 18), e2, mode="reference").plan()[0]))
for cls in (pmp.AStar3D, pmp.GBFS3D):
    out.append((cls.__name__, cls((1, 1, 1), (12, 10, 8), e3, reference_order=True).plan()[0]))
out.append(("Dijkstra3D", pmp.Dijkstra3D((1, 1, 1), (12, 10, 8), e3, reference_order=True).plan()[0]))
for cls in (pmp.ThetaStar3D, pmp.LazyThetaStar3D, pmp.JPS3D):
    out.append((cls.__name__, cls((1, 1, 1), (12, 10, 8), e3, mode="reference").plan()[0]))
for cls, s, g, env in ((pmp.LPAStar, (2, 2), (28, 18), env2d()), (pmp.LPAStar3D, (1, 1, 1), (12, 10, 8), env3d()),
                       (pmp.DStarLite, (2, 2), (28, 18), env2d())):
    p = cls(s, g, env, reference_order=True)
    cost, path, _ = p.plan()
    for i in (3, 6):
        if len(path) > i + 2:
            cost, path, _ = p.apply_change(tuple(path[i]))
    out.append((cls.__name__, cost))
for cls, s, g, env in ((pmp.DStar, (2, 2), (28, 18), env2d()), (pmp.DStar3D, (1, 1, 1), (12, 10, 8), env3d())):
    p = cls(s, g, env, reference_order=True)
    cost, path, _ = p.plan()
    if len(path) > 6:
        cost, path = p.apply_dynamic_obstacles([tuple(path[4])])
    out.append((cls.__name__, cost))
for name, c in out:
    print(name, c)
