"""Golden vectors for the JPS row, written from the UNMODIFIED
reference in the build container (needs /root/reference):

  * JPS (2D, jps.py) cost and jump-point path for seeded queries on the 2D fixture grids;
  * JPS3D and AStar3D costs on seeded random 3D maps (JPS3D prunes too much: the comparison is
    printed and stored).

    python tools/make_golden_jps.py        ->  tests/golden/jps.npz
"""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tools"))
import ref_harness  # noqa: E402
from tests import golden_io as gio  # noqa: E402

OUT = os.path.join(ROOT, "tests", "golden")


def pack(occ):
    return np.packbits(np.ascontiguousarray(occ.T).reshape(-1), bitorder="little")


def main():
    pmp = ref_harness.load_reference()
    ref_harness.fast_grid_map(pmp)
    sys.setrecursionlimit(50000)
    rng = np.random.default_rng(17)
    out = {}
    c2 = gio.Cases("cases2d.npz", 2)
    grids = list(c2.grids[:3])
    q2, paths = [], []
    for gi, occ in enumerate(grids):
        env = pmp.Grid(*occ.shape)
        env.update(set(map(tuple, np.argwhere(occ).tolist())))
        free = np.argwhere(~occ)
        done = 0
        while done < 14:
            i, j = rng.integers(0, len(free), size=2)
            s, g = tuple(int(v) for v in free[i]), tuple(int(v) for v in free[j])
            if s == g:
                continue
            try:
                cost, path, _ = pmp.JPS(start=s, goal=g, env=env).plan()
            except RecursionError:
                continue
            q2.append((gi, s[0], s[1], g[0], g[1], float(cost) if path else float("inf")))
            paths.append(np.array(path, dtype=np.int32).reshape(-1, 2))
            done += 1
        out[f"g{gi}_shape"] = np.array(occ.shape, dtype=np.int32)
        out[f"g{gi}_bits"] = pack(occ)
    out["n_grids"] = np.array(len(grids))
    out["jps2d_queries"] = np.array(q2, dtype=np.float64)
    out["jps2d_path_off"] = np.cumsum([0] + [len(p) for p in paths]).astype(np.int64)
    out["jps2d_path_dat"] = np.concatenate(paths) if paths else np.zeros((0, 2), dtype=np.int32)
    # 3D
    q3 = []
    dims = (16, 13, 10)
    occs = []
    for t in range(3):
        env = pmp.Grid3D(*dims)
        obs = env.obstacles
        for x in range(1, dims[0] - 1):
            for y in range(1, dims[1] - 1):
                for z in range(1, dims[2] - 1):
                    if rng.random() < 0.14:
                        obs.add((x, y, z))
        env.update(obs)
        occ = np.zeros(dims, dtype=bool)
        for p in obs:
            if all(0 <= c < d for c, d in zip(p, dims)):
                occ[p] = True
        occs.append(occ)
        free = np.argwhere(~occ)
        for _ in range(8):
            i, j = rng.integers(0, len(free), size=2)
            s, g = tuple(int(v) for v in free[i]), tuple(int(v) for v in free[j])
            if s == g:
                continue
            cj, pj, _ = pmp.JPS3D(start=s, goal=g, env=env).plan()
            ca, pa, _ = pmp.AStar3D(start=s, goal=g, env=env).plan()
            q3.append((t,) + s + g + (float(cj) if pj else float("inf"), float(ca) if pa else float("inf")))
        out[f"h{t}_shape"] = np.array(dims, dtype=np.int32)
        out[f"h{t}_bits"] = pack(occ)
    out["n_grids3d"] = np.array(len(occs))
    out["jps3d_queries"] = np.array(q3, dtype=np.float64)
    a = out["jps3d_queries"]
    fin = np.isfinite(a[:, 7]) & np.isfinite(a[:, 8])
    print("2D queries", len(q2), "3D queries", len(q3), "JPS3D longer than AStar3D in",
          int((a[fin, 7] > a[fin, 8] * (1 + 1e-9)).sum()), "of", int(fin.sum()))
    np.savez_compressed(os.path.join(OUT, "jps.npz"), **out)
    print("wrote", os.path.join(OUT, "jps.npz"))


if __name__ == "__main__":
    main()
