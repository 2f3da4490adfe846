"""Readers for the committed fixtures under tests/golden/ (written by tools/make_golden.py
from the unmodified reference)."""
import os

import numpy as np

GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")


def unpack_occ(bits, shape):
    """packed little-endian bits in z-major (x fastest) order -> bool array indexed [x,y(,z)]"""
    n = int(np.prod(shape))
    flat = np.unpackbits(np.asarray(bits, dtype=np.uint8), bitorder="little")[:n].astype(bool)
    return flat.reshape(tuple(shape)[::-1]).T.copy()


class Ragged:
    def __init__(self, z, prefix):
        self.off = z[prefix + "_off"]
        self.dat = z[prefix + "_dat"]

    def __getitem__(self, i):
        return self.dat[self.off[i]:self.off[i + 1]]


def load(name):
    return np.load(os.path.join(GOLDEN, name), allow_pickle=False)


MOTION_SETS = {}


def motions_for(ndim, mset):
    """motion subset (offsets, costs) in the reference's order for a fixture's motion-set id"""
    from oracle import oracle
    base = oracle.MOTIONS_2D if ndim == 2 else oracle.MOTIONS_3D
    cost = oracle.motion_costs(base)
    if (ndim == 2 and mset == 8) or (ndim == 3 and mset == 26):
        keep = np.ones(len(base), dtype=bool)
    elif mset in (4, 6):
        keep = cost == 1.0
    elif mset == 18:
        keep = cost < 1.5
    else:
        raise ValueError(mset)
    return base[keep].copy(), cost[keep].copy()


class Cases:
    """cases2d.npz / cases3d.npz"""

    def __init__(self, fname, ndim):
        z = load(fname)
        self.ndim = ndim
        self.meta = z["meta"]
        self.cost = z["cost"]
        self.found = z["found"]
        self.grids = [unpack_occ(z[f"grid{i}_bits"], z[f"grid{i}_shape"]) for i in range(int(z["n_grids"]))]
        self.grid_names = [str(s) for s in z["grid_names"]]
        self.path = Ragged(z, "path")
        self.e_cell = Ragged(z, "ecell")
        self.e_g = Ragged(z, "eg")
        self.e_h = Ragged(z, "eh")
        self.e_parent = Ragged(z, "eparent")

    def __len__(self):
        return len(self.meta)

    def case(self, i):
        gi, algo, heur, mset = (int(v) for v in self.meta[i][:4])
        start = tuple(int(v) for v in self.meta[i][4:4 + self.ndim])
        goal = tuple(int(v) for v in self.meta[i][7:7 + self.ndim])
        return dict(gi=gi, occ=self.grids[gi], algo=algo, heur=heur, mset=mset, start=start, goal=goal,
                    cost=float(self.cost[i]), found=bool(self.found[i]),
                    path=[tuple(int(v) for v in p[:self.ndim]) for p in self.path[i]],
                    e_cell=self.e_cell[i][:, 0], e_g=self.e_g[i][:, 0], e_h=self.e_h[i][:, 0],
                    e_parent=self.e_parent[i][:, 0])
