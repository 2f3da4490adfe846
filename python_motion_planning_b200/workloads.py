"""Synthetic inputs of the BASELINE.json configurations; seeded numpy
generators, no reference code involved."""
import numpy as np


def seal(occ):
    """shell walls: every face of the box occupied (examples/scenarios.py:4-20 shell_walls)"""
    for ax in range(occ.ndim):
        sl = [slice(None)] * occ.ndim
        sl[ax] = 0
        occ[tuple(sl)] = True
        sl[ax] = -1
        occ[tuple(sl)] = True
    return occ


def readme_grid():
    """config 1: README.md:62-76 -- Grid(51, 31) with four walls, (5,5) -> (45,25)"""
    occ = seal(np.zeros((51, 31), dtype=bool))
    occ[10:21, 15] = True
    occ[20, 0:15] = True
    occ[30, 15:30] = True
    occ[40, 0:16] = True
    return occ, (5, 5), (45, 25)


def random2d(n=1024, density=0.30, nq=4096, seed=0):
    """config 3: n x n, `density` random obstacles, border occupied; nq (start, goal) pairs drawn
    uniformly from free cells (unreachable pairs are kept)."""
    occ = np.random.default_rng(seed).random((n, n)) < density
    seal(occ)
    free = np.argwhere(~occ)
    idx = np.random.default_rng(seed + 1).integers(0, len(free), size=(nq, 2))
    return occ, free[idx[:, 0]].astype(np.int32), free[idx[:, 1]].astype(np.int32)


def boxes3d(n=512, n_boxes=None, edge=(4, 32), seed=0, bubble=2):
    """config 4: n^3 voxels, shell walls + random axis-aligned boxes (4096 at n = 512, scaled by
    volume otherwise), edges uniform in [4, 32]; radius-2 bubbles carved at (1,1,1) and
    (n-2,n-2,n-2) as examples/scenarios.py:182-207 does."""
    if n_boxes is None:
        n_boxes = max(8, int(round(4096 * (n / 512.0) ** 3)))
    rng = np.random.default_rng(seed)
    occ = np.zeros((n, n, n), dtype=bool)
    c = rng.integers(1, n - 1, size=(n_boxes, 3))
    e = rng.integers(edge[0], edge[1] + 1, size=(n_boxes, 3))
    for (x, y, z), (a, b, d) in zip(c, e):
        occ[x:x + a, y:y + b, z:z + d] = True
    r = bubble
    occ[1:2 + r, 1:2 + r, 1:2 + r] = False
    occ[n - 2 - r:n - 1, n - 2 - r:n - 1, n - 2 - r:n - 1] = False
    return seal(occ)


def replicated_boxes3d(base=128, reps=12, n_boxes=512, edge=(2, 12), seed=0):
    """config 5: obstacle boxes generated at base^3 and replicated `reps` times per axis
    (1536^3 at the defaults), shell walls around the whole grid.  Returns (base block, full
    shape); materialise a z-range with `replicated_slab`."""
    rng = np.random.default_rng(seed)
    blk = np.zeros((base, base, base), dtype=bool)
    c = rng.integers(0, base, size=(n_boxes, 3))
    e = rng.integers(edge[0], edge[1] + 1, size=(n_boxes, 3))
    for (x, y, z), (a, b, d) in zip(c, e):
        blk[x:x + a, y:y + b, z:z + d] = True
    blk[0:4, 0:4, 0:4] = False
    return blk, (base * reps,) * 3


def replicated_slab(blk, shape, z0, z1):
    """occupancy [x, y, z0:z1] of the replicated grid, with the global shell walls"""
    base = blk.shape[0]
    reps = shape[0] // base
    zs = np.arange(z0, z1)
    plane = np.tile(blk[:, :, zs % base], (reps, reps, 1))
    plane[0, :, :] = plane[-1, :, :] = True
    plane[:, 0, :] = plane[:, -1, :] = True
    plane[:, :, zs == 0] = True
    plane[:, :, zs == shape[2] - 1] = True
    return plane


def pack_rows(occ):
    """dense occupancy [x, y, z] -> bit rows [z][y][ceil(X/32)] of uint32 (Grid3D.update_packed)"""
    X = occ.shape[0]
    b = np.packbits(np.ascontiguousarray(occ.transpose(2, 1, 0)), axis=-1, bitorder="little")
    pad = (-b.shape[-1]) % 4
    if pad:
        b = np.concatenate([b, np.zeros(b.shape[:-1] + (pad,), dtype=np.uint8)], axis=-1)
    return np.ascontiguousarray(b).view(np.uint32).reshape(occ.shape[2], occ.shape[1], (X + 31) // 32)


def replicated_packed(blk, shape, z0, z1):
    """replicated_slab(blk, shape, z0, z1) as bit rows [z1 - z0][Y][X / 32], built from the packed
    block without materialising a byte per cell (base must be a multiple of 32)"""
    base = blk.shape[0]
    reps = shape[0] // base
    assert base % 32 == 0 and shape[0] == shape[1] == base * reps
    pb = pack_rows(blk)                                    # [base z][base y][base / 32]
    zs = np.arange(z0, z1)
    rows = np.tile(pb[zs % base], (1, reps, reps))         # [nz][Y][X / 32]
    rows[:, :, 0] |= np.uint32(1)                          # x = 0 wall
    rows[:, :, -1] |= np.uint32(0x80000000)                # x = X - 1 wall
    rows[:, 0, :] = np.uint32(0xFFFFFFFF)
    rows[:, -1, :] = np.uint32(0xFFFFFFFF)
    rows[zs == 0] = np.uint32(0xFFFFFFFF)
    rows[zs == shape[2] - 1] = np.uint32(0xFFFFFFFF)
    return rows


def door3d(dims=(26, 20, 16), door_size=2, start=(1, 1, 1), goal=None, bubble=1):
    """config 2: the default scene of examples/3d_example.py -- two rooms split by a wall at
    x = X//2 with a door_size^2 door at the centre (examples/scenarios.py:28-47), shell walls,
    radius-1 bubbles carved at start and goal (3d_example.py:84-86).  Returns (occ, dims)."""
    X, Y, Z = dims
    goal = goal or (X - 2, Y - 2, Z - 2)
    occ = seal(np.zeros(dims, dtype=bool))
    x0 = X // 2
    occ[x0, 1:Y - 1, 1:Z - 1] = True
    ym, zm = Y // 2, Z // 2
    lo, hi = -(door_size // 2), door_size - (door_size // 2)
    occ[x0, ym + lo:ym + hi, zm + lo:zm + hi] = False
    for c in (start, goal):
        sl = tuple(slice(max(1, c[i] - bubble), min(dims[i] - 1, c[i] + bubble + 1)) for i in range(3))
        occ[sl] = False
    return occ, dims
