"""Grid / Grid3D / Node / Node3D with the reference's Python surface, backed by a device bitmask.

Mirrors /root/reference/src/python_motion_planning/utils/environment/{env.py:14-80,
env3d.py:14-103, node.py:8-84, node3d.py:8-96}: same constructor arguments, attributes
(`x_range`, `y_range`, `z_range`, `eps`, `motions`, `obstacles`, `grid_map`) and `update()`.
Additions: `update()` also accepts a dense occupancy array / torch tensor indexed [x, y(, z)]
(the reference's set-of-tuples cannot represent 512^3 grids), `obstacles` is then materialised
lazily, and `obstacles_tree` (a scipy cKDTree the reference rebuilds on every update but never
queries on the graph-search path) is built on first access only.
"""
import math
import weakref
from math import sqrt

import itertools
import os

import numpy as np

from . import _lib


class Node(object):
    """Search record (node.py:8-84): `current`, `parent`, `g`, `h`; equality/hash by coordinate;
    `+` applies a motion; `<` orders by f = g + h, then h."""

    def __init__(self, current: tuple, parent: tuple = None, g: float = 0, h: float = 0) -> None:
        self.current = current
        self.parent = parent
        self.g = g
        self.h = h

    def __add__(self, node):
        assert isinstance(node, Node)
        return Node((self.x + node.x, self.y + node.y), self.parent, self.g + node.g, self.h)

    def __eq__(self, node) -> bool:
        if not isinstance(node, Node):
            return False
        return self.current == node.current

    def __ne__(self, node) -> bool:
        return not self.__eq__(node)

    def __lt__(self, node) -> bool:
        assert isinstance(node, Node)
        return self.g + self.h < node.g + node.h or \
            (self.g + self.h == node.g + node.h and self.h < node.h)

    def __hash__(self) -> int:
        return hash(self.current)

    def __str__(self) -> str:
        return "Node({}, {}, {}, {})".format(self.current, self.parent, self.g, self.h)

    __repr__ = __str__

    @property
    def x(self):
        return self.current[0]

    @property
    def y(self):
        return self.current[1]

    @property
    def px(self):
        return self.parent[0] if self.parent else None

    @property
    def py(self):
        return self.parent[1] if self.parent else None


class Node3D(object):
    """3D search record (node3d.py:8-96)."""

    def __init__(self, current: tuple, parent: tuple = None, g: float = 0, h: float = 0) -> None:
        self.current = current
        self.parent = parent
        self.g = g
        self.h = h

    def __add__(self, node):
        assert isinstance(node, Node3D)
        return Node3D((self.x + node.x, self.y + node.y, self.z + node.z), self.parent,
                      self.g + node.g, self.h)

    def __eq__(self, node) -> bool:
        if not isinstance(node, Node3D):
            return False
        return self.current == node.current

    def __ne__(self, node) -> bool:
        return not self.__eq__(node)

    def __lt__(self, node) -> bool:
        assert isinstance(node, Node3D)
        return self.g + self.h < node.g + node.h or \
            (self.g + self.h == node.g + node.h and self.h < node.h)

    def __hash__(self) -> int:
        return hash(self.current)

    def __str__(self) -> str:
        return "Node({}, {}, {}, {})".format(self.current, self.parent, self.g, self.h)

    __repr__ = __str__

    @property
    def x(self):
        return self.current[0]

    @property
    def y(self):
        return self.current[1]

    @property
    def z(self):
        return self.current[2]

    @property
    def px(self):
        return self.parent[0] if self.parent else None

    @property
    def py(self):
        return self.parent[1] if self.parent else None

    @property
    def pz(self):
        return self.parent[2] if self.parent else None


def _destroy_handle(handle):
    try:
        _lib.lib().pmp_grid_destroy(handle)
    except Exception:
        pass


class DeviceGrid:
    """Owner of one pmp_grid handle (occupancy bitmask + motion table on the current device)."""

    def __init__(self, ndim, dims):
        import ctypes
        self.torch = _lib.require_cuda()
        self.L = _lib.lib()
        self.ndim = ndim
        self.dims = tuple(int(d) for d in dims)
        d3 = (ctypes.c_int32 * 3)(*(list(self.dims) + [1] * (3 - ndim)))
        h = ctypes.c_void_p()
        _lib.check(self.L.pmp_grid_create(ndim, d3, ctypes.byref(h)), "pmp_grid_create")
        self.handle = h
        self._fin = weakref.finalize(self, _destroy_handle, h)
        lay = (ctypes.c_int64 * 8)()
        _lib.check(self.L.pmp_grid_layout(h, lay), "pmp_grid_layout")
        self.padded = (int(lay[0]), int(lay[1]), int(lay[2]))
        self.tile = (int(lay[3]), int(lay[4]), int(lay[5]))
        self.ntiles = int(lay[6])
        self.words_per_row = int(lay[7])
        self.pcells = self.padded[0] * self.padded[1] * self.padded[2]
        self.device = self.torch.device("cuda", self.torch.cuda.current_device())
        self._motion_key = None
        self.h2d_bytes = 0   # bytes uploaded by the last set_* call (bench accounting)

    def stream(self):
        return self.torch.cuda.current_stream().cuda_stream

    def set_cells(self, coords):
        """coords: (n, ndim) integer array of obstacle cells (host)."""
        t = self.torch
        a = np.zeros((len(coords), 3), dtype=np.int32)
        if len(coords):
            a[:, :self.ndim] = np.asarray(coords, dtype=np.int32).reshape(len(coords), self.ndim)
        host = t.from_numpy(a).pin_memory()
        dev = host.to(self.device, non_blocking=True)
        self.h2d_bytes = a.nbytes
        _lib.check(self.L.pmp_grid_set_cells(self.handle, dev.data_ptr(), len(a), self.stream()),
                   "pmp_grid_set_cells")
        self._keep = dev

    def set_dense(self, occ):
        """occ: array / tensor indexed [x, y(, z)], non-zero = obstacle.  One-byte element types
        are uploaded as they are (pinned host memory -> device) and packed by a transposing
        kernel; nothing is converted or transposed on the host."""
        t = self.torch
        if isinstance(occ, t.Tensor):
            if tuple(occ.shape) != self.dims:
                raise ValueError(f"occupancy shape {tuple(occ.shape)} != grid {self.dims}")
            o = occ if occ.element_size() == 1 else (occ != 0)
            o = o.contiguous()
            self.h2d_bytes = 0 if o.is_cuda else o.numel() * o.element_size()
            dev = o.to(self.device, non_blocking=True)
            dev = dev.view(t.uint8)
        else:
            a = np.asarray(occ)
            if a.shape != self.dims:
                raise ValueError(f"occupancy shape {a.shape} != grid {self.dims}")
            if a.dtype.itemsize != 1:
                a = a != 0
            a = np.ascontiguousarray(a).view(np.uint8)
            host = t.from_numpy(a)
            dev = host.pin_memory().to(self.device, non_blocking=True)
            self.h2d_bytes = a.nbytes
        _lib.check(self.L.pmp_grid_set_dense_xyz(self.handle, dev.data_ptr(), self.stream()),
                   "pmp_grid_set_dense_xyz")
        self._keep = dev

    def set_packed(self, rows):
        """rows: uint32 array / tensor [Z][Y][ceil(X/32)] (2D: [Y][ceil(X/32)]), bit i of word w <->
        x = 32 w + i, 1 = obstacle -- occupancy the caller already holds as bits: 1/8 of the upload
        of the dense form.  Host arrays go pinned host memory -> device."""
        t = self.torch
        X, Y = self.dims[0], self.dims[1]
        Z = self.dims[2] if self.ndim == 3 else 1
        want = Z * Y * ((X + 31) // 32)
        if isinstance(rows, t.Tensor):
            if rows.numel() != want or rows.element_size() != 4:
                raise ValueError(f"packed occupancy: expected {want} 32-bit words")
            if (not rows.is_cuda and rows.is_pinned() and rows.is_contiguous() and self.ndim == 3 and X % 32 == 0
                    and self.padded[1] == Y and self.padded[2] == Z and not os.environ.get("PMP_B200_NO_STREAMED_UPLOAD")):
                # pinned host rows in the grid's own layout: streamed upload -- the copy engine fills the bitmask chunk
                # by chunk while a unit-weight field search launched next already runs behind it (k_bfs re-queues a
                # tile until the chunks it needs are there); every other entry point waits for the whole upload first
                if getattr(self, "_copy_stream", None) is None:
                    self._copy_stream = t.cuda.Stream(device=self.device)
                _lib.check(self.L.pmp_grid_set_packed_host_streamed(self.handle, rows.data_ptr(), int(os.environ.get("PMP_B200_STREAM_PLANES", "32")), self.stream(),
                                                                    self._copy_stream.cuda_stream),
                           "pmp_grid_set_packed_host_streamed")
                self.h2d_bytes = rows.numel() * 4
                self._keep = rows
                return
            self.h2d_bytes = 0 if rows.is_cuda else rows.numel() * 4
            dev = rows.contiguous().to(self.device, non_blocking=True)
        else:
            a = np.ascontiguousarray(rows)
            if a.size != want or a.dtype.itemsize != 4:
                raise ValueError(f"packed occupancy: expected {want} 32-bit words")
            dev = t.from_numpy(a.view(np.int32)).pin_memory().to(self.device, non_blocking=True)
            self.h2d_bytes = a.nbytes
        _lib.check(self.L.pmp_grid_set_packed(self.handle, dev.data_ptr(), self.stream()), "pmp_grid_set_packed")
        self._keep = dev

    def set_motions(self, offsets, costs, cut=False):
        """motion table + edge rule (cut = True: diagonal moves may cut corners, the JPS graph)"""
        key = (tuple(map(tuple, offsets)), tuple(costs))
        if key != self._motion_key:
            off = np.zeros((len(offsets), 3), dtype=np.int32)
            for i, o in enumerate(offsets):
                off[i, :len(o)] = o
            c = np.ascontiguousarray(costs, dtype=np.float64)
            _lib.check(self.L.pmp_grid_set_motions(self.handle, len(off), off.ctypes.data, c.ctypes.data),
                       "pmp_grid_set_motions")
            self._motion_key = key
        if bool(cut) != getattr(self, "_cut", False):
            _lib.check(self.L.pmp_grid_set_corner_cutting(self.handle, 1 if cut else 0),
                       "pmp_grid_set_corner_cutting")
            self._cut = bool(cut)

    def read_bits(self):
        """bool array [x, y(, z)] of the device occupancy (box only)."""
        n = self.words_per_row * self.padded[1] * self.padded[2]
        w = np.zeros(n, dtype=np.uint32)
        self.torch.cuda.synchronize()
        _lib.check(self.L.pmp_grid_read_bits(self.handle, w.ctypes.data, n), "pmp_grid_read_bits")
        bits = np.unpackbits(w.view(np.uint8), bitorder="little").reshape(
            self.padded[2], self.padded[1], self.words_per_row * 32)
        X, Y = self.dims[0], self.dims[1]
        Z = self.dims[2] if self.ndim == 3 else 1
        box = bits[:Z, :Y, :X].astype(bool)
        return box.T.copy() if self.ndim == 3 else box[0].T.copy()

    def count_free(self):
        import ctypes
        n = ctypes.c_int64(0)
        self.torch.cuda.synchronize()
        _lib.check(self.L.pmp_grid_count_free(self.handle, ctypes.byref(n)), "pmp_grid_count_free")
        return n.value


class _EnvBase:
    """Shared device-sync logic of Grid and Grid3D."""

    _ndim = 2
    # sets up to this size are re-uploaded on every plan(), so in-place edits of
    # `env.obstacles` without update() behave like the reference (planners alias the set,
    # graph_search.py:26-28); larger sets are synced by update() only
    _ALWAYS_SYNC_LIMIT = 200_000

    def _dims(self):
        return (self.x_range, self.y_range) + ((self.z_range,) if self._ndim == 3 else ())

    def _init_state(self):
        self._obstacles = None
        self._packed = None       # uint32 bit rows when update_packed() was used
        self._dense = None        # numpy / torch occupancy when update() got an array
        self._tree = None
        self._device = None
        self._device_version = -1
        self._version = 0
        self._synced_set = None   # copy of the obstacle set as last uploaded (small sets: detects in-place edits)

    def update_packed(self, rows):
        """Occupancy as bit rows [Z][Y][ceil(X/32)] of uint32 (bit i of word w <-> x = 32 w + i, 1 =
        obstacle; DeviceGrid.set_packed) -- the 1-bit-per-cell form of update() for callers that
        hold their map as bits already (a 1536^3 grid is 432 MiB instead of 3.4 GiB).  No
        counterpart in the reference (env.py:78-80 takes a set of tuples)."""
        self._packed = rows
        self._dense = None
        self._obstacles = None
        self._tree = None
        self._version += 1

    def _unpack(self):
        r = self._packed
        if not isinstance(r, np.ndarray):
            r = r.detach().cpu().numpy()
        dims = self._dims()
        X, Y = dims[0], dims[1]
        Z = dims[2] if self._ndim == 3 else 1
        bits = np.unpackbits(np.ascontiguousarray(r).view(np.uint8).reshape(Z, Y, -1), axis=-1, bitorder="little")[:, :, :X]
        occ = bits.astype(bool).transpose(2, 1, 0)
        return occ if self._ndim == 3 else occ[:, :, 0]

    @property
    def obstacles(self):
        if self._obstacles is None and self._packed is not None:
            self._obstacles = set(map(tuple, np.argwhere(self._unpack()).tolist()))
        if self._obstacles is None and self._dense is not None:
            d = self._dense
            if not isinstance(d, np.ndarray):
                d = d.detach().cpu().numpy()
            self._obstacles = set(map(tuple, np.argwhere(d != 0).tolist()))
        return self._obstacles

    @obstacles.setter
    def obstacles(self, value):
        self._obstacles = value
        self._dense = None
        self._tree = None
        self._version += 1

    @property
    def obstacles_tree(self):
        if self._tree is None:
            from scipy.spatial import cKDTree
            self._tree = cKDTree(np.array(list(self.obstacles)))
        return self._tree

    def update(self, obstacles):
        """env.py:78-80 / env3d.py:101-103.  `obstacles`: a set of coordinate tuples (the
        reference's form) or a dense occupancy array / torch tensor indexed [x, y(, z)]."""
        if isinstance(obstacles, (set, frozenset, list, tuple)):
            self._obstacles = obstacles if isinstance(obstacles, set) else set(obstacles)
            self._dense = None
        else:
            if tuple(obstacles.shape) != self._dims():
                raise ValueError(f"occupancy shape {tuple(obstacles.shape)} != {self._dims()}")
            self._dense = obstacles
            self._obstacles = None
        self._packed = None
        self._tree = None
        self._version += 1

    def occupancy(self):
        """dense bool array [x, y(, z)] of the box (host)."""
        if self._packed is not None:
            return self._unpack().copy()
        if self._dense is not None:
            d = self._dense
            if not isinstance(d, np.ndarray):
                d = d.detach().cpu().numpy()
            return d != 0
        occ = np.zeros(self._dims(), dtype=bool)
        if self._obstacles:
            a = np.array(list(self._obstacles), dtype=np.int64).reshape(-1, self._ndim)
            ok = np.all((a >= 0) & (a < np.array(self._dims())), axis=1)
            a = a[ok]
            occ[tuple(a.T)] = True
        return occ

    def device_grid(self, force=False, sync=True):
        """The DeviceGrid mirroring this environment, uploading the occupancy if it changed (sync=False: only
        create / return the handle -- callers that want their own small uploads queued ahead of a large streamed
        occupancy upload call it once without and once with sync)."""
        if self._device is None:
            self._device = DeviceGrid(self._ndim, self._dims())
        dg = self._device
        if not sync:
            return dg
        small_set = self._dense is None and self._packed is None and len(self._obstacles) <= self._ALWAYS_SYNC_LIMIT
        if small_set and not force and self._device_version == self._version and self._obstacles == self._synced_set:
            return dg          # in-place edits are looked for on every plan(); nothing changed since the last upload
        if force or small_set or self._device_version != self._version:
            if self._packed is not None:
                dg.set_packed(self._packed)
            elif self._dense is not None:
                dg.set_dense(self._dense)
            else:
                self._synced_set = set(self._obstacles) if small_set else None
                n = len(self._obstacles)
                coords = np.fromiter(itertools.chain.from_iterable(self._obstacles), dtype=np.int32,
                                     count=n * self._ndim).reshape(n, self._ndim)
                dg.set_cells(coords)
            self._device_version = self._version
        return dg


class Env(_EnvBase):
    """2D workspace (env.py:14-39)."""

    def __init__(self, x_range: int, y_range: int, eps: float = 1e-6) -> None:
        self.x_range = x_range
        self.y_range = y_range
        self.eps = eps
        self._init_state()

    @property
    def grid_map(self) -> set:
        return {(i, j) for i in range(self.x_range) for j in range(self.y_range)}


class Grid(Env):
    """Discrete 2D grid map (env.py:41-80): 8 motions with cost 1 / sqrt(2), border walls."""

    _ndim = 2

    def __init__(self, x_range: int, y_range: int) -> None:
        super().__init__(x_range, y_range)
        self.motions = [Node((-1, 0), None, 1, None), Node((-1, 1), None, sqrt(2), None),
                        Node((0, 1), None, 1, None), Node((1, 1), None, sqrt(2), None),
                        Node((1, 0), None, 1, None), Node((1, -1), None, sqrt(2), None),
                        Node((0, -1), None, 1, None), Node((-1, -1), None, sqrt(2), None)]
        self.init()

    def init(self) -> None:
        x, y = self.x_range, self.y_range
        obstacles = set()
        for i in range(x):
            obstacles.add((i, 0))
            obstacles.add((i, y - 1))
        for i in range(y):
            obstacles.add((0, i))
            obstacles.add((x - 1, i))
        self.update(obstacles)


class Env3D(_EnvBase):
    """3D workspace (env3d.py:14-41)."""

    _ndim = 3

    def __init__(self, x_range: int, y_range: int, z_range: int, eps: float = 1e-6) -> None:
        self.x_range = x_range
        self.y_range = y_range
        self.z_range = z_range
        self.eps = eps
        self._init_state()

    @property
    def grid_map(self):
        return BoxCells(self.x_range, self.y_range, self.z_range)


class BoxCells:
    """`grid_map` of a 3D box: behaves like the reference's set of all cells (env3d.py:35-37)
    for membership, len and iteration, without materialising X*Y*Z tuples per access."""

    def __init__(self, x, y, z):
        self._r = (x, y, z)

    def __contains__(self, p):
        try:
            return len(p) == 3 and all(isinstance(v, (int, np.integer)) and 0 <= v < r
                                       for v, r in zip(p, self._r))
        except TypeError:
            return False

    def __len__(self):
        return self._r[0] * self._r[1] * self._r[2]

    def __iter__(self):
        return ((i, j, o) for i in range(self._r[0]) for j in range(self._r[1])
                for o in range(self._r[2]))


class Grid3D(Env3D):
    """Discrete 3D voxel grid (env3d.py:43-103): 26 motions (1 / sqrt 2 / sqrt 3) and the
    reference's default walls, including its quirk of writing the side walls at z - 1
    (env3d.py:86-92)."""

    def __init__(self, x_range: int, y_range: int, z_range: int, walls: bool = True) -> None:
        super().__init__(x_range, y_range, z_range)
        s2, s3 = sqrt(2), sqrt(3)
        self.motions = [
            Node3D((-1, 0, 0), None, 1, None), Node3D((-1, 1, 0), None, s2, None),
            Node3D((0, 1, 0), None, 1, None), Node3D((1, 1, 0), None, s2, None),
            Node3D((1, 0, 0), None, 1, None), Node3D((1, -1, 0), None, s2, None),
            Node3D((0, -1, 0), None, 1, None), Node3D((-1, -1, 0), None, s2, None),
            Node3D((0, 0, 1), None, 1, None), Node3D((0, 0, -1), None, 1, None),
            Node3D((-1, 0, 1), None, s2, None), Node3D((-1, 1, 1), None, s3, None),
            Node3D((0, 1, 1), None, s2, None), Node3D((1, 1, 1), None, s3, None),
            Node3D((1, 0, 1), None, s2, None), Node3D((1, -1, 1), None, s3, None),
            Node3D((0, -1, 1), None, s2, None), Node3D((-1, -1, 1), None, s3, None),
            Node3D((-1, 0, -1), None, s2, None), Node3D((-1, 1, -1), None, s3, None),
            Node3D((0, 1, -1), None, s2, None), Node3D((1, 1, -1), None, s3, None),
            Node3D((1, 0, -1), None, s2, None), Node3D((1, -1, -1), None, s3, None),
            Node3D((0, -1, -1), None, s2, None), Node3D((-1, -1, -1), None, s3, None)]
        if walls:
            self.init()
        else:                      # huge grids: the caller supplies a dense array via update()
            self.update(set())

    def init(self) -> None:
        x, y, z = self.x_range, self.y_range, self.z_range
        obstacles = set()
        for _z in range(z):
            for i in range(x):
                obstacles.add((i, 0, _z - 1))
                obstacles.add((i, y - 1, _z - 1))
            for i in range(y):
                obstacles.add((0, i, _z - 1))
                obstacles.add((x - 1, i, _z - 1))
        for _x in range(x):
            for _y in range(y):
                obstacles.add((_x, _y, 0))
                obstacles.add((_x, _y, z - 1))
        self.update(obstacles)


def motion_table(env, ndim):
    """(offsets, costs) of env.motions in device form.  The 2D planners accumulate `motion.g`
    (node.py:41); the 3D planners use the Euclidean length of the offset (a_star3d.py:67,
    planner3d.py:22-27)."""
    offsets, costs = [], []
    for m in env.motions:
        off = tuple(int(v) for v in m.current)
        offsets.append(off)
        if ndim == 2:
            costs.append(float(m.g))
        else:
            costs.append(math.sqrt(sum(v * v for v in off)))
    return offsets, costs
