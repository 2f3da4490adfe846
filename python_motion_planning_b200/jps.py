"""JPS / JPS3D.  Two modes, like the Theta* classes (anyangle.py):

* `mode="reference"` -- jump point search as the reference runs it (`pmp_bestfirst` with algo JPS,
  csrc/pmp_jps_core.h: JPS.jump / detectForceNeighbor, jps.py:85-165; JPS3D.jump /
  _has_forced_neighbor, jps3d.py:99-246), one warp per query, one whole jump per lane.  cost, the
  jump-point path, len(expand) and every expand record equal the reference's -- including the
  places where the reference is not optimal: its 2D jump() never calls isCollision (diagonal moves
  cut corners) and its 3D forced-neighbour rules prune too much (longer than AStar3D in a third of
  the seeded queries, tools/make_golden_jps.py).  Default up to 128 Ki padded cells.
* `mode="wavefront"` -- for large maps: jump point search is a pruning technique for sequential A*;
  the tiled wavefront of k_sssp is this package's accelerator for the same job.  `JPS` runs it with
  the edge rule of the reference's jump() (`pmp_grid_set_corner_cutting`), so the cost equals the
  reference's (the optimum of that graph; tests/golden/jps.npz) and the path holds the turning
  points of the grid path (the reference lists its jump points: same polyline, possibly more
  vertices).  `JPS3D` returns the optimum: cost <= the reference's.  Default above 128 Ki cells.
"""
from .environment import Node, Node3D
from .planner import GraphSearcher, GraphSearcher3D, _plan_bestfirst
from . import _lib


def turning_points(path):
    """endpoints + every cell where the step direction changes"""
    if len(path) <= 2:
        return list(path)
    out = [path[0]]
    for a, b, c in zip(path[:-2], path[1:-1], path[2:]):
        if tuple(y - x for x, y in zip(a, b)) != tuple(y - x for x, y in zip(b, c)):
            out.append(b)
    out.append(path[-1])
    return out


REFERENCE_MODE_MAX_CELLS = 1 << 17


class _JpsMixin:
    def __init__(self, start, goal, env, heuristic_type: str = "euclidean", *, mode: str = None) -> None:
        super().__init__(start, goal, env, heuristic_type)
        self.mode = mode   # "reference" | "wavefront" | None: by grid size (REFERENCE_MODE_MAX_CELLS)

    def plan(self) -> tuple:
        dg = self._prepare()
        mode = self.mode or ("reference" if dg.pcells <= REFERENCE_MODE_MAX_CELLS else "wavefront")
        if mode == "reference":
            return _plan_bestfirst(self, _lib.JPS)
        if mode != "wavefront":
            raise ValueError("mode must be 'reference', 'wavefront' or None")
        cost, path, expand = self._plan_sssp()
        return cost, turning_points(path), expand


class JPS(_JpsMixin, GraphSearcher):
    """jps.py:14-165"""
    _algo = _lib.ASTAR
    _cut = True            # the graph jump() walks: only the target cell must be free
    _node_cls = Node

    def __str__(self) -> str:
        return "Jump Point Search(JPS)"

    def run(self):
        cost, path, expand = self.plan()
        self.plot.animation(path, str(self), cost, expand)


class JPS3D(_JpsMixin, GraphSearcher3D):
    """jps3d.py:26-261"""
    _algo = _lib.ASTAR
    _node_cls = Node3D

    def __str__(self) -> str:
        return "Jump Point Search (JPS) 3D"

    def run(self):
        cost, path, expand = self.plan()
        self.plot.animation(path, str(self), cost, expand)
