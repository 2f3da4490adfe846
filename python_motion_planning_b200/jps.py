"""JPS / JPS3D entry points.

Jump point search is a pruning technique for sequential A*; the tiled wavefront of k_sssp is this
package's accelerator for the same job, so the classes below keep the reference's names, signatures
and return conventions and run the ordinary device search:

* `JPS` (global_planner/graph_search/jps.py:14-165).  The reference's jump() / detectForceNeighbor
  never call isCollision: diagonal moves may cut corners, i.e. JPS searches a different graph than
  AStar.  The device search is run with that edge rule (`pmp_grid_set_corner_cutting`), so the cost
  equals the reference's (it is the optimum of that graph; committed reference outputs in
  tests/golden/jps.npz).  The returned path holds the turning points of the grid path (goal ->
  start, like jps.py:63 / a_star.py:98-117); the reference lists its jump points, which also
  include forced-neighbour cells on straight runs -- same polyline, possibly more vertices.
* `JPS3D` (jps3d.py:26-261) uses isCollision like AStar3D, but its 3D forced-neighbour rules prune
  too much: on seeded random maps it returned a longer path than AStar3D in a third of the queries
  (tools/make_golden_jps.py prints the comparison).  Here it returns the optimum: cost <= the
  reference's, path = turning points, start -> goal (jps3d.py:248-261).
"""
from .environment import Node, Node3D
from .planner import GraphSearcher, GraphSearcher3D
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


class JPS(GraphSearcher):
    """jps.py:14-165"""
    _algo = _lib.ASTAR
    _cut = True            # the graph jump() walks: only the target cell must be free
    _node_cls = Node

    def __str__(self) -> str:
        return "Jump Point Search(JPS)"

    def plan(self) -> tuple:
        cost, path, expand = self._plan_sssp()
        return cost, turning_points(path), expand

    def run(self):
        cost, path, expand = self.plan()
        self.plot.animation(path, str(self), cost, expand)


class JPS3D(GraphSearcher3D):
    """jps3d.py:26-261"""
    _algo = _lib.ASTAR
    _node_cls = Node3D

    def __str__(self) -> str:
        return "Jump Point Search (JPS) 3D"

    def plan(self) -> tuple:
        cost, path, expand = self._plan_sssp()
        return cost, turning_points(path), expand

    def run(self):
        cost, path, expand = self.plan()
        self.plot.animation(path, str(self), cost, expand)
