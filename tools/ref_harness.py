"""Import the UNMODIFIED reference (python_motion_planning at /root/reference) in this
container, for pinning the oracle and generating golden vectors.

TEST/FIXTURE INFRASTRUCTURE ONLY.  Nothing on the product path, in bench.py or in the
`-m gpu` tests may import this module: /root/reference does not exist on the GPU box.

The reference constructs a plotter in every planner constructor
(src/python_motion_planning/utils/planner/planner.py:20, planner3d.py:20) and imports
matplotlib / pyvista / cvxopt / osqp at module scope; none is installed here, so they are
replaced by attribute-absorbing stub modules.  No arithmetic of the
graph-search path lives in those modules.
"""
import importlib.machinery
import logging
import sys
import types
from unittest import mock

REFERENCE_ROOT = "/root/reference"

_STUBS = (
    "matplotlib", "matplotlib.pyplot", "matplotlib.patches", "matplotlib.animation",
    "matplotlib.colors", "matplotlib.cm", "mpl_toolkits", "mpl_toolkits.mplot3d",
    "mpl_toolkits.mplot3d.art3d", "pyvista", "cvxopt", "osqp",
)


class _Stub(types.ModuleType):
    def __init__(self, name):
        super().__init__(name)
        self.__path__ = []
        self.__spec__ = importlib.machinery.ModuleSpec(name, None, is_package=True)

    def __getattr__(self, item):
        if item.startswith("__") and item.endswith("__"):
            raise AttributeError(item)
        value = mock.MagicMock(name=f"{self.__name__}.{item}")
        setattr(self, item, value)
        return value


def load_reference():
    """Returns the imported `python_motion_planning` package of the reference."""
    for name in _STUBS:
        if name not in sys.modules:
            sys.modules[name] = _Stub(name)
    for p in (REFERENCE_ROOT + "/src", REFERENCE_ROOT + "/examples"):
        if p not in sys.path:
            sys.path.insert(0, p)
    import python_motion_planning as pmp  # noqa: E402
    logging.disable(logging.CRITICAL)     # the 3D planners log every 1000 expansions
    return pmp


class BoxMap:
    """O(1) stand-in for Env3D.grid_map (env3d.py:35-37 rebuilds a set of every cell on each
    access; Dijkstra3D.getNeighbor reads it once per expansion, dijkstra3d.py:99).  Only used
    for membership tests (dijkstra3d.py:106), so this changes no arithmetic."""

    def __init__(self, x, y, z):
        self.r = (x, y, z)

    def __contains__(self, p):
        return 0 <= p[0] < self.r[0] and 0 <= p[1] < self.r[1] and 0 <= p[2] < self.r[2]

    def __iter__(self):   # LPAStar3D / DStar3D build their node map from it (lpa_star3d.py:68, d_star3d.py:87)
        return ((x, y, z) for x in range(self.r[0]) for y in range(self.r[1]) for z in range(self.r[2]))


def fast_grid_map(pmp):
    """Patch Grid3D.grid_map in THIS process with the O(1) membership object."""
    from python_motion_planning.utils.environment.env3d import Env3D
    Env3D.grid_map = property(lambda self: BoxMap(self.x_range, self.y_range, self.z_range))
