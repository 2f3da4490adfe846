"""Import path of the reference's `python_motion_planning.utils` (utils/__init__.py:1-25) for the names on
this package's path: `from python_motion_planning_b200.utils import Grid, SearchFactory` works like
examples/global_examples.py:9 of the reference.  Map / Point / Pose / Robot / CurveFactory /
ControlFactory belong to subsystems outside the graph-search path and are not provided."""
from ..environment import Env, Env3D, Grid, Grid3D, Node, Node3D
from ..planner import HeadlessPlot as Plot
from ..planner import HeadlessPlot as Plot3D
from ..planner import Planner, Planner3D, SearchFactory

__all__ = ["Env", "Env3D", "Grid", "Grid3D", "Node", "Node3D", "Plot", "Plot3D", "Planner", "Planner3D",
           "SearchFactory"]
