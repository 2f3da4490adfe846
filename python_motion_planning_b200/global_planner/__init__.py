"""Import path of the reference's `python_motion_planning.global_planner` (graph search only)."""
from .graph_search import *  # noqa: F401,F403
