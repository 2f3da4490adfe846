"""Import path of the reference's `python_motion_planning.global_planner.graph_search`
(global_planner/graph_search/__init__.py:1-46), as used by examples/algorithms.py:1 of the reference.
VoronoiPlanner(3D) and SThetaStar are outside SURVEY.md section 8 and are not provided."""
from ...anyangle import LazyThetaStar, LazyThetaStar3D, ThetaStar, ThetaStar3D
from ...jps import JPS, JPS3D
from ...planner import AStar, AStar3D, Dijkstra, Dijkstra3D, GBFS, GBFS3D, GraphSearcher, GraphSearcher3D
from ...replan import DStar, DStar3D, DStarLite, LPAStar, LPAStar3D

__all__ = ["AStar", "AStar3D", "Dijkstra", "Dijkstra3D", "GBFS", "GBFS3D", "JPS", "JPS3D", "DStar", "DStar3D",
           "LPAStar", "LPAStar3D", "DStarLite", "ThetaStar", "ThetaStar3D", "LazyThetaStar", "LazyThetaStar3D",
           "GraphSearcher", "GraphSearcher3D"]
