"""B200-native grid graph search behind python_motion_planning's planner API.

    import python_motion_planning_b200 as pmp
    env = pmp.Grid(51, 31); ...; env.update(obstacles)
    cost, path, expand = pmp.AStar(start=(5, 5), goal=(45, 25), env=env).plan()

Importing the package needs neither the shared library nor a GPU; every compute call does and
raises loudly otherwise (there is no CPU fallback).
"""
from .environment import Env, Env3D, Grid, Grid3D, Node, Node3D
from .planner import (AStar, AStar3D, Dijkstra, Dijkstra3D, GBFS, GBFS3D, GraphSearcher,
                      GraphSearcher3D, Planner, Planner3D, SearchFactory, ExpandView,
                      set_reference_order)
from .batch import plan_batch, plan_batch_reference, plan_batch_sharded, cost_field, shard_range
from .replan import LPAStar, LPAStar3D, DStar, DStar3D, DStarLite, DynamicField, ReplanSession
from .jps import JPS, JPS3D
from .anyangle import ThetaStar, ThetaStar3D, LazyThetaStar, LazyThetaStar3D, line_of_sight
from ._lib import PmpError

__all__ = ["Env", "Env3D", "Grid", "Grid3D", "Node", "Node3D", "AStar", "AStar3D", "Dijkstra",
           "Dijkstra3D", "GBFS", "GBFS3D", "GraphSearcher", "GraphSearcher3D", "Planner",
           "Planner3D", "SearchFactory", "ExpandView", "plan_batch", "plan_batch_reference", "plan_batch_sharded", "shard_range", "cost_field", "PmpError",
           "LPAStar", "LPAStar3D", "DStar", "DStar3D", "DStarLite", "DynamicField", "ReplanSession",
           "set_reference_order", "JPS", "JPS3D", "ThetaStar", "ThetaStar3D", "LazyThetaStar", "LazyThetaStar3D", "line_of_sight"]
