"""Time the UNMODIFIED reference planners staged under oracle/_ref (oracle/stage_ref.py) -- TEST /
BASELINE INFRASTRUCTURE ONLY; only bench.py's cpu_baseline leg and tests may import this.

The reference imports matplotlib / pyvista / cvxopt / osqp at module scope and builds a plotter
in every planner constructor (utils/planner/planner.py:20, planner3d.py:20); none of them is
installed, so attribute-absorbing stub modules stand in.  No arithmetic of
the graph-search path lives in those modules.  Only plan() is called and timed, exactly as
examples/benchmark.py:79-83 does."""
import importlib.machinery
import logging
import os
import statistics
import sys
import time
import types
from unittest import mock

HERE = os.path.dirname(os.path.abspath(__file__))
REF_SRC = os.path.join(HERE, "_ref", "src")
_STUBS = ("matplotlib", "matplotlib.pyplot", "matplotlib.patches", "matplotlib.animation", "matplotlib.colors",
          "matplotlib.cm", "mpl_toolkits", "mpl_toolkits.mplot3d", "mpl_toolkits.mplot3d.art3d", "pyvista", "cvxopt",
          "osqp")


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


def load():
    if not os.path.isdir(os.path.join(REF_SRC, "python_motion_planning")):
        raise RuntimeError("oracle/_ref is not staged (python oracle/stage_ref.py in the build container)")
    for name in _STUBS:
        if name not in sys.modules:
            sys.modules[name] = _Stub(name)
    if REF_SRC not in sys.path:
        sys.path.insert(0, REF_SRC)
    import python_motion_planning as ref
    logging.disable(logging.CRITICAL)
    return ref


def _median_ms(fn, reps):
    ts = []
    out = None
    for _ in range(reps):
        t0 = time.perf_counter()
        out = fn()
        ts.append((time.perf_counter() - t0) * 1e3)
    return statistics.median(ts), out


def time_c1_c2():
    """C1: README example, AStar.plan() (README.md:62-76); C2: examples/3d_example.py defaults,
    AStar3D.plan() on Grid3D(26,20,16) with the 'door' scene.  Single-threaded, like the reference."""
    ref = load()
    env = ref.Grid(51, 31)
    obs = env.obstacles
    for i in range(10, 21):
        obs.add((i, 15))
    for i in range(15):
        obs.add((20, i))
    for i in range(15, 30):
        obs.add((30, i))
    for i in range(16):
        obs.add((40, i))
    env.update(obs)
    ms1, r1 = _median_ms(lambda: ref.AStar(start=(5, 5), goal=(45, 25), env=env).plan(), 10)
    import numpy as np
    from python_motion_planning_b200 import workloads
    occ, dims = workloads.door3d()
    from python_motion_planning.utils.environment.env3d import Grid3D
    from python_motion_planning.global_planner.graph_search.a_star3d import AStar3D
    e3 = Grid3D(*dims)
    e3.update(set(map(tuple, np.argwhere(occ).tolist())))
    ms2, r2 = _median_ms(lambda: AStar3D(start=(1, 1, 1), goal=(24, 18, 14), env=e3).plan(), 5)
    k1, k2 = 8, 26
    return {"kind": "reference-python", "cores": 1,
            "c1": {"what": "README Grid(51,31) AStar.plan(), median of 10", "ms": ms1, "cost": float(r1[0]),
                   "expanded": len(r1[2]), "gteps": k1 * len(r1[2]) / (ms1 * 1e-3) / 1e9},
            "c2": {"what": "Grid3D(26,20,16) 'door' AStar3D.plan(), median of 5", "ms": ms2, "cost": float(r2[0]),
                   "expanded": len(r2[2]), "gteps": k2 * len(r2[2]) / (ms2 * 1e-3) / 1e9}}


if __name__ == "__main__":
    import json
    sys.path.insert(0, os.path.dirname(HERE))
    print(json.dumps(time_c1_c2()))
