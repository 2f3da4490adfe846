"""ctypes binding of libpmp_b200.so (include/pmp_b200.h).

There is no CPU fallback: if the shared library or a CUDA device is missing, every compute
entry point raises.  PyTorch is used only to own device buffers and streams.
"""
import ctypes
import os
import subprocess

_HERE = os.path.dirname(os.path.abspath(__file__))
# PMP_B200_LIB selects an alternative build of the same library (tuning experiments)
LIB_PATH = os.environ.get("PMP_B200_LIB") or os.path.join(_HERE, "libpmp_b200.so")
CSRC = os.path.join(_HERE, "csrc")

ABI_VERSION = 3                              # PMP_B200_ABI_VERSION this binding was written against
ASTAR, DIJKSTRA, GBFS = 0, 1, 2
THETA_STAR, LAZY_THETA_STAR, JPS = 3, 4, 5    # pmp_bestfirst only
BESTFIRST_HEAP_REC_BYTES = 40
H_EUCLIDEAN, H_MANHATTAN = 0, 1
W_F64, W_U32 = 0, 1
ST_RUNNING, ST_EXHAUSTED, ST_REACHED, ST_BAD_QUERY, ST_REPLAN_OVERFLOW, ST_ERROR = 0, 1, 2, 3, 4, 5


def check_status(status, what="search"):
    """raise PmpError when a status word (or array of them) reports an aborted search -- a queue
    wait timed out inside the kernel, so whatever the field holds must not be used"""
    import numpy as np
    if np.any(np.asarray(status) == ST_ERROR):
        raise PmpError(f"{what}: the search kernel aborted (queue wait timed out); no result")

# every symbol include/pmp_b200.h declares (tests/test_abi.py checks the header against this)
SYMBOLS = (
    "pmp_last_error", "pmp_abi_version", "pmp_device_count",
    "pmp_grid_create", "pmp_grid_destroy", "pmp_grid_layout", "pmp_grid_set_cells",
    "pmp_grid_set_dense", "pmp_grid_set_dense_xyz", "pmp_grid_set_packed", "pmp_grid_set_packed_host_streamed", "pmp_grid_edit_cells", "pmp_grid_read_bits", "pmp_grid_count_free",
    "pmp_grid_set_motions", "pmp_grid_set_corner_cutting", "pmp_sssp_sizes", "pmp_sssp", "pmp_field_finalize",
    "pmp_count_settled", "pmp_extract_paths", "pmp_verify_field", "pmp_gbfs", "pmp_bestfirst",
    "pmp_lpa_star", "pmp_d_star",
    "pmp_slab_begin", "pmp_slab_run", "pmp_slab_merge", "pmp_replan_begin", "pmp_replan_update",
    "pmp_line_of_sight", "pmp_shortcut_paths",
    "pmp_bfs_slab_create", "pmp_bfs_slab_destroy", "pmp_bfs_slab_export", "pmp_bfs_slab_connect",
    "pmp_bfs_slab_connect_local", "pmp_bfs_slab_begin", "pmp_bfs_slab_run", "pmp_bfs_slab_run_group",
    "pmp_bfs_slab_buffers", "pmp_cost_field_u32_static",
)


class PmpError(RuntimeError):
    pass


class SsspParams(ctypes.Structure):
    _fields_ = [("algo", ctypes.c_int32), ("heuristic", ctypes.c_int32), ("weight", ctypes.c_int32),
                ("full_field", ctypes.c_int32), ("window", ctypes.c_double),
                ("max_rounds", ctypes.c_int32), ("reserved", ctypes.c_int32)]


def build(force=False, verbose=False):
    """Compile csrc/*.cu for sm_100a into libpmp_b200.so (nvcc cross-compiles without a GPU)."""
    srcs = [os.path.join(CSRC, f) for f in os.listdir(CSRC) if f.endswith((".cu", ".cuh", ".h"))]
    srcs.append(os.path.join(_HERE, "..", "include", "pmp_b200.h"))
    if (not force and os.path.exists(LIB_PATH)
            and all(os.path.getmtime(LIB_PATH) >= os.path.getmtime(s) for s in srcs)):
        return LIB_PATH
    cmd = ["make", "-C", CSRC, "-j4"] + (["-B"] if force else [])
    res = subprocess.run(cmd, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True)
    if verbose or res.returncode != 0:
        print(res.stdout)
    if res.returncode != 0:
        raise PmpError("building libpmp_b200.so failed")
    return LIB_PATH


_lib = None


def lib():
    """The loaded library.  Raises PmpError when it is absent (no fallback path exists)."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise PmpError(f"{LIB_PATH} not found: run `python -c 'import __graft_entry__ as g; g.build()'` "
                       "(or make -C python_motion_planning_b200/csrc). There is no CPU fallback.")
    L = ctypes.CDLL(LIB_PATH)
    vp, i32, i64, dbl = ctypes.c_void_p, ctypes.c_int, ctypes.c_int64, ctypes.c_double
    L.pmp_last_error.restype = ctypes.c_char_p
    L.pmp_last_error.argtypes = []
    L.pmp_abi_version.restype = i32
    if L.pmp_abi_version() != ABI_VERSION:
        raise PmpError(f"{LIB_PATH} has ABI version {L.pmp_abi_version()}, this package needs {ABI_VERSION}: "
                       "rebuild it (python -c 'import __graft_entry__ as g; g.build()')")
    L.pmp_device_count.restype = i32
    sig = {
        "pmp_grid_create": [i32, vp, ctypes.POINTER(vp)],
        "pmp_grid_destroy": [vp],
        "pmp_grid_layout": [vp, vp],
        "pmp_grid_set_cells": [vp, vp, i64, vp],
        "pmp_grid_set_dense": [vp, vp, vp],
        "pmp_grid_set_dense_xyz": [vp, vp, vp],
        "pmp_grid_set_packed": [vp, vp, vp],
        "pmp_grid_set_packed_host_streamed": [vp, vp, i32, vp, vp],
        "pmp_grid_edit_cells": [vp, vp, i64, i32, vp],
        "pmp_grid_read_bits": [vp, vp, i64],
        "pmp_grid_count_free": [vp, vp],
        "pmp_grid_set_motions": [vp, i32, vp, vp],
        "pmp_grid_set_corner_cutting": [vp, i32],
        "pmp_sssp_sizes": [vp, i64, i32, vp, vp],
        "pmp_sssp": [vp, ctypes.POINTER(SsspParams), i64, vp, vp, vp, vp, i64, vp, vp, vp, vp],
        "pmp_field_finalize": [vp, i32, i64, vp],
        "pmp_count_settled": [vp, i32, i64, vp, vp, vp, i32, vp, vp],
        "pmp_extract_paths": [vp, i32, i64, vp, vp, vp, vp, i64, vp, vp, vp],
        "pmp_verify_field": [vp, i32, vp, vp, i32, i32, vp, vp],
        "pmp_gbfs": [vp, i32, i64, vp, vp, vp, vp, i64, vp, i64, vp, vp, vp, vp],
        "pmp_bestfirst": [vp, i32, i32, i64, vp, vp, vp, i64, vp, vp, vp, vp, i64, vp, i64, vp, vp, vp, vp],
        "pmp_lpa_star": [vp, i32, i32, i32, vp, vp, vp, vp, vp, vp, i64, vp, vp, i64, vp, i64, vp, vp, vp, vp],
        "pmp_d_star": [vp, i32, vp, vp, vp, vp, vp, vp, i64, vp, vp, i64, vp, i64, vp, vp, vp, vp],
        "pmp_slab_begin": [vp, ctypes.POINTER(SsspParams), vp, vp, vp, i64, vp],
        "pmp_slab_run": [vp, ctypes.POINTER(SsspParams), vp, vp, i64, dbl, vp, vp],
        "pmp_slab_merge": [vp, i32, vp, vp, i64, i32, i32, vp, vp, vp],
        "pmp_bfs_slab_create": [vp, i32, i32, ctypes.POINTER(vp)],
        "pmp_bfs_slab_destroy": [vp],
        "pmp_bfs_slab_export": [vp, vp, vp],
        "pmp_bfs_slab_connect": [vp, vp, vp],
        "pmp_bfs_slab_connect_local": [vp, i32],
        "pmp_bfs_slab_begin": [vp, vp, vp, vp],
        "pmp_bfs_slab_run": [vp, vp, vp, vp],
        "pmp_bfs_slab_run_group": [vp, i32, vp],
        "pmp_bfs_slab_buffers": [vp, vp, vp, vp, vp],
        "pmp_cost_field_u32_static": [vp, vp, vp, vp, vp, vp],
        "pmp_line_of_sight": [vp, i64, vp, vp, vp, vp],
        "pmp_shortcut_paths": [vp, i64, vp, i64, vp, vp, vp, vp, vp],
        "pmp_replan_begin": [vp, ctypes.POINTER(SsspParams), vp, vp, vp, vp, i64, vp, vp, vp, vp],
        "pmp_replan_update": [vp, ctypes.POINTER(SsspParams), vp, vp, vp, vp, i64, vp, vp, i64, vp, i64,
                              vp, vp, vp, vp],
    }
    for name, argtypes in sig.items():
        fn = getattr(L, name)
        fn.restype = i32
        fn.argtypes = argtypes
    _lib = L
    return L


def check(rc, what):
    if rc != 0:
        raise PmpError(f"{what} failed ({rc}): {lib().pmp_last_error().decode(errors='replace')}")


def require_cuda():
    """torch + a visible CUDA device, or a loud failure."""
    import torch
    if not torch.cuda.is_available():
        raise PmpError("python_motion_planning_b200 needs a CUDA device (B200, sm_100a); "
                       "there is no CPU fallback")
    return torch
