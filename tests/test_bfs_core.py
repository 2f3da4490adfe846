"""CPU check of the unit-weight field kernel (csrc/pmp_bfs_core.cuh) on a SIMT emulator.

tests/bfs_host.cpp compiles the kernel header UNCHANGED with g++ against tests/simt_host.h
(every lane a fiber, warp collectives as rendezvous, warps interleaved at random at every atomic)
and compares the field with a plain queue BFS over the six unit moves -- what the reference's
Dijkstra3D computes with env.motions filtered to cost 1 (dijkstra3d.py:39-126).  Single-slab and
multi-slab ("ranks" sharing the address space: the peer-memory protocol of the multi-GPU path,
incl. distributed termination) cases, awkward shapes (x extent not a multiple of 32, thin grids),
dense random obstacles, boxes, serpentine corridors longer than one recording pass.
The kernel itself is checked on the GPU in tests/test_gpu_parity.py / tests/test_bfs_gpu.py.
"""
import os
import subprocess

import pytest

HERE = os.path.dirname(os.path.abspath(__file__))
SRC = os.path.join(HERE, "bfs_host.cpp")
DEPS = [SRC, os.path.join(HERE, "simt_host.h"),
        os.path.join(HERE, "..", "python_motion_planning_b200", "csrc", "pmp_bfs_core.cuh"),
        os.path.join(HERE, "..", "python_motion_planning_b200", "csrc", "pmp_griddev.h")]
EXE = os.path.join(HERE, "bfs_host.bin")


@pytest.fixture(scope="module")
def exe():
    newest = max(os.path.getmtime(f) for f in DEPS)
    if not os.path.exists(EXE) or os.path.getmtime(EXE) < newest:
        subprocess.run(["g++", "-O1", "-std=c++17", "-Wall", "-Wno-unknown-pragmas", "-Wno-unused-function",
                        "-Wno-unused-variable", "-o", EXE, SRC], check=True)
    return EXE


def run(exe, *a):
    r = subprocess.run([exe] + [str(v) for v in a], stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True,
                       timeout=600)
    assert r.returncode == 0 and r.stdout.startswith("OK"), r.stdout[-2000:]
    return dict(kv.split("=") for kv in r.stdout.split()[1:])


#            X    Y   Z  seed dens ranks warps mode
SINGLE = [(40, 24, 20, 1, 10, 1, 2, 0),
          (48, 24, 20, 2, 20, 1, 3, 0),       # x extent = 1.5 tiles: the half-filled last x tile
          (70, 30, 40, 3, 35, 1, 4, 0),       # dense random obstacles: many re-activations
          (33, 9, 9, 4, 5, 1, 1, 0),
          (100, 50, 30, 5, 10, 1, 4, 1),
          (45, 17, 23, 7, 0, 1, 2, 3),        # empty box
          (40, 24, 20, 8, 0, 1, 2, 2),        # serpentine: in-tile paths beyond one pass of 128 levels
          (96, 40, 24, 9, 0, 1, 4, 2)]
MULTI = [(40, 24, 40, 1, 10, 2, 2, 0),
         (48, 24, 64, 2, 20, 3, 3, 0),
         (70, 30, 40, 3, 35, 2, 4, 0),
         (33, 9, 33, 4, 5, 4, 1, 0),
         (64, 64, 64, 6, 20, 8, 2, 1),        # 8 slabs of one tile row each
         (40, 16, 72, 9, 0, 4, 3, 2)]


@pytest.mark.parametrize("cfg", SINGLE)
def test_field_single_slab(exe, cfg):
    st = run(exe, *cfg)
    assert int(st["bad"]) == 0
    assert int(st["lowered"]) >= int(st["reached"])


@pytest.mark.parametrize("cfg", MULTI)
def test_field_multi_slab(exe, cfg):
    st = run(exe, *cfg)
    assert int(st["bad"]) == 0


@pytest.mark.parametrize("cfg", [(64, 24, 64, 31, 20, 1, 3, 0, 3), (40, 16, 48, 32, 10, 1, 2, 1, 4), (96, 16, 40, 33, 0, 1, 4, 2, 3)])
def test_streamed_occupancy(exe, cfg):
    """GridDev.ready: the occupancy arrives chunk by chunk (a copy-engine fiber) while the search runs; a tile
    waits for the chunks its rows lie in and never reads words that have not landed (they hold garbage)"""
    st = run(exe, *cfg)
    assert int(st["bad"]) == 0


def test_interleavings(exe):
    """the same case under different warp interleavings (the seed drives the scheduler too)"""
    for seed in range(20, 28):
        run(exe, 64, 24, 24, seed, 25, 2, 3, 0)


def test_edge_candidates_from_two_faces(exe):
    """dense random obstacles on boxes one tile wide in y / z and several tiles long in x: most candidate cells of the
    lane-0 / lane-31 columns sit on an edge and get a level from an x face AND from a y / z face -- the bit-sliced
    min that merges the two seed sets (bfs_process_tile) decides every one of them"""
    for seed, dens in ((41, 30), (42, 40), (43, 20), (44, 35)):
        st = run(exe, 128, 8, 8, seed, dens, 1, 3, 0)
        assert int(st["bad"]) == 0
        st = run(exe, 96, 16, 8, seed, dens, 1, 4, 0)
        assert int(st["bad"]) == 0


def test_streamed_put_back_keeps_the_count(exe):
    """a slow upload (small chunks) with several warps: tiles whose planes are missing go back into the ring many times;
    the in-flight count must survive that (the search ends, nothing is left pending)"""
    for seed in (51, 52):
        st = run(exe, 64, 16, 96, seed, 15, 1, 4, 0, 2)
        assert int(st["bad"]) == 0


@pytest.mark.parametrize("cfg", [(40, 9, 8, 140, 5, 1, 4, 3, 4), (40, 8, 24, 143, 5, 1, 3, 2, 4)])
def test_put_backs_are_not_activations(exe, cfg):
    """tiny grids behind a slow upload: hundreds of put-backs for a handful of tiles.  The activation budget (64 x
    tiles) must not count them -- and the put-back count can pass a ticket that was drawn earlier (found by fuzzing:
    an unsigned difference wrapped and tripped PMP_ST_ERROR)"""
    st = run(exe, *cfg)
    assert int(st["bad"]) == 0
