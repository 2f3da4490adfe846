// pmp_griddev.h -- the device-side description of a grid (no CUDA dependency: also included by
// the CPU test harness of the unit-weight field kernel, tests/bfs_host.cpp).
#pragma once
#include <stdint.h>

typedef unsigned long long u64;
typedef uint32_t u32;

struct GridDev {
    int ndim;
    int X, Y, Z;       // box
    int PX, PY, PZ;    // padded to whole tiles
    int TX, TY, TZ;    // tile
    int ntx, nty, ntz; // tiles per axis
    int wpr;           // 32-bit words per bitmask row
    u32 ntiles;
    u64 pcells;
    const u32 *bits; // occupancy, 1 = blocked; row (y,z) at ((z*PY)+y)*wpr
    // streamed upload in progress (pmp_grid_set_packed_host_streamed): ready[c] != 0 once the planes
    // [c << ready_shift, (c + 1) << ready_shift) of `bits` have arrived; nullptr: everything is there
    const u32 *ready;
    int ready_shift;
};

