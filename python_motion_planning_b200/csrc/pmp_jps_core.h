// pmp_jps_core.h -- the jump functions of the reference's jump point search, for the reference-order
// search (pmp_bestfirst with algo BF_JPS; see pmp_bestfirst_core.h for the heap / CLOSED part, which
// JPS shares with A*: jps.py:38-83 is a_star.py:39-83 with getNeighbor replaced by jump(), and
// jps3d.py:42-92 is a_star3d.py:33-78 likewise).
//
//   2D  JPS.jump / detectForceNeighbor      global_planner/graph_search/jps.py:85-165
//   3D  JPS3D.jump / _mk_qnode / _has_forced_neighbor   jps3d.py:99-246
//
// Both are written recursively in the reference; only the truth value of the inner calls is used,
// so they are restated as loops:
//   * 2D: a diagonal jump asks at every step whether a straight jump along x or y finds anything
//     (jps.py:113-118); straight jumps do not recurse.
//   * 3D: jump(node, d) for a 2-axis d calls jump(nxt, d) on itself (jps3d.py:152-157, the
//     "plane-diagonal projection" of a plane diagonal is the same direction), hence it returns the
//     FIRST step whenever anything lies further along the ray, else None; a 3-axis d asks the three
//     axis and the three plane-diagonal rays at every step.  jps3d_exists() is that truth value.
// ENV supplies, besides the members listed in pmp_bestfirst_core.h:
//   bool obst(x, y, z)    cell is in env.obstacles (inside the box and occupied)
//   bool inbox(x, y, z)   cell is in env.grid_map
//   bool legal_dir(dx, dy, dz, x, y, z)   not isCollision((x,y,z), (x,y,z)+d), target inside the box
// Free of CUDA intrinsics; tests/bestfirst_host.cpp compiles it with g++.
#pragma once

#include "pmp_bestfirst_core.h"

#define BF_JPS 5

// ---- 2D ----

// detectForceNeighbor, jps.py:132-165
template <class ENV>
BF_HD static inline bool jps2d_forced(const ENV &env, int x, int y, int xd, int yd) {
    if (xd && !yd) {
        if (env.obst(x, y + 1, 0) && !env.obst(x + xd, y + 1, 0)) return true;
        if (env.obst(x, y - 1, 0) && !env.obst(x + xd, y - 1, 0)) return true;
    }
    if (!xd && yd) {
        if (env.obst(x + 1, y, 0) && !env.obst(x + 1, y + yd, 0)) return true;
        if (env.obst(x - 1, y, 0) && !env.obst(x - 1, y + yd, 0)) return true;
    }
    if (xd && yd) {
        if (env.obst(x - xd, y, 0) && !env.obst(x - xd, y + yd, 0)) return true;
        if (env.obst(x, y - yd, 0) && !env.obst(x + xd, y - yd, 0)) return true;
    }
    return false;
}

// truth value of jump(node at (x, y), straight motion (xd, yd)): jps.py:96-121 without the
// diagonal branch.  Leaving the box ends the jump (the reference's maps are walled, env.py:68-74).
template <class ENV>
BF_HD static inline bool jps2d_straight(const BfState &S, const ENV &env, int x, int y, int xd, int yd) {
    for (;;) {
        x += xd; y += yd;
        if (!env.inbox(x, y, 0) || env.obst(x, y, 0)) return false;
        if (x == S.gx && y == S.gy) return true;
        if (jps2d_forced(env, x, y, xd, yd)) return true;
    }
}

// jump(node, motion m) as called from plan() (jps.py:62-69): false = None or already in CLOSED.
// g is accumulated one Node.__add__ at a time (node.py:41), as the recursion does.
template <class ENV>
BF_HD static inline bool jps2d_eval(const BfState &S, const ENV &env, const BfRec &node, int m, double w,
                                    BfRec &out) {
    int x, y, z, xd, yd, zd;
    bf_coords(S, node.cell, x, y, z);
    env.offset(m, xd, yd, zd);
    if (!xd && !yd) return false;
    double g = node.g;
    for (;;) {
        x += xd; y += yd;
        g = g + w;
        if (!env.inbox(x, y, 0) || env.obst(x, y, 0)) return false;    // jps.py:102-103
        if (x == S.gx && y == S.gy) break;                              // jps.py:106-107
        if (xd && yd && (jps2d_straight(S, env, x, y, xd, 0) || jps2d_straight(S, env, x, y, 0, yd)))
            break;                                                      // jps.py:110-115
        if (jps2d_forced(env, x, y, xd, yd)) break;                     // jps.py:118-119
    }
    const uint32_t nc = bf_cell(S, x, y, 0);
    if (S.cpar[nc] != BF_NONE) return false;                            // jps.py:65
    out.cell = nc;
    out.parent = node.cell;
    out.g = g;
    out.h = bf_h(S.heuristic, S.gx - x, S.gy - y, 0);
    out.f = out.g + out.h;
    out.counter = 0;
    out.pad = 0;
    return true;
}

// ---- 3D ----

// _has_forced_neighbor, jps3d.py:163-246.  blocked(p) = p in obstacles; free(p) = p in grid_map and
// not in obstacles.
template <class ENV>
BF_HD static inline bool jps3d_forced(const ENV &env, int x, int y, int z, int dx, int dy, int dz) {
#define JB(a, b, c) env.obst((a), (b), (c))
#define JF(a, b, c) (env.inbox((a), (b), (c)) && !env.obst((a), (b), (c)))
    const int nz = (dx != 0) + (dy != 0) + (dz != 0);
    if (nz == 1) {
        if (dx) {
            for (int o = -1; o <= 1; o += 2) if (JB(x, y + o, z) && JF(x + dx, y + o, z)) return true;
            for (int o = -1; o <= 1; o += 2) if (JB(x, y, z + o) && JF(x + dx, y, z + o)) return true;
        } else if (dy) {
            for (int o = -1; o <= 1; o += 2) if (JB(x + o, y, z) && JF(x + o, y + dy, z)) return true;
            for (int o = -1; o <= 1; o += 2) if (JB(x, y, z + o) && JF(x, y + dy, z + o)) return true;
        } else {
            for (int o = -1; o <= 1; o += 2) if (JB(x + o, y, z) && JF(x + o, y, z + dz)) return true;
            for (int o = -1; o <= 1; o += 2) if (JB(x, y + o, z) && JF(x, y + o, z + dz)) return true;
        }
        return false;
    }
    if (nz == 2) {
        if (dx && dy) return (JB(x - dx, y, z) && JF(x - dx, y + dy, z)) || (JB(x, y - dy, z) && JF(x + dx, y - dy, z));
        if (dx && dz) return (JB(x - dx, y, z) && JF(x - dx, y, z + dz)) || (JB(x, y, z - dz) && JF(x + dx, y, z - dz));
        return (JB(x, y - dy, z) && JF(x, y - dy, z + dz)) || (JB(x, y, z - dz) && JF(x, y + dy, z - dz));
    }
    if (nz == 3)
        return (JB(x - dx, y, z) && JF(x - dx, y + dy, z)) || (JB(x, y - dy, z) && JF(x + dx, y - dy, z)) ||
               (JB(x, y, z - dz) && JF(x + dx, y, z - dz));
    return false;
#undef JB
#undef JF
}

// truth value of jump(node at (x, y, z), d) for a 1-axis d: jps3d.py:108-134
template <class ENV>
BF_HD static inline bool jps3d_exists1(const BfState &S, const ENV &env, int x, int y, int z, int dx, int dy, int dz) {
    for (;;) {
        if (!env.legal_dir(dx, dy, dz, x, y, z)) return false;
        x += dx; y += dy; z += dz;
        if (x == S.gx && y == S.gy && z == S.gz) return true;
        if (jps3d_forced(env, x, y, z, dx, dy, dz)) return true;
    }
}

// ... for a 2-axis d: some cell along the legal ray is the goal, has a forced neighbour, or starts a
// successful axis jump (jps3d.py:137-157; the call on the same direction makes it "anything further")
template <class ENV>
BF_HD static inline bool jps3d_exists2(const BfState &S, const ENV &env, int x, int y, int z, int dx, int dy, int dz) {
    for (;;) {
        if (!env.legal_dir(dx, dy, dz, x, y, z)) return false;
        x += dx; y += dy; z += dz;
        if (x == S.gx && y == S.gy && z == S.gz) return true;
        if (jps3d_forced(env, x, y, z, dx, dy, dz)) return true;
        if (dx && jps3d_exists1(S, env, x, y, z, dx, 0, 0)) return true;
        if (dy && jps3d_exists1(S, env, x, y, z, 0, dy, 0)) return true;
        if (dz && jps3d_exists1(S, env, x, y, z, 0, 0, dz)) return true;
    }
}

// JPS3D.jump(node, motion m) + the CLOSED dominance test of plan() (jps3d.py:79-85)
template <class ENV>
BF_HD static inline bool jps3d_eval(const BfState &S, const ENV &env, const BfRec &node, int m, BfRec &out) {
    int x0, y0, z0, dx, dy, dz;
    bf_coords(S, node.cell, x0, y0, z0);
    env.offset(m, dx, dy, dz);
    const int nz = (dx != 0) + (dy != 0) + (dz != 0);
    if (nz == 0) return false;
    int x = x0, y = y0, z = z0;
    if (nz == 1) {
        for (;;) {
            if (!env.legal_dir(dx, dy, dz, x, y, z)) return false;
            x += dx; y += dy; z += dz;
            if (x == S.gx && y == S.gy && z == S.gz) break;
            if (jps3d_forced(env, x, y, z, dx, dy, dz)) break;
        }
    } else if (nz == 2) {
        if (!jps3d_exists2(S, env, x, y, z, dx, dy, dz)) return false;
        x += dx; y += dy; z += dz;                                      // the first step
    } else {
        for (;;) {
            if (!env.legal_dir(dx, dy, dz, x, y, z)) return false;
            x += dx; y += dy; z += dz;
            if (x == S.gx && y == S.gy && z == S.gz) break;
            if (jps3d_forced(env, x, y, z, dx, dy, dz)) break;
            if (jps3d_exists1(S, env, x, y, z, dx, 0, 0) || jps3d_exists1(S, env, x, y, z, 0, dy, 0) ||
                jps3d_exists1(S, env, x, y, z, 0, 0, dz)) break;
            if (jps3d_exists2(S, env, x, y, z, dx, dy, 0) || jps3d_exists2(S, env, x, y, z, dx, 0, dz) ||
                jps3d_exists2(S, env, x, y, z, 0, dy, dz)) break;
        }
    }
    // _mk_qnode (jps3d.py:159-161): g = from_node.g + dist(from_node, at_node)
    const uint32_t nc = bf_cell(S, x, y, z);
    const double g = node.g + bf_h(BF_H_EUCLIDEAN, x - x0, y - y0, z - z0);
    if (S.cpar[nc] != BF_NONE && g >= S.cg[nc]) return false;           // jps3d.py:83-84
    out.cell = nc;
    out.parent = node.cell;
    out.g = g;
    out.h = bf_h(S.heuristic, S.gx - x, S.gy - y, S.gz - z);
    out.f = out.g + out.h;
    out.counter = 0;
    out.pad = 0;
    return true;
}
