// pmp_incr_core.h -- the reference's incremental planners in the reference's own processing order
// (pmp_lpa_star / pmp_d_star in csrc/pmp_bestfirst.cu).
//
//   LPA*   LPAStar      global_planner/graph_search/lpa_star.py:62-230
//          LPAStar3D    lpa_star3d.py:57-225
//   D* Lite DStarLite   d_star_lite.py:34-187 (LPA* searched from the goal, keys shifted by km)
//   D*     DStar        d_star.py:60-291
//          DStar3D      d_star3d.py:68-280
//
// pmp_replan_* (pmp_replan.cuh) repairs the resident cost field in parallel and returns the
// optimum.  What the reference returns differs in ways that follow from its data structures: the
// OPEN "queue" is a Python list searched with min() and edited with list.remove(), so ties are
// broken by list position; LPA* walks back from the goal greedily by g alone; D* follows back
// pointers whose reverse edge may be illegal under the directed corner rule (cost = inf).  This
// header restates those structures literally -- list order, heapq.heappush on a list that is not
// a heap, first-minimum scans -- so cost, path and len(EXPAND) ("Visited Nodes") equal the
// reference's, also where the reference is not optimal.  State lives in caller-owned arrays and
// survives between calls (apply_change / apply_dynamic_obstacles).  Sequential by nature: one
// thread per planner; meant for the reference's own grid sizes.  Free of CUDA intrinsics;
// tests/bestfirst_host.cpp compiles it with g++.
#pragma once

#include "pmp_bestfirst_core.h"

// ---------------------------------------------------------------------------------------------
// LPA*
// ---------------------------------------------------------------------------------------------
struct LpaState {
    double *g, *rhs, *k1, *k2; // map[cell].g / .rhs / .key[0] / .key[1]
    uint32_t *U;               // the list self.U (cells, in list order)
    int64_t un, ucap;
    uint32_t *expand;          // self.EXPAND
    int64_t n_expand, expand_cap;
    BfState S;                 // geometry, heuristic, start / goal (heap / CLOSED members unused)
    int family;                // 2: LPAStar, 3: LPAStar3D
    int status;                // BF_RUNNING while searching; BF_REACHED / BF_EXHAUSTED / BF_OVERFLOW
    int dlite;                 // 1: DStarLite -- rhs = 0 at the goal, keys measured to the start, + km
    double km;
};

// the node whose rhs is pinned to 0 / the node the search runs towards
BF_HD static inline uint32_t lpa_anchor(const LpaState &L) { return L.dlite ? L.S.goal_cell : L.S.start_cell; }
BF_HD static inline uint32_t lpa_target(const LpaState &L) { return L.dlite ? L.S.start_cell : L.S.goal_cell; }

BF_HD static inline bool lpa_key_lt(double a1, double a2, double b1, double b2) { // list < list
    return a1 < b1 || (a1 == b1 && a2 < b2);
}

// calculateKey, lpa_star.py:181-192 / lpa_star3d.py:172-175 / d_star_lite.py:155-166
BF_HD static inline void lpa_calc_key(const LpaState &L, uint32_t c, double &k1, double &k2) {
    int x, y, z, tx, ty, tz;
    bf_coords(L.S, c, x, y, z);
    bf_coords(L.S, lpa_target(L), tx, ty, tz);
    const double m = L.g[c] < L.rhs[c] ? L.g[c] : L.rhs[c]; // min(node.g, node.rhs)
    k1 = m + bf_h(L.S.heuristic, tx - x, ty - y, tz - z);
    if (L.dlite) k1 = k1 + L.km;
    k2 = m;
}

BF_HD static inline int64_t lpa_find(const LpaState &L, uint32_t c) {
    for (int64_t i = 0; i < L.un; ++i)
        if (L.U[i] == c) return i;
    return -1;
}

BF_HD static inline void lpa_remove_at(LpaState &L, int64_t i) { // list.remove: shift left
    for (int64_t j = i; j + 1 < L.un; ++j) L.U[j] = L.U[j + 1];
    --L.un;
}

// heapq.heappush on self.U (which list.remove has long stopped being a heap): append + _siftdown
BF_HD static inline void lpa_heappush(LpaState &L, uint32_t c) {
    if (L.un >= L.ucap) { L.status = BF_OVERFLOW; return; }
    int64_t pos = L.un++;
    L.U[pos] = c;
    while (pos > 0) {
        const int64_t parent = (pos - 1) >> 1;
        const uint32_t pc = L.U[parent];
        if (lpa_key_lt(L.k1[c], L.k2[c], L.k1[pc], L.k2[pc])) { L.U[pos] = pc; pos = parent; continue; }
        break;
    }
    L.U[pos] = c;
}

// updateVertex, lpa_star.py:159-179 / lpa_star3d.py:156-170
template <class ENV>
BF_HD static inline void lpa_update_vertex(LpaState &L, const ENV &env, uint32_t c) {
    if (c != lpa_anchor(L)) {
        int x, y, z;
        bf_coords(L.S, c, x, y, z);
        double best = INFINITY; // the 2D class raises on an empty neighbour list; 3D: inf
        for (int m = 0; m < env.k(); ++m) { // getNeighbor: the endpoint must be in the map and free
            int dx, dy, dz;
            env.offset(m, dx, dy, dz);
            const int nx = x + dx, ny = y + dy, nz = z + dz;
            if (!env.inbox(nx, ny, nz) || env.obst(nx, ny, nz)) continue;
            // n.g + self.cost(n, node): inf when the edge n -> node collides (graph_search.py:46-59)
            const double w = env.legal_dir(-dx, -dy, -dz, nx, ny, nz) ? env.len(m) : INFINITY;
            const double v = L.g[bf_cell(L.S, nx, ny, nz)] + w;
            if (v < best) best = v;
        }
        L.rhs[c] = best;
    }
    const int64_t i = lpa_find(L, c);
    if (i >= 0) lpa_remove_at(L, i);
    if (L.g[c] != L.rhs[c]) {
        lpa_calc_key(L, c, L.k1[c], L.k2[c]);
        lpa_heappush(L, c);
    }
}

// __init__ (lpa_star.py:62-75): the caller has filled g, rhs with +inf
BF_HD static inline void lpa_init(LpaState &L) {
    L.un = 0;
    L.n_expand = 0;
    L.status = BF_RUNNING;
    L.km = 0.0;
    const uint32_t a = lpa_anchor(L);                       // d_star_lite.py:34-52: the goal
    L.rhs[a] = 0.0;
    lpa_calc_key(L, a, L.k1[a], L.k2[a]);
    lpa_heappush(L, a);
}

// the part of apply_change / OnPress after the obstacle set was edited (lpa_star3d.py:104-130,
// lpa_star.py:112-127): EXPAND cleared, the freed cell relaxed, its neighbours updated
template <class ENV>
BF_HD static inline void lpa_apply_change(LpaState &L, const ENV &env, int x, int y, int z, bool freed) {
    L.n_expand = 0;
    L.status = BF_RUNNING;
    if (!env.inbox(x, y, z)) return;
    if (freed) lpa_update_vertex(L, env, bf_cell(L.S, x, y, z));
    for (int m = 0; m < env.k(); ++m) {
        int dx, dy, dz;
        env.offset(m, dx, dy, dz);
        const int nx = x + dx, ny = y + dy, nz = z + dz;
        if (!env.inbox(nx, ny, nz) || env.obst(nx, ny, nz)) continue;
        lpa_update_vertex(L, env, bf_cell(L.S, nx, ny, nz));
    }
}

// computeShortestPath, lpa_star.py:135-157 / lpa_star3d.py:137-154 / d_star_lite.py:104-131
template <class ENV>
BF_HD static inline void lpa_compute(LpaState &L, const ENV &env) {
    const uint32_t goal = lpa_target(L);
    while (L.status == BF_RUNNING) {
        if (L.un == 0) { L.status = BF_EXHAUSTED; return; } // 3D: loop ends; 2D: min([]) raises
        int64_t bi = 0; // min(self.U, key=...): the first minimum in list order
        for (int64_t i = 1; i < L.un; ++i) {
            const uint32_t a = L.U[i], b = L.U[bi];
            if (lpa_key_lt(L.k1[a], L.k2[a], L.k1[b], L.k2[b])) bi = i;
        }
        const uint32_t c = L.U[bi];
        double g1, g2;
        lpa_calc_key(L, goal, g1, g2);
        if (!lpa_key_lt(L.k1[c], L.k2[c], g1, g2) && L.rhs[goal] == L.g[goal]) return;
        lpa_remove_at(L, bi);
        if (L.n_expand >= L.expand_cap) { L.status = BF_OVERFLOW; return; }
        L.expand[L.n_expand++] = c;
        if (L.dlite) { // "affected by obstacles": the stored key is out of date (d_star_lite.py:116-118)
            double n1, n2;
            lpa_calc_key(L, c, n1, n2);
            if (lpa_key_lt(L.k1[c], L.k2[c], n1, n2)) {
                L.k1[c] = n1; L.k2[c] = n2;
                lpa_heappush(L, c);
                continue;
            }
        }
        if (L.g[c] > L.rhs[c]) L.g[c] = L.rhs[c];
        else { L.g[c] = INFINITY; lpa_update_vertex(L, env, c); }
        int x, y, z;
        bf_coords(L.S, c, x, y, z);
        for (int m = 0; m < env.k(); ++m) {
            int dx, dy, dz;
            env.offset(m, dx, dy, dz);
            const int nx = x + dx, ny = y + dy, nz = z + dz;
            if (!env.inbox(nx, ny, nz) || env.obst(nx, ny, nz)) continue;
            lpa_update_vertex(L, env, bf_cell(L.S, nx, ny, nz));
        }
    }
}

// extractPath, lpa_star.py:209-230 / lpa_star3d.py:196-225: greedy descent on g from the goal.
// Cells are written goal first; *plen = 0: no path.
template <class ENV>
BF_HD static inline void lpa_extract(const LpaState &L, const ENV &env, uint32_t *path, int64_t cap,
                                     int32_t *plen, double *cost) {
    uint32_t c = L.S.goal_cell;
    int64_t n = 0, count = 0;
    double total = 0.0;
    if (n < cap) path[n] = c;
    ++n;
    while (c != L.S.start_cell) {
        int x, y, z;
        bf_coords(L.S, c, x, y, z);
        bool any = false;
        double bg = 0.0, bw = 0.0;
        uint32_t bc = 0;
        for (int m = 0; m < env.k(); ++m) {
            int dx, dy, dz;
            env.offset(m, dx, dy, dz);
            const int nx = x + dx, ny = y + dy, nz = z + dz;
            if (!env.inbox(nx, ny, nz) || env.obst(nx, ny, nz)) continue;
            if (!env.legal_dir(dx, dy, dz, x, y, z)) continue;   // not isCollision(node, n)
            const uint32_t nc = bf_cell(L.S, nx, ny, nz);
            if (!any || L.g[nc] < bg) { any = true; bg = L.g[nc]; bc = nc; bw = env.len(m); }
        }
        if (!any) { *plen = 0; *cost = total; return; }          // 3D: "disconnected"; 2D: min([]) raises
        total += bw;
        c = bc;
        if (L.family == 2) { // the cell is appended before the count check (lpa_star.py:222-228)
            if (n < cap) path[n] = c;
            ++n;
            if (++count == 1000) { *plen = 0; *cost = total; return; }
        } else {
            if (n < cap) path[n] = c;
            ++n;
            if (++count >= 100000) { *plen = 0; *cost = total; return; }
        }
    }
    *plen = n <= cap ? (int32_t)n : -1;
    *cost = total;
}

// one greedy step of the extraction walks: the free, collision-free neighbour with the smallest g
// (first in motion order on ties); false when there is none
template <class ENV>
BF_HD static inline bool lpa_greedy_step(const LpaState &L, const ENV &env, uint32_t c, uint32_t &next, double &w) {
    int x, y, z;
    bf_coords(L.S, c, x, y, z);
    bool any = false;
    double bg = 0.0;
    for (int m = 0; m < env.k(); ++m) {
        int dx, dy, dz;
        env.offset(m, dx, dy, dz);
        const int nx = x + dx, ny = y + dy, nz = z + dz;
        if (!env.inbox(nx, ny, nz) || env.obst(nx, ny, nz)) continue;
        if (!env.legal_dir(dx, dy, dz, x, y, z)) continue;
        const uint32_t nc = bf_cell(L.S, nx, ny, nz);
        if (!any || L.g[nc] < bg) { any = true; bg = L.g[nc]; next = nc; w = env.len(m); }
    }
    return any;
}

// DStarLite.extractPath, d_star_lite.py:168-187: greedy descent from the start; cells start first
template <class ENV>
BF_HD static inline void dlite_extract(const LpaState &L, const ENV &env, uint32_t *path, int64_t cap,
                                       int32_t *plen, double *cost) {
    uint32_t c = L.S.start_cell, nc = 0;
    int64_t n = 0, count = 0;
    double total = 0.0, w = 0.0;
    if (n < cap) path[n] = c;
    ++n;
    while (c != L.S.goal_cell) {
        if (!lpa_greedy_step(L, env, c, nc, w)) { *plen = 0; *cost = total; return; }
        if (n < cap) path[n] = nc;
        ++n;
        total += w;
        c = nc;
        if (++count == 1000) { *plen = 0; *cost = total; return; }
    }
    *plen = n <= cap ? (int32_t)n : -1;
    *cost = total;
}

// DStarLite.OnPress, d_star_lite.py:55-102: walk greedily from the start; after the first step the
// clicked cell is toggled (env.toggle returns true when an obstacle was removed), km takes the
// heuristic between the new and the old position, the vertices around the change are updated and
// the field repaired; the walk then continues on the repaired field.  Cells start first.
template <class ENV>
BF_HD static inline void dlite_press(LpaState &L, ENV &env, int px, int py, int pz, uint32_t *path, int64_t cap,
                                     int32_t *plen, double *cost) {
    L.n_expand = 0;
    L.status = BF_RUNNING;
    uint32_t c = L.S.start_cell, nc = 0;
    int64_t n = 0, count = 0;
    double total = 0.0, w = 0.0;
    bool update_start = true;
    if (n < cap) path[n] = c;
    ++n;
    while (c != L.S.goal_cell) {
        if (!lpa_greedy_step(L, env, c, nc, w)) { *plen = 0; *cost = total; return; } // min([]) raises in the reference
        if (n < cap) path[n] = nc;
        ++n;
        total += w;
        c = nc;
        if (++count >= 100000) { *plen = 0; *cost = total; return; }                  // the reference has no guard
        if (update_start) {
            update_start = false;
            int x, y, z, sx, sy, sz;
            bf_coords(L.S, c, x, y, z);
            bf_coords(L.S, L.S.start_cell, sx, sy, sz);
            L.km = bf_h(L.S.heuristic, sx - x, sy - y, sz - z);
            const bool freed = env.toggle(px, py, pz);
            const int ne = (int)L.n_expand;
            lpa_apply_change(L, env, px, py, pz, freed);
            L.n_expand = ne;                               // EXPAND was cleared once, at the top
            lpa_compute(L, env);
            if (L.status != BF_RUNNING) { *plen = 0; *cost = total; return; }
        }
    }
    *plen = n <= cap ? (int32_t)n : -1;
    *cost = total;
}

// ---------------------------------------------------------------------------------------------
// D*
// ---------------------------------------------------------------------------------------------
#define DS_NEW 0
#define DS_OPEN 1
#define DS_CLOSED 2
#define DS_GUARD 4000000 // processState calls per entry point (the 2D plan() spins forever on an unreachable start)

struct DsState {
    double *h, *k;     // map[cell].h / .k
    uint32_t *parent;  // map[cell].parent (BF_NONE = None)
    uint8_t *t;        // DS_NEW / DS_OPEN / DS_CLOSED
    uint32_t *OPEN;    // the list self.OPEN (2D: may hold a cell more than once, d_star.py:242)
    int64_t on, ocap;
    uint32_t *expand;  // self.EXPAND
    int64_t n_expand, expand_cap;
    BfState S;
    int family;
    int status;
};

// insert, d_star.py:228-242 / d_star3d.py:236-249
BF_HD static inline void ds_insert(DsState &D, uint32_t c, double h_new) {
    const bool was_open = D.t[c] == DS_OPEN;
    if (D.t[c] == DS_NEW) D.k[c] = h_new;
    else if (was_open) D.k[c] = D.k[c] < h_new ? D.k[c] : h_new;
    else D.k[c] = D.h[c] < h_new ? D.h[c] : h_new;
    D.h[c] = h_new;
    D.t[c] = DS_OPEN;
    // 3D: "if node not in self.OPEN" (d_star3d.py:248-249).  There a node is in the list exactly while
    // its tag is OPEN (insert appends on the way in, delete removes on the way out), so the tag
    // answers the membership test without scanning the list.
    if (D.family == 3 && was_open) return;
    if (D.on >= D.ocap) { D.status = BF_OVERFLOW; return; }
    D.OPEN[D.on++] = c;
}

// min_state: the first minimum of k in list order; -1 when OPEN is empty
BF_HD static inline int64_t ds_min_index(const DsState &D) {
    if (D.on == 0) return -1;
    int64_t bi = 0;
    for (int64_t i = 1; i < D.on; ++i)
        if (D.k[D.OPEN[i]] < D.k[D.OPEN[bi]]) bi = i;
    return bi;
}

// processState, d_star.py:150-211 / d_star3d.py:167-220.  Returns the new minimum k, -1: OPEN empty
// (want_min = false: only the sign is wanted -- plan() -- and the second scan of the list is skipped).
template <class ENV>
BF_HD static inline double ds_process_state(DsState &D, const ENV &env, bool want_min = true) {
    const int64_t bi = ds_min_index(D);
    if (bi < 0) return -1.0;
    const uint32_t c = D.OPEN[bi];
    if (D.n_expand >= D.expand_cap) { D.status = BF_OVERFLOW; return -1.0; }
    D.expand[D.n_expand++] = c;
    const double k_old = D.k[c];
    if (D.t[c] == DS_OPEN) D.t[c] = DS_CLOSED;          // delete(): the first occurrence leaves the list --
    {                                                     // that is bi: duplicates of a node share its k, and
        for (int64_t j = bi; j + 1 < D.on; ++j) D.OPEN[j] = D.OPEN[j + 1]; // min() returned the first of them
        --D.on;
    }
    int x, y, z;
    bf_coords(D.S, c, x, y, z);
    // getNeighbor: in the map and not isCollision(node, n) (d_star.py:276-291, d_star3d.py:265-280)
    uint32_t nb[32];
    double nw[32];
    int nn = 0;
    for (int m = 0; m < env.k(); ++m) {
        int dx, dy, dz;
        env.offset(m, dx, dy, dz);
        if (!env.inbox(x + dx, y + dy, z + dz) || !env.legal_dir(dx, dy, dz, x, y, z)) continue;
        nb[nn] = bf_cell(D.S, x + dx, y + dy, z + dz);
        nw[nn] = env.len(m);                              // self.cost(node, node_n): the edge is free
        ++nn;
    }
    if (k_old < D.h[c]) {                                 // RAISE
        for (int i = 0; i < nn; ++i) {
            const uint32_t n = nb[i];
            if (D.h[n] <= k_old && D.h[c] > D.h[n] + nw[i]) { D.parent[c] = n; D.h[c] = D.h[n] + nw[i]; }
        }
    }
    if (k_old == D.h[c]) {                                // LOWER
        for (int i = 0; i < nn; ++i) {
            const uint32_t n = nb[i];
            const double v = D.h[c] + nw[i];
            if (D.t[n] == DS_NEW || (D.parent[n] == c && D.h[n] != v) || (D.parent[n] != c && D.h[n] > v)) {
                D.parent[n] = c;
                ds_insert(D, n, v);
            }
        }
    } else {
        for (int i = 0; i < nn; ++i) {
            const uint32_t n = nb[i];
            const double v = D.h[c] + nw[i];
            if (D.t[n] == DS_NEW || (D.parent[n] == c && D.h[n] != v)) {
                D.parent[n] = c;
                ds_insert(D, n, v);
            } else if (D.parent[n] != c && D.h[n] > v) {
                ds_insert(D, c, D.h[c]);
            } else if (D.parent[n] != c && D.h[c] > D.h[n] + nw[i] && D.t[n] == DS_CLOSED && D.h[n] > k_old) {
                ds_insert(D, n, D.h[n]);
            }
        }
    }
    if (!want_min) return D.on == 0 ? -1.0 : 0.0;
    const int64_t mi = ds_min_index(D);
    return mi < 0 ? -1.0 : D.k[D.OPEN[mi]];
}

// __init__ (d_star.py:52-70, d_star3d.py:68-95): the caller has filled h, k with +inf, parent with
// BF_NONE, t with DS_NEW
BF_HD static inline void ds_init(DsState &D) {
    D.on = 0;
    D.n_expand = 0;
    D.status = BF_RUNNING;
    D.h[D.S.goal_cell] = 0.0;
    ds_insert(D, D.S.goal_cell, 0.0);
}

// plan(), d_star.py:75-88 / d_star3d.py:102-111 (without extractPath)
template <class ENV>
BF_HD static inline void ds_plan(DsState &D, const ENV &env) {
    for (int64_t it = 0; it < DS_GUARD && D.status == BF_RUNNING; ++it) {
        const double kmin = ds_process_state(D, env, false);
        if (D.family == 3 && kmin < 0) break;
        if (D.t[D.S.start_cell] == DS_CLOSED) break;
        if (D.family == 2 && D.on == 0) { D.status = BF_EXHAUSTED; break; } // the reference would spin here
    }
}

// self.cost(a, b): inf when the edge a -> b collides (graph_search.py:46-59)
template <class ENV>
BF_HD static inline double ds_cost(const DsState &D, const ENV &env, uint32_t a, uint32_t b) {
    int ax, ay, az, bx, by, bz;
    bf_coords(D.S, a, ax, ay, az);
    bf_coords(D.S, b, bx, by, bz);
    const int dx = bx - ax, dy = by - ay, dz = bz - az;
    const bool unit = dx >= -1 && dx <= 1 && dy >= -1 && dy <= 1 && dz >= -1 && dz <= 1;
    if (unit && !env.legal_dir(dx, dy, dz, ax, ay, az)) return INFINITY;
    return bf_h(BF_H_EUCLIDEAN, dx, dy, dz);
}

// extractPath, d_star.py:136-148 / d_star3d.py:153-165: back pointers from the start; cells are
// written start first.  The 3D class stops at a node without a parent and returns what it has.
template <class ENV>
BF_HD static inline void ds_extract(const DsState &D, const ENV &env, uint32_t *path, int64_t cap,
                                    int32_t *plen, double *cost) {
    double total = 0.0;
    int64_t n = 0;
    uint32_t c = D.S.start_cell;
    if (n < cap) path[n] = c;
    ++n;
    while (c != D.S.goal_cell) {
        const uint32_t p = D.parent[c];
        if (p == BF_NONE || n > cap) break;               // 2D: KeyError in the reference
        total += ds_cost(D, env, c, p);
        c = p;
        if (n < cap) path[n] = c;
        ++n;
    }
    *plen = n <= cap ? (int32_t)n : -1;
    *cost = total;
}

// modify, d_star.py:257-269 / d_star3d.py:252-262
template <class ENV>
BF_HD static inline void ds_modify(DsState &D, const ENV &env, uint32_t c, uint32_t p) {
    if (D.t[c] == DS_CLOSED) ds_insert(D, c, D.h[p] + ds_cost(D, env, c, p));
    for (int64_t it = 0; it < DS_GUARD && D.status == BF_RUNNING; ++it) {
        const double kmin = ds_process_state(D, env);
        if (kmin < 0 || kmin >= D.h[c]) break;            // 2D has no "< 0" exit: it would spin
    }
}

// The walk of OnPress (d_star.py:113-127) / apply_dynamic_obstacles (d_star3d.py:127-151) after the
// obstacle set was edited: follow the back pointers from the start, repair where an edge collides.
// Cells are written start first; the 2D walk does not append the goal.
template <class ENV>
BF_HD static inline void ds_replan(DsState &D, const ENV &env, uint32_t *path, int64_t cap, int32_t *plen,
                                   double *cost) {
    D.n_expand = 0;
    D.status = BF_RUNNING;
    double total = 0.0;
    int64_t n = 0, guard = 0;
    uint32_t c = D.S.start_cell;
    while (c != D.S.goal_cell && D.status == BF_RUNNING && ++guard < DS_GUARD) {
        const uint32_t p = D.parent[c];
        if (p == BF_NONE) {
            if (D.family == 3) ds_modify(D, env, c, D.S.goal_cell);
            break;
        }
        if (ds_cost(D, env, c, p) == INFINITY) { ds_modify(D, env, c, p); continue; } // isCollision(node, parent)
        if (n < cap) path[n] = c;
        ++n;
        total += ds_cost(D, env, c, p);
        c = p;
    }
    if (c == D.S.goal_cell && D.family == 3) {
        if (n < cap) path[n] = c;
        ++n;
    }
    *plen = n <= cap ? (int32_t)n : -1;
    *cost = total;
}
