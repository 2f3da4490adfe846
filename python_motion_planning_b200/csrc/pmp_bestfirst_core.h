// pmp_bestfirst_core.h -- the sequential part of the reference-order best-first searches
// (pmp_bestfirst, csrc/pmp_bestfirst.cu): OPEN heap, CLOSED table, neighbour evaluation, path walk.
//
// The reference planners are defined by their pop order: which of several equal-cost paths is
// returned, the content and order of `expand` and "Visited Nodes" (examples/benchmark.py:85) all
// follow from the heap discipline.  This header restates that discipline once, free of CUDA
// intrinsics, for the five planners that share it (paths relative to
// /root/reference/src/python_motion_planning/global_planner/graph_search/):
//
//   A*          a_star.py:39-117          a_star3d.py:33-106
//   Dijkstra    dijkstra.py:36-85         dijkstra3d.py:39-143
//   GBFS        gbfs.py:36-86             gbfs3d.py:34-110
//   Theta*      theta_star.py:44-108      theta_star3d.py:38-107
//   Lazy Theta* lazy_theta_star.py:44-118 lazy_theta_star3d.py:41-133
//
// FAMILY 2 = the 2D classes: OPEN holds Node objects ordered by Node.__lt__
// (utils/environment/node.py:51-54: f, then h -- not a total order, so CPython heapq's sift
// order decides), a node is closed once, the neighbour loop stops when it meets the goal.
// FAMILY 3 = the 3D classes: OPEN holds (f, h, counter, node) tuples, a cell is re-closed when a
// strictly better g is popped.
//
// The kernel runs this code on lane 0 of a warp (bf_pop / bf_push / bf_close) and bf_eval on one
// lane per motion; tests/bestfirst_host.cpp compiles the same header with g++ to check the logic
// on the CPU against outputs of the reference (test infrastructure, not a product path).
#pragma once

#include <math.h>
#include <stdint.h>

#ifdef __CUDACC__
#define BF_HD __host__ __device__
#else
#define BF_HD
#endif

// algo ids: the first three equal pmp_algo in include/pmp_b200.h
#define BF_ASTAR 0
#define BF_DIJKSTRA 1
#define BF_GBFS 2
#define BF_THETA 3
#define BF_LAZY_THETA 4

#define BF_H_EUCLIDEAN 0
#define BF_H_MANHATTAN 1

#define BF_NONE 0xFFFFFFFFu

// bf_pop results
#define BF_SKIP 0
#define BF_EXPAND 1
#define BF_DONE 2

// status
#define BF_RUNNING 0
#define BF_REACHED 1
#define BF_EXHAUSTED 2
#define BF_OVERFLOW (-2)

struct BfRec { // one OPEN entry; 40 bytes (PMP_BESTFIRST_HEAP_REC_BYTES)
    double f, h, g;
    uint32_t cell, parent; // padded cell indices
    uint32_t counter;      // FAMILY 3: itertools.count() value of the push
    uint32_t pad;
};

struct BfState {
    BfRec *heap;
    int64_t hn, heap_cap;
    double *cg, *ch;   // CLOSED[cell].g / .h
    uint32_t *cpar;    // CLOSED[cell].parent; BF_NONE = cell not in CLOSED
    uint32_t *order;   // keys of CLOSED in insertion order (dict order)
    int64_t n_order, order_cap;
    uint32_t counter;
    int algo, heuristic;
    int PX, PY;        // x and y extents of the padded cell index
    int gx, gy, gz;
    uint32_t start_cell, goal_cell;
    int status;
};

BF_HD static inline void bf_coords(const BfState &S, uint32_t cell, int &x, int &y, int &z) {
    x = (int)(cell % (uint32_t)S.PX);
    y = (int)((cell / (uint32_t)S.PX) % (uint32_t)S.PY);
    z = (int)(cell / ((uint32_t)S.PX * (uint32_t)S.PY));
}
BF_HD static inline uint32_t bf_cell(const BfState &S, int x, int y, int z) {
    return (uint32_t)(((int64_t)z * S.PY + y) * S.PX + x);
}

// GraphSearcher.h (graph_search.py:30-44: math.hypot / manhattan), GraphSearcher3D.h
// (graph_search_3d.py:30-49: math.sqrt(dx**2 + dy**2 + dz**2) / manhattan).  On integer offsets
// both equal the correctly rounded sqrt of the exact integer sum (checked exhaustively for offsets
// below 2100 against CPython 3.12 in the build container).
BF_HD static inline double bf_h(int heuristic, int dx, int dy, int dz) {
    dx = dx < 0 ? -dx : dx; dy = dy < 0 ? -dy : dy; dz = dz < 0 ? -dz : dz;
    if (heuristic == BF_H_MANHATTAN) return (double)(dx + dy + dz);
    return sqrt((double)((long long)dx * dx + (long long)dy * dy + (long long)dz * dz));
}
// Planner.dist (utils/planner/planner.py:22-23), Planner3D.dist (planner3d.py:22-27)
BF_HD static inline double bf_dist(const BfState &S, uint32_t a, uint32_t b) {
    int ax, ay, az, bx, by, bz;
    bf_coords(S, a, ax, ay, az);
    bf_coords(S, b, bx, by, bz);
    return bf_h(BF_H_EUCLIDEAN, ax - bx, ay - by, az - bz);
}

template <int FAMILY>
BF_HD static inline bool bf_lt(const BfRec &a, const BfRec &b) {
    if (FAMILY == 2) return a.f < b.f || (a.f == b.f && a.h < b.h); // node.py:51-54
    if (a.f != b.f) return a.f < b.f;                               // tuple order, a_star3d.py:40
    if (a.h != b.h) return a.h < b.h;
    return a.counter < b.counter;
}

template <int FAMILY>
BF_HD static inline void bf_sift_down(BfRec *a, int64_t startpos, int64_t pos) { // heapq._siftdown
    const BfRec item = a[pos];
    while (pos > startpos) {
        const int64_t parent = (pos - 1) >> 1;
        const BfRec pr = a[parent];
        if (bf_lt<FAMILY>(item, pr)) { a[pos] = pr; pos = parent; continue; }
        break;
    }
    a[pos] = item;
}

template <int FAMILY>
BF_HD static inline void bf_sift_up(BfRec *a, int64_t n, int64_t pos) { // heapq._siftup
    const int64_t startpos = pos;
    const BfRec item = a[pos];
    int64_t child = 2 * pos + 1;
    while (child < n) {
        const int64_t right = child + 1;
        BfRec cr = a[child];
        if (right < n) {
            const BfRec rr = a[right];
            if (!bf_lt<FAMILY>(cr, rr)) { child = right; cr = rr; }
        }
        a[pos] = cr;
        pos = child;
        child = 2 * pos + 1;
    }
    a[pos] = item;
    bf_sift_down<FAMILY>(a, startpos, pos);
}

// heapq.heappush; FAMILY 3 stamps the tuple's counter.  Returns true when the FAMILY 2 neighbour
// loop has to stop because the goal was pushed (a_star.py:75-77, theta_star.py:83-85).
template <int FAMILY>
BF_HD static inline bool bf_push(BfState &S, BfRec r) {
    if (S.hn >= S.heap_cap) { S.status = BF_OVERFLOW; return true; }
    r.counter = S.counter++;
    r.pad = 0;
    S.heap[S.hn] = r;
    ++S.hn;
    bf_sift_down<FAMILY>(S.heap, 0, S.hn - 1);
    return FAMILY == 2 && r.cell == S.goal_cell;
}

// CLOSED[node.current] = node: a new key is appended, an existing key keeps its place and takes
// the new value (dict semantics; a_star.py:82, a_star3d.py:52)
BF_HD static inline void bf_close(BfState &S, const BfRec &node) {
    if (S.cpar[node.cell] == BF_NONE) {
        if (S.n_order >= S.order_cap) { S.status = BF_OVERFLOW; return; }
        S.order[S.n_order++] = node.cell;
    }
    S.cg[node.cell] = node.g;
    S.ch[node.cell] = node.h;
    S.cpar[node.cell] = node.parent;
}

template <int FAMILY>
BF_HD static inline void bf_start(BfState &S, int sx, int sy, int sz) {
    BfRec s;
    s.cell = S.start_cell;
    s.parent = S.start_cell;
    s.g = 0.0;
    if (FAMILY == 2) { // Planner.__init__, planner.py:15: Node(start, start, 0, 0)
        s.h = 0.0;
        s.f = 0.0;
    } else {           // a_star3d.py:35-40, dijkstra3d.py:33-34,55, gbfs3d.py:28-29,46
        s.h = S.algo == BF_DIJKSTRA ? 0.0 : bf_h(S.heuristic, S.gx - sx, S.gy - sy, S.gz - sz);
        s.f = S.algo == BF_GBFS ? s.h : s.g + s.h;
    }
    S.hn = 0;
    S.counter = 0;
    S.n_order = 0;
    S.status = BF_RUNNING;
    bf_push<FAMILY>(S, s);
}

// One heappop and everything up to the neighbour loop.  ENV supplies
//   bool legal(int m, int x, int y, int z)   edge (x,y,z) -> (x,y,z) + motion m exists
//   bool los(ax, ay, az, bx, by, bz)         lineOfSight(a, b) of the family
//   int  k()                                 number of motions
//   void offset(m, dx, dy, dz), double len(m)
template <int FAMILY, class ENV>
BF_HD static inline int bf_pop(BfState &S, const ENV &env, BfRec &node) {
    if (S.hn == 0) { S.status = BF_EXHAUSTED; return BF_DONE; }
    { // heapq.heappop
        const BfRec last = S.heap[--S.hn];
        if (S.hn > 0) { node = S.heap[0]; S.heap[0] = last; bf_sift_up<FAMILY>(S.heap, S.hn, 0); }
        else node = last;
    }
    if (S.algo == BF_LAZY_THETA) {
        // "set vertex" (lazy_theta_star.py:60-70, lazy_theta_star3d.py:60-71): the optimistic
        // parent is checked now; without line of sight the best CLOSED neighbour takes over
        const uint32_t p = node.parent;
        if (S.cpar[p] != BF_NONE) {
            int px, py, pz, x, y, z;
            bf_coords(S, p, px, py, pz);
            bf_coords(S, node.cell, x, y, z);
            if (!env.los(px, py, pz, x, y, z)) {
                node.g = INFINITY;
                for (int m = 0; m < env.k(); ++m) {
                    if (!env.legal(m, x, y, z)) continue;
                    int dx, dy, dz;
                    env.offset(m, dx, dy, dz);
                    const uint32_t nb = bf_cell(S, x + dx, y + dy, z + dz);
                    if (S.cpar[nb] == BF_NONE) continue;
                    const double cand = S.cg[nb] + env.len(m);
                    if (node.g > cand) { node.g = cand; node.parent = nb; }
                }
            }
        }
    }
    if (FAMILY == 2) {
        if (S.cpar[node.cell] != BF_NONE) return BF_SKIP;             // a_star.py:56-58
        if (node.cell == S.goal_cell) {                               // a_star.py:60-64
            bf_close(S, node);
            if (S.status == BF_RUNNING) S.status = BF_REACHED;
            return BF_DONE;
        }
        return BF_EXPAND;
    }
    if (S.algo == BF_GBFS) {
        if (S.cpar[node.cell] != BF_NONE) return BF_SKIP;             // gbfs3d.py:55-56
    } else if (S.cpar[node.cell] != BF_NONE && node.g >= S.cg[node.cell]) {
        return BF_SKIP;                                               // a_star3d.py:47-50
    }
    bf_close(S, node);                                                // a_star3d.py:52
    if (S.status != BF_RUNNING) return BF_DONE;
    if (node.cell == S.goal_cell) { S.status = BF_REACHED; return BF_DONE; } // a_star3d.py:60-63
    return BF_EXPAND;
}

// Neighbour m of the popped node: false when the reference does not push it.  `w` is what one
// step adds to g: motion.g for FAMILY 2 (Node.__add__, node.py:41), dist(node, nbr) for FAMILY 3
// (a_star3d.py:67).  Reads CLOSED only.
template <int FAMILY, class ENV>
BF_HD static inline bool bf_eval(const BfState &S, const ENV &env, const BfRec &node, int m, double w,
                                 BfRec &out) {
    int x, y, z, dx, dy, dz;
    bf_coords(S, node.cell, x, y, z);
    if (!env.legal(m, x, y, z)) return false;   // getNeighbor: a_star.py:85-96, a_star3d.py:80-89,
    env.offset(m, dx, dy, dz);                  // dijkstra3d.py:89-126, theta_star3d.py:109-133
    const int nx = x + dx, ny = y + dy, nz = z + dz;
    const uint32_t nc = bf_cell(S, nx, ny, nz);
    const bool closed = S.cpar[nc] != BF_NONE;
    out.cell = nc;
    out.parent = node.cell;
    out.counter = 0;
    out.pad = 0;
    if (S.algo == BF_GBFS) {
        if (closed) return false;                                     // gbfs.py:70-71, gbfs3d.py:74-75
        out.g = 0.0;                                                  // gbfs.py:75
        out.h = bf_h(S.heuristic, S.gx - nx, S.gy - ny, S.gz - nz);
        out.f = out.h;
        return true;
    }
    const double g1 = node.g + w;
    if (FAMILY == 2) {
        if (closed) return false;                                     // a_star.py:68-69
    } else if (closed && g1 >= S.cg[nc]) {
        return false;                                                 // a_star3d.py:68-70
    }
    out.g = g1;
    out.h = S.algo == BF_DIJKSTRA ? 0.0 : bf_h(S.heuristic, S.gx - nx, S.gy - ny, S.gz - nz);
    if (S.algo == BF_THETA || S.algo == BF_LAZY_THETA) {
        // path 2 through the parent of the popped node, if that parent is in CLOSED
        // (theta_star.py:76-79,96-108; theta_star3d.py:85-88,101-107; the lazy variants skip
        // the line-of-sight test here: lazy_theta_star.py:110-118, lazy_theta_star3d.py:124-133)
        const uint32_t p = node.parent;
        if (S.cpar[p] != BF_NONE) {
            bool visible = true;
            if (S.algo == BF_THETA) {
                int px, py, pz;
                bf_coords(S, p, px, py, pz);
                visible = env.los(nx, ny, nz, px, py, pz);            // lineOfSight(node_c, node_p)
            }
            if (visible) {
                const double alt = S.cg[p] + bf_dist(S, nc, p);
                if (FAMILY == 2 ? alt <= out.g : alt < out.g) { out.g = alt; out.parent = p; }
            }
        }
    }
    out.f = out.g + out.h;
    return true;
}

// extractPath (a_star.py:98-117, a_star3d.py:91-106): CLOSED parents from the goal back to the
// start, cost += dist(node, parent).  Cells are written goal first; *plen = -1 when cap is too small.
BF_HD static inline void bf_extract(const BfState &S, uint32_t *path, int64_t cap, int32_t *plen,
                                    double *cost) {
    int64_t n = 0;
    double c = 0.0;
    uint32_t cell = S.goal_cell;
    for (;;) {
        if (n < cap) path[n] = cell;
        ++n;
        if (cell == S.start_cell) break;
        if (n > S.n_order) { *plen = -1; *cost = c; return; } // a walk cannot exceed len(CLOSED)
        const uint32_t p = S.cpar[cell];
        c += bf_dist(S, cell, p);
        cell = p;
    }
    *plen = n <= cap ? (int32_t)n : -1;
    *cost = c;
}
