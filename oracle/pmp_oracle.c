/*
 * pmp_oracle.c -- CPU restatement (plain C) of the reference's A-star / Dijkstra / GBFS on Grid and
 * Grid3D.  TEST INFRASTRUCTURE (parity oracle + timed CPU baseline); see pmp_oracle.h.
 *
 * Parity status: PINNED against the reference's 3d_pathfinding_results.csv and against the
 * unmodified reference run in the build container (tests/test_oracle_golden.py).
 *
 * Reference paths below are relative to /root/reference/src/python_motion_planning/.
 *
 * What is restated, and where it comes from:
 *   - binary heap with the exact sift order of CPython's heapq (the 2D planners keep Node
 *     objects in a heapq ordered by Node.__lt__, utils/environment/node.py:51-54, which is not
 *     a total order, so pop order depends on heapq's sift algorithm; the 3D planners keep
 *     (f, h, counter, node) tuples, a_star3d.py:40, which is a total order);
 *   - the expansion loops: global_planner/graph_search/a_star.py:39-83, dijkstra.py:36-85,
 *     gbfs.py:36-86, a_star3d.py:33-78, dijkstra3d.py:39-87, gbfs3d.py:34-82;
 *   - neighbour generation: a_star.py:85-96, a_star3d.py:80-89, dijkstra3d.py:89-126,
 *     gbfs3d.py:84-93;
 *   - edge legality: graph_search.py:61-87, graph_search_3d.py:66-107;
 *   - heuristics: graph_search.py:30-44, graph_search_3d.py:30-49;
 *   - path extraction and the reported cost: a_star.py:98-117, a_star3d.py:91-106.
 *
 * Floating point: every sum is an IEEE double evaluated in the reference's order.
 * math.hypot(dx, dy) on integer arguments equals sqrt((double)(dx*dx+dy*dy)) bit for bit for
 * all 0 <= dx,dy < 2100 (checked exhaustively in the build container against CPython 3.12),
 * and math.sqrt is the correctly rounded IEEE sqrt, as is C's.
 */
#include "pmp_oracle.h"

#include <math.h>
#include <stdlib.h>
#include <string.h>
#include <pthread.h>
#include <unistd.h>

typedef struct {
    double f, h, g;
    int64_t cell, parent;
    uint64_t counter;
} rec_t;

typedef struct {
    rec_t *a;
    int64_t n, cap;
    int total_order; /* 0: Node.__lt__ (2D family); 1: tuple (f, h, counter) (3D family) */
} heap_t;

/* Node.__lt__, utils/environment/node.py:51-54: f = g + h, ties broken by smaller h.
 * Tuple comparison for the 3D heaps (a_star3d.py:40,75; dijkstra3d.py:55,84; gbfs3d.py:46,79). */
static inline int rec_lt(const heap_t *hp, const rec_t *x, const rec_t *y)
{
    if (!hp->total_order)
        return x->f < y->f || (x->f == y->f && x->h < y->h);
    if (x->f != y->f) return x->f < y->f;
    if (x->h != y->h) return x->h < y->h;
    return x->counter < y->counter;
}

/* CPython Lib/heapq.py _siftdown */
static void sift_down(heap_t *hp, int64_t startpos, int64_t pos)
{
    rec_t newitem = hp->a[pos];
    while (pos > startpos) {
        int64_t parentpos = (pos - 1) >> 1;
        if (rec_lt(hp, &newitem, &hp->a[parentpos])) {
            hp->a[pos] = hp->a[parentpos];
            pos = parentpos;
            continue;
        }
        break;
    }
    hp->a[pos] = newitem;
}

/* CPython Lib/heapq.py _siftup */
static void sift_up(heap_t *hp, int64_t pos)
{
    int64_t endpos = hp->n, startpos = pos;
    rec_t newitem = hp->a[pos];
    int64_t childpos = 2 * pos + 1;
    while (childpos < endpos) {
        int64_t rightpos = childpos + 1;
        if (rightpos < endpos && !rec_lt(hp, &hp->a[childpos], &hp->a[rightpos]))
            childpos = rightpos;
        hp->a[pos] = hp->a[childpos];
        pos = childpos;
        childpos = 2 * pos + 1;
    }
    hp->a[pos] = newitem;
    sift_down(hp, startpos, pos);
}

static int heap_push(heap_t *hp, const rec_t *r)
{
    if (hp->n == hp->cap) {
        int64_t ncap = hp->cap ? hp->cap * 2 : 1024;
        rec_t *na = (rec_t *)realloc(hp->a, (size_t)ncap * sizeof(rec_t));
        if (!na) return -1;
        hp->a = na;
        hp->cap = ncap;
    }
    hp->a[hp->n++] = *r;
    sift_down(hp, 0, hp->n - 1);
    return 0;
}

static rec_t heap_pop(heap_t *hp)
{
    rec_t last = hp->a[--hp->n];
    if (hp->n > 0) {
        rec_t ret = hp->a[0];
        hp->a[0] = last;
        sift_up(hp, 0);
        return ret;
    }
    return last;
}

typedef struct {
    int ndim;
    int64_t X, Y, Z;
    const uint8_t *occ;
    int escaped;
} grid_t;

/* membership test of `self.obstacles` (graph_search.py:72; graph_search_3d.py:81) on the
 * dense box; outside the box counts as blocked and is recorded (see header). */
static inline int occ_at(grid_t *g, int64_t x, int64_t y, int64_t z)
{
    if (x < 0 || y < 0 || z < 0 || x >= g->X || y >= g->Y || z >= g->Z) {
        g->escaped = 1;
        return 1;
    }
    return g->occ[(z * g->Y + y) * g->X + x] != 0;
}

static inline int in_box(const grid_t *g, int64_t x, int64_t y, int64_t z)
{
    return x >= 0 && y >= 0 && z >= 0 && x < g->X && y < g->Y && z < g->Z;
}

static inline int64_t imin(int64_t a, int64_t b) { return a < b ? a : b; }
static inline int64_t imax(int64_t a, int64_t b) { return a > b ? a : b; }

/* GraphSearcher.isCollision, graph_search.py:61-87 */
static int collision2d(grid_t *g, int64_t x1, int64_t y1, int64_t x2, int64_t y2)
{
    if (occ_at(g, x1, y1, 0) || occ_at(g, x2, y2, 0)) return 1;
    if (x1 != x2 && y1 != y2) {
        int64_t s1x, s1y, s2x, s2y;
        if (x2 - x1 == y1 - y2) {
            s1x = imin(x1, x2); s1y = imin(y1, y2);
            s2x = imax(x1, x2); s2y = imax(y1, y2);
        } else {
            s1x = imin(x1, x2); s1y = imax(y1, y2);
            s2x = imax(x1, x2); s2y = imin(y1, y2);
        }
        if (occ_at(g, s1x, s1y, 0) || occ_at(g, s2x, s2y, 0)) return 1;
    }
    return 0;
}

/* GraphSearcher3D.isCollision, graph_search_3d.py:66-107 */
static int collision3d(grid_t *g, int64_t x1, int64_t y1, int64_t z1,
                       int64_t x2, int64_t y2, int64_t z2)
{
    if (occ_at(g, x1, y1, z1) || occ_at(g, x2, y2, z2)) return 1;
    int64_t dx = x2 - x1, dy = y2 - y1, dz = z2 - z1;
    int64_t m = imax(imax(llabs(dx), llabs(dy)), llabs(dz));
    if (m > 1) return 0;                                   /* :87-88 */
    int changes = (dx != 0) + (dy != 0) + (dz != 0);
    if (changes <= 1) return 0;                            /* :92-93 */
    if (changes == 2) {                                    /* :96-102 */
        if (dx != 0 && dy != 0)
            if (occ_at(g, x1 + dx, y1, z1) | occ_at(g, x1, y1 + dy, z1)) return 1;
        if (dx != 0 && dz != 0)
            if (occ_at(g, x1 + dx, y1, z1) | occ_at(g, x1, y1, z1 + dz)) return 1;
        if (dy != 0 && dz != 0)
            if (occ_at(g, x1, y1 + dy, z1) | occ_at(g, x1, y1, z1 + dz)) return 1;
        return 0;
    }
    /* three axes: the three face neighbours of the SOURCE cell only (:103-105) */
    return occ_at(g, x1 + dx, y1, z1) | occ_at(g, x1, y1 + dy, z1) | occ_at(g, x1, y1, z1 + dz);
}

int pmp_oracle_is_collision(int ndim, const int32_t *dims, const uint8_t *occ,
                            const int32_t *p1, const int32_t *p2)
{
    grid_t g = { ndim, dims[0], dims[1], ndim == 3 ? dims[2] : 1, occ, 0 };
    if (ndim == 2) return collision2d(&g, p1[0], p1[1], p2[0], p2[1]);
    return collision3d(&g, p1[0], p1[1], p1[2], p2[0], p2[1], p2[2]);
}

/* GraphSearcher.h graph_search.py:30-44 (math.hypot / manhattan);
 * GraphSearcher3D.h graph_search_3d.py:30-49 (math.sqrt(dx**2+dy**2+dz**2) / manhattan) */
static inline double heuristic_val(int heuristic, int64_t dx, int64_t dy, int64_t dz)
{
    dx = llabs(dx); dy = llabs(dy); dz = llabs(dz);
    if (heuristic == PMP_ORACLE_MANHATTAN) return (double)(dx + dy + dz);
    return sqrt((double)(dx * dx + dy * dy + dz * dz));
}

typedef struct {
    double *g, *h;
    int64_t *parent, *pos;   /* pos: index in `order`, -1 = not in CLOSED */
    int64_t *order, n_order;
} closed_t;

static int closed_alloc(closed_t *c, int64_t cells)
{
    c->g = (double *)malloc((size_t)cells * sizeof(double));
    c->h = (double *)malloc((size_t)cells * sizeof(double));
    c->parent = (int64_t *)malloc((size_t)cells * sizeof(int64_t));
    c->pos = (int64_t *)malloc((size_t)cells * sizeof(int64_t));
    c->order = (int64_t *)malloc((size_t)cells * sizeof(int64_t));
    c->n_order = 0;
    if (!c->g || !c->h || !c->parent || !c->pos || !c->order) return -1;
    for (int64_t i = 0; i < cells; ++i) c->pos[i] = -1;
    return 0;
}

static void closed_free(closed_t *c)
{
    free(c->g); free(c->h); free(c->parent); free(c->pos); free(c->order);
}

/* dict assignment CLOSED[node.current] = node: a new key is appended, an existing key keeps
 * its position and takes the new value (a_star.py:82; a_star3d.py:52). */
static inline void closed_set(closed_t *c, const rec_t *r)
{
    if (c->pos[r->cell] < 0) {
        c->pos[r->cell] = c->n_order;
        c->order[c->n_order++] = r->cell;
    }
    c->g[r->cell] = r->g;
    c->h[r->cell] = r->h;
    c->parent[r->cell] = r->parent;
}

int pmp_oracle_plan(int ndim, const int32_t *dims, const uint8_t *occ,
                    int n_motions, const int32_t *motions, const double *motion_g,
                    int algo, int heuristic,
                    const int32_t *start, const int32_t *goal,
                    double *cost_out,
                    int32_t *path_out, int64_t path_cap, int64_t *path_len,
                    int64_t *expand_cell, double *expand_g, double *expand_h,
                    int64_t *expand_parent, int64_t expand_cap, int64_t *n_expand,
                    double *g_field, int32_t *flags_out)
{
    if ((ndim != 2 && ndim != 3) || !dims || !occ || !motions || !start) return -1;
    if (algo < 0 || algo > 2) return -1;
    if (!goal && algo != PMP_ORACLE_DIJKSTRA) return -1;
    if (ndim == 2 && !motion_g) return -1;

    grid_t G = { ndim, dims[0], dims[1], ndim == 3 ? dims[2] : 1, occ, 0 };
    const int64_t X = G.X, Y = G.Y, Z = G.Z, cells = X * Y * Z;
    const int64_t sx = start[0], sy = start[1], sz = ndim == 3 ? start[2] : 0;
    if (!in_box(&G, sx, sy, sz)) return -1;
    int64_t gx = -1, gy = -1, gz = -1, goal_cell = -1;
    if (goal) {
        gx = goal[0]; gy = goal[1]; gz = ndim == 3 ? goal[2] : 0;
        if (!in_box(&G, gx, gy, gz)) return -1;
        goal_cell = (gz * Y + gy) * X + gx;
    }
    const int64_t start_cell = (sz * Y + sy) * X + sx;

    closed_t C;
    if (closed_alloc(&C, cells)) { closed_free(&C); return -2; }
    heap_t H = { NULL, 0, 0, ndim == 3 };
    uint64_t counter = 0;
    int reached = 0, rc = 0;

    rec_t s;
    memset(&s, 0, sizeof s);
    s.cell = start_cell; s.parent = start_cell; s.g = 0.0;
    if (ndim == 2) {
        /* Planner.__init__, utils/planner/planner.py:15: Node(start, start, 0, 0) */
        s.h = 0.0; s.f = 0.0;
    } else {
        /* a_star3d.py:37-40; dijkstra3d.py:33-34,55; gbfs3d.py:28-29,46 */
        s.h = (algo == PMP_ORACLE_DIJKSTRA) ? 0.0 : heuristic_val(heuristic, gx - sx, gy - sy, gz - sz);
        s.f = (algo == PMP_ORACLE_GBFS) ? s.h : s.g + s.h;
        s.counter = counter++;
    }
    if (heap_push(&H, &s)) { rc = -2; goto done; }

    while (H.n > 0) {
        rec_t node = heap_pop(&H);
        const int64_t cx = node.cell % X, cy = (node.cell / X) % Y, cz = node.cell / (X * Y);

        if (ndim == 2) {
            /* a_star.py:54-64 (dijkstra.py:51-61, gbfs.py:51-61 identical) */
            if (C.pos[node.cell] >= 0) continue;
            if (node.cell == goal_cell) {
                closed_set(&C, &node);
                reached = 1;
                break;
            }
            /* getNeighbor builds the full list first (a_star.py:95-96); the loop over it may
             * break early at the goal (a_star.py:75-77) -- occupancy lookups have no side
             * effects, so generation and consumption are fused here. */
            for (int m = 0; m < n_motions; ++m) {
                const int64_t nx = cx + motions[3 * m], ny = cy + motions[3 * m + 1];
                if (collision2d(&G, cx, cy, nx, ny)) continue;
                if (!in_box(&G, nx, ny, 0)) continue;
                const int64_t ncell = ny * X + nx;
                if (C.pos[ncell] >= 0) continue;                       /* a_star.py:68-69 */
                rec_t n;
                n.cell = ncell; n.parent = node.cell; n.counter = 0;
                n.g = node.g + motion_g[m];                            /* node.py:41 */
                if (algo == PMP_ORACLE_DIJKSTRA) n.h = 0.0;            /* dijkstra.py:74 */
                else n.h = heuristic_val(heuristic, gx - nx, gy - ny, 0);  /* a_star.py:72 */
                if (algo == PMP_ORACLE_GBFS) n.g = 0.0;                /* gbfs.py:75 */
                n.f = n.g + n.h;
                if (heap_push(&H, &n)) { rc = -2; goto done; }
                if (ncell == goal_cell) break;                         /* a_star.py:75-77 */
            }
            closed_set(&C, &node);                                     /* a_star.py:82 */
        } else {
            if (algo == PMP_ORACLE_GBFS) {
                if (C.pos[node.cell] >= 0) continue;                   /* gbfs3d.py:55-56 */
            } else {
                /* a_star3d.py:47-50; dijkstra3d.py:61-63: re-expansion on strictly better g */
                if (C.pos[node.cell] >= 0 && node.g >= C.g[node.cell]) continue;
            }
            closed_set(&C, &node);                                     /* a_star3d.py:52 */
            if (node.cell == goal_cell) { reached = 1; break; }        /* a_star3d.py:60-63 */

            for (int m = 0; m < n_motions; ++m) {
                const int64_t dx = motions[3 * m], dy = motions[3 * m + 1], dz = motions[3 * m + 2];
                const int64_t nx = cx + dx, ny = cy + dy, nz = cz + dz;
                if (algo == PMP_ORACLE_DIJKSTRA) {
                    /* Dijkstra3D.getNeighbor, dijkstra3d.py:102-124 */
                    if (!in_box(&G, nx, ny, nz)) continue;             /* :106-107 */
                    if (occ_at(&G, nx, ny, nz)) continue;              /* :110-111 */
                    if ((dx != 0 && occ_at(&G, cx + dx, cy, cz)) ||
                        (dy != 0 && occ_at(&G, cx, cy + dy, cz)) ||
                        (dz != 0 && occ_at(&G, cx, cy, cz + dz))) continue;  /* :114-118 */
                }
                if (collision3d(&G, cx, cy, cz, nx, ny, nz)) continue; /* a_star3d.py:86-89 */
                if (!in_box(&G, nx, ny, nz)) continue;
                const int64_t ncell = (nz * Y + ny) * X + nx;
                rec_t n;
                n.cell = ncell; n.parent = node.cell;
                if (algo == PMP_ORACLE_GBFS) {
                    if (C.pos[ncell] >= 0) continue;                   /* gbfs3d.py:74-75 */
                    n.h = heuristic_val(heuristic, gx - nx, gy - ny, gz - nz);
                    n.g = 0.0; n.f = n.h;                              /* gbfs3d.py:77-79 */
                } else {
                    /* a_star3d.py:67: node.g + dist(node, node_n); planner3d.py:22-27 */
                    const double tg = node.g + sqrt((double)(dx * dx + dy * dy + dz * dz));
                    if (C.pos[ncell] >= 0 && tg >= C.g[ncell]) continue;   /* a_star3d.py:68-70 */
                    n.g = tg;
                    n.h = (algo == PMP_ORACLE_DIJKSTRA) ? 0.0
                          : heuristic_val(heuristic, gx - nx, gy - ny, gz - nz);
                    n.f = tg + n.h;                                    /* a_star3d.py:75 */
                }
                n.counter = counter++;
                if (heap_push(&H, &n)) { rc = -2; goto done; }
            }
        }
    }

    /* extractPath: a_star.py:98-117 (goal->start, cost += hypot) /
     * a_star3d.py:91-106 (same walk, then reversed) */
    {
        double cost = 0.0;
        int64_t n_path = 0;
        if (reached) {
            int64_t cell = goal_cell;
            for (;;) {
                if (path_out) {
                    if (n_path >= path_cap) { rc = -3; goto done; }
                    path_out[3 * n_path] = (int32_t)(cell % X);
                    path_out[3 * n_path + 1] = (int32_t)((cell / X) % Y);
                    path_out[3 * n_path + 2] = (int32_t)(cell / (X * Y));
                }
                ++n_path;
                if (cell == start_cell) break;
                const int64_t p = C.parent[cell];
                const int64_t ddx = cell % X - p % X, ddy = (cell / X) % Y - (p / X) % Y,
                              ddz = cell / (X * Y) - p / (X * Y);
                cost += sqrt((double)(ddx * ddx + ddy * ddy + ddz * ddz));
                cell = p;
            }
            if (ndim == 3 && path_out) {                               /* a_star3d.py:105 */
                for (int64_t i = 0, j = n_path - 1; i < j; ++i, --j)
                    for (int k = 0; k < 3; ++k) {
                        int32_t t = path_out[3 * i + k];
                        path_out[3 * i + k] = path_out[3 * j + k];
                        path_out[3 * j + k] = t;
                    }
            }
        } else {
            cost = INFINITY;   /* 3D: float("inf") (a_star3d.py:78); 2D returns [] (a_star.py:83) */
        }
        if (cost_out) *cost_out = cost;
        if (path_len) *path_len = n_path;
    }

    if (n_expand) *n_expand = C.n_order;
    if (expand_cell || expand_g || expand_h || expand_parent) {
        if (C.n_order > expand_cap) { rc = -3; goto done; }
        for (int64_t i = 0; i < C.n_order; ++i) {
            const int64_t c = C.order[i];
            if (expand_cell) expand_cell[i] = c;
            if (expand_g) expand_g[i] = C.g[c];
            if (expand_h) expand_h[i] = C.h[c];
            if (expand_parent) expand_parent[i] = C.parent[c];
        }
    }
    if (g_field) {
        for (int64_t i = 0; i < cells; ++i) g_field[i] = C.pos[i] >= 0 ? C.g[i] : INFINITY;
    }

done:
    if (flags_out) *flags_out = (reached ? PMP_ORACLE_REACHED : 0) | (G.escaped ? PMP_ORACLE_ESCAPED : 0);
    free(H.a);
    closed_free(&C);
    return rc;
}

int pmp_oracle_max_threads(void)
{
    long n = sysconf(_SC_NPROCESSORS_ONLN);
    return n > 0 ? (int)n : 1;
}

typedef struct {
    int ndim; const int32_t *dims; const uint8_t *occ;
    int n_motions; const int32_t *motions; const double *motion_g;
    int algo, heuristic; int64_t nq;
    const int32_t *starts, *goals;
    double *cost_out; int64_t *n_expand_out, *path_len_out; int32_t *flags_out;
    int64_t next;   /* work counter, __atomic */
    int rc;
} batch_t;

static void *batch_worker(void *arg)
{
    batch_t *b = (batch_t *)arg;
    for (;;) {
        const int64_t q = __atomic_fetch_add(&b->next, 1, __ATOMIC_RELAXED);
        if (q >= b->nq) break;
        double cost = 0.0;
        int64_t n_exp = 0, plen = 0;
        int32_t flags = 0;
        int rc = pmp_oracle_plan(b->ndim, b->dims, b->occ, b->n_motions, b->motions, b->motion_g,
                                 b->algo, b->heuristic, b->starts + 3 * q, b->goals + 3 * q,
                                 &cost, NULL, 0, &plen, NULL, NULL, NULL, NULL, 0, &n_exp, NULL,
                                 &flags);
        if (b->cost_out) b->cost_out[q] = cost;
        if (b->n_expand_out) b->n_expand_out[q] = n_exp;
        if (b->path_len_out) b->path_len_out[q] = plen;
        if (b->flags_out) b->flags_out[q] = flags;
        if (rc) __atomic_store_n(&b->rc, rc, __ATOMIC_RELAXED);
    }
    return NULL;
}

int pmp_oracle_plan_batch(int ndim, const int32_t *dims, const uint8_t *occ,
                          int n_motions, const int32_t *motions, const double *motion_g,
                          int algo, int heuristic, int64_t nq,
                          const int32_t *starts, const int32_t *goals,
                          double *cost_out, int64_t *n_expand_out, int64_t *path_len_out,
                          int32_t *flags_out, int n_threads)
{
    if (n_threads <= 0) n_threads = pmp_oracle_max_threads();
    if (n_threads > nq) n_threads = (int)(nq > 0 ? nq : 1);
    batch_t b = { ndim, dims, occ, n_motions, motions, motion_g, algo, heuristic, nq,
                  starts, goals, cost_out, n_expand_out, path_len_out, flags_out, 0, 0 };
    pthread_t *th = (pthread_t *)malloc((size_t)n_threads * sizeof(pthread_t));
    if (!th) return -2;
    int started = 0;
    for (int i = 1; i < n_threads; ++i)
        if (pthread_create(&th[started], NULL, batch_worker, &b) == 0) ++started;
    batch_worker(&b);
    for (int i = 0; i < started; ++i) pthread_join(th[i], NULL);
    free(th);
    return b.rc;
}

/* ---------------------------------------------------------------------------------------------
 * Line of sight (any-angle planners).
 * ------------------------------------------------------------------------------------------- */

/* ThetaStar.lineOfSight, global_planner/graph_search/theta_star.py:110-171.  The reference runs
 * this Bresenham variant on Python floats (s_x = (x2 - x1) / d_x, tau = (d_y - d_x) / 2); every
 * quantity is an integer or a half-integer, so it is restated on integers with e and tau doubled
 * (e2 = 2 e, tau2 = 2 tau): the comparisons e > tau, e < tau are unchanged. */
static int los2d(grid_t *g, int64_t x1, int64_t y1, int64_t x2, int64_t y2)
{
    if (occ_at(g, x1, y1, 0) || occ_at(g, x2, y2, 0)) return 0;        /* :122, :128-131 */
    const int64_t d_x = llabs(x2 - x1), d_y = llabs(y2 - y1);
    const int64_t s_x = (x2 > x1) - (x2 < x1), s_y = (y2 > y1) - (y2 < y1);
    int64_t x = x1, y = y1, e2 = 0;
    int64_t guard = 4 * (d_x + d_y) + 8;
    if (d_x > d_y) {                                                    /* :140-153 */
        const int64_t tau2 = d_y - d_x;
        while (x != x2) {
            if (e2 > tau2) { x += s_x; e2 -= 2 * d_y; }
            else if (e2 < tau2) { y += s_y; e2 += 2 * d_x; }
            else { x += s_x; y += s_y; e2 += 2 * (d_x - d_y); }
            if (occ_at(g, x, y, 0)) return 0;
            if (--guard < 0) return 0;
        }
    } else {                                                            /* :155-169 */
        const int64_t tau2 = d_x - d_y;
        while (y != y2) {
            if (e2 > tau2) { y += s_y; e2 -= 2 * d_x; }
            else if (e2 < tau2) { x += s_x; e2 += 2 * d_y; }
            else { x += s_x; y += s_y; e2 += 2 * (d_y - d_x); }
            if (occ_at(g, x, y, 0)) return 0;
            if (--guard < 0) return 0;
        }
    }
    return 1;
}

/* ThetaStar3D.lineOfSight, theta_star3d.py:139-215 (LazyThetaStar3D.lineOfSight,
 * lazy_theta_star3d.py:158-235, is the same walk): integer 3D Bresenham along the dominant
 * axis; in_bounds() = membership in env.grid_map = the box. */
static int los3d(grid_t *g, int64_t x0, int64_t y0, int64_t z0, int64_t x1, int64_t y1, int64_t z1)
{
    if (occ_at(g, x0, y0, z0) || occ_at(g, x1, y1, z1)) return 0;      /* :148-149 */
    const int64_t dx = llabs(x1 - x0), dy = llabs(y1 - y0), dz = llabs(z1 - z0);
    const int64_t sx = x1 >= x0 ? 1 : -1, sy = y1 >= y0 ? 1 : -1, sz = z1 >= z0 ? 1 : -1;
    int64_t x = x0, y = y0, z = z0;
    if (dx >= dy && dx >= dz) {                                         /* :164-178 */
        int64_t err_y = dx / 2, err_z = dx / 2;
        while (x != x1) {
            x += sx; err_y -= dy; err_z -= dz;
            if (err_y < 0) { y += sy; err_y += dx; }
            if (err_z < 0) { z += sz; err_z += dx; }
            if (occ_at(g, x, y, z)) return 0;
        }
        return 1;
    }
    if (dy >= dx && dy >= dz) {                                         /* :180-194 */
        int64_t err_x = dy / 2, err_z = dy / 2;
        while (y != y1) {
            y += sy; err_x -= dx; err_z -= dz;
            if (err_x < 0) { x += sx; err_x += dy; }
            if (err_z < 0) { z += sz; err_z += dy; }
            if (occ_at(g, x, y, z)) return 0;
        }
        return 1;
    }
    int64_t err_x = dz / 2, err_y = dz / 2;                             /* :197-214 */
    while (z != z1) {
        z += sz; err_x -= dx; err_y -= dy;
        if (err_x < 0) { x += sx; err_x += dz; }
        if (err_y < 0) { y += sy; err_y += dz; }
        if (occ_at(g, x, y, z)) return 0;
    }
    return 1;
}

int pmp_oracle_line_of_sight(int ndim, const int32_t *dims, const uint8_t *occ, int64_t n,
                             const int32_t *a, const int32_t *b, uint8_t *out)
{
    if ((ndim != 2 && ndim != 3) || !dims || !occ || n < 0 || (n && (!a || !b || !out))) return -1;
    grid_t g = { ndim, dims[0], dims[1], ndim == 3 ? dims[2] : 1, occ, 0 };
    for (int64_t i = 0; i < n; ++i) {
        const int32_t *p = a + 3 * i, *q = b + 3 * i;
        out[i] = (uint8_t)(ndim == 2 ? los2d(&g, p[0], p[1], q[0], q[1])
                                     : los3d(&g, p[0], p[1], p[2], q[0], q[1], q[2]));
    }
    return 0;
}
