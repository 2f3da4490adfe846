/*
 * pmp_oracle.h -- CPU restatement of the reference's grid graph-search hot path.
 *
 * TEST INFRASTRUCTURE.  This is the parity oracle for python_motion_planning_b200: a plain-C
 * restatement of the algorithms in /root/reference/src/python_motion_planning (pure Python).
 * Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / `--impl reference` legs
 * may load it.  The product path (python_motion_planning_b200/) never does.
 *
 * Parity status: PINNED.  tests/test_oracle_golden.py checks this file against
 *   (1) the 1 500 known answers of the reference's own 3d_pathfinding_results.csv
 *       (Distance, Visited Nodes) and
 *   (2) outputs of the unmodified reference run in the build container
 *       (tools/make_golden.py -> the .npz files under tests/golden/): cost, path, expansion order, g, h.
 *
 * Every function cites the reference file:line it restates (paths relative to
 * /root/reference/src/python_motion_planning/).
 */
#ifndef PMP_ORACLE_H
#define PMP_ORACLE_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

enum { PMP_ORACLE_ASTAR = 0, PMP_ORACLE_DIJKSTRA = 1, PMP_ORACLE_GBFS = 2 };
enum { PMP_ORACLE_EUCLIDEAN = 0, PMP_ORACLE_MANHATTAN = 1 };

/* flags_out bits */
enum { PMP_ORACLE_REACHED = 1, PMP_ORACLE_ESCAPED = 2 };

/*
 * One plan() call of the reference.
 *
 * ndim = 2 selects the 2D planner family (a_star.py / dijkstra.py / gbfs.py),
 * ndim = 3 the 3D family (a_star3d.py / dijkstra3d.py / gbfs3d.py); they differ in heap
 * record, re-expansion rule, edge weight source and path order exactly as the reference does.
 *
 * dims      : {X, Y, Z} (Z = 1 for ndim 2)
 * occ       : dense occupancy, index (z*Y + y)*X + x, non-zero = obstacle.  Cells outside the
 *             box count as obstacles and raise PMP_ORACLE_ESCAPED in flags_out when touched
 *             (the reference's A-star/GBFS do not bound-check; with the default walls they never
 *             leave the box, so a set ESCAPED bit marks an input outside the oracle's domain).
 * motions   : n_motions x 3 integer offsets in env.motions order (env.py:52-55, env3d.py:56-70)
 * motion_g  : n_motions edge costs; used by the 2D family only (node.py:41); the 3D family
 *             recomputes sqrt(dx^2+dy^2+dz^2) (a_star3d.py:67, planner3d.py:22-27)
 * goal      : may be NULL only with algo = DIJKSTRA: run to exhaustion (full cost-to-go field)
 * path_out  : path_cap x 3 coordinates, in the family's order (2D goal->start, 3D start->goal)
 * expand_*  : CLOSED.values() in insertion order (cell index, g, h, parent cell index)
 * g_field   : optional dense array (cells) receiving CLOSED g, +inf where not closed
 *
 * returns 0, or <0 on bad arguments / out of memory / output capacity exceeded.
 */
int pmp_oracle_plan(int ndim, const int32_t *dims, const uint8_t *occ,
                    int n_motions, const int32_t *motions, const double *motion_g,
                    int algo, int heuristic,
                    const int32_t *start, const int32_t *goal,
                    double *cost_out,
                    int32_t *path_out, int64_t path_cap, int64_t *path_len,
                    int64_t *expand_cell, double *expand_g, double *expand_h,
                    int64_t *expand_parent, int64_t expand_cap, int64_t *n_expand,
                    double *g_field, int32_t *flags_out);

/*
 * Batch of independent plan() calls on one grid (pthreads over queries; n_threads <= 0 = one per online core).
 * starts/goals : nq x 3.  cost_out, n_expand_out, path_len_out, flags_out : nq.
 * Paths are not returned (cost/expansion counts only) -- used as the timed CPU baseline.
 */
int pmp_oracle_plan_batch(int ndim, const int32_t *dims, const uint8_t *occ,
                          int n_motions, const int32_t *motions, const double *motion_g,
                          int algo, int heuristic, int64_t nq,
                          const int32_t *starts, const int32_t *goals,
                          double *cost_out, int64_t *n_expand_out, int64_t *path_len_out,
                          int32_t *flags_out, int n_threads);

/* ThetaStar.lineOfSight (theta_star.py:110-171, ndim 2) / ThetaStar3D.lineOfSight
 * (theta_star3d.py:139-215, ndim 3) for n ordered pairs a[i] -> b[i] (n x 3 each);
 * out[i] = 1 if the reference returns True.  Cells outside the box count as blocked. */
int pmp_oracle_line_of_sight(int ndim, const int32_t *dims, const uint8_t *occ, int64_t n,
                             const int32_t *a, const int32_t *b, uint8_t *out);

/* GraphSearcher.isCollision (graph_search.py:61-87) / GraphSearcher3D.isCollision
 * (graph_search_3d.py:66-107) for one ordered pair of cells; 1 = collision. */
int pmp_oracle_is_collision(int ndim, const int32_t *dims, const uint8_t *occ,
                            const int32_t *p1, const int32_t *p2);

int pmp_oracle_max_threads(void);

#ifdef __cplusplus
}
#endif
#endif
