/*
 * pmp_b200.h -- C ABI of libpmp_b200.so: B200 (sm_100a) grid graph search behind the
 * python_motion_planning planner API.
 *
 * The reference (Slenderman00/python_motion_planning) is pure Python and has no FFI; its seam
 * is the planner class contract.  Each entry point below names the reference code it replaces
 * (paths relative to /root/reference/src/python_motion_planning/).  The Python host layer
 * (python_motion_planning_b200/) binds these with ctypes; INTEGRATION.md shows the stub a
 * reference maintainer would add.
 *
 * Conventions
 *   - every function returns 0 on success, <0 on error; pmp_last_error() gives the text of the
 *     last error raised on the calling thread.  No exceptions cross the ABI.
 *   - "dev" pointers are CUDA device pointers owned by the caller (the Python layer carries
 *     them in torch tensors); "host" pointers are ordinary host memory.
 *   - `stream` is a cudaStream_t passed as void* (NULL = legacy default stream).  Calls are
 *     asynchronous on that stream unless stated otherwise.
 *   - coordinates are int32 triples (x, y, z); z = 0 for 2D grids.
 *   - "no path" is data, not an error: cost = +inf, reached = 0.
 *   - a handle is not thread-safe; use one handle per host thread / stream.
 *
 * Cell indexing.  The device grid is padded up to whole tiles: dims (X, Y, Z) -> (PX, PY, PZ),
 * multiples of the tile (32 x 32 in 2D, 16 x 8 x 8 in 3D).  A "padded cell index" is
 * (z * PY + y) * PX + x.  Cost fields (g) are dense arrays over padded cells, one per query.
 * Cells outside the original box are permanently blocked (the reference's default walls seal
 * the box: utils/environment/env.py:68-74, env3d.py:84-97).
 */
#ifndef PMP_B200_H
#define PMP_B200_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

/* 2: + pmp_bestfirst, pmp_lpa_star, pmp_d_star (additive; every version-1 entry point is unchanged) */
#define PMP_B200_ABI_VERSION 3

typedef struct pmp_grid pmp_grid; /* opaque: occupancy bitmask + motion table on one device */

enum pmp_algo {
    PMP_ASTAR = 0, PMP_DIJKSTRA = 1, PMP_GBFS = 2,
    PMP_THETA_STAR = 3, PMP_LAZY_THETA_STAR = 4, PMP_JPS = 5 /* pmp_bestfirst only */
};
enum pmp_heuristic { PMP_H_EUCLIDEAN = 0, PMP_H_MANHATTAN = 1 };
enum pmp_weight { PMP_W_F64 = 0, PMP_W_U32 = 1 };

/* per-query status written by pmp_sssp */
enum pmp_status {
    PMP_ST_RUNNING = 0,
    PMP_ST_EXHAUSTED = 1, /* every reachable cell settled; goal (if any) not reached */
    PMP_ST_REACHED = 2,   /* g[goal] is final */
    PMP_ST_BAD_QUERY = 3, /* start or goal outside the box or on an obstacle */
    PMP_ST_REPLAN_OVERFLOW = 4, /* pmp_replan_update: scratch too small for the invalidated region */
    PMP_ST_ERROR = 5      /* the search did not finish (a queue wait timed out): the result is invalid */
};

const char *pmp_last_error(void);
int pmp_abi_version(void);
/* number of CUDA devices visible; <0 if the CUDA runtime cannot initialise */
int pmp_device_count(void);

/* ---- environment: replaces Grid / Grid3D occupancy (utils/environment/env.py:41-80,
 *      env3d.py:43-103: a Python set of coordinate tuples + a cKDTree rebuilt per update) ---- */

/* Creates an empty (all-free inside the box) grid on the current CUDA device.
 * ndim 2 or 3; dims = {X, Y, Z} (Z ignored for ndim 2).  Installs the reference's default
 * motion table (env.py:52-55: 8 moves, 1 / sqrt 2; env3d.py:56-70: 26 moves, 1 / sqrt 2 / sqrt 3). */
int pmp_grid_create(int ndim, const int32_t *dims, pmp_grid **out);
int pmp_grid_destroy(pmp_grid *g);
/* out[0..2] = padded dims, out[3..5] = tile dims, out[6] = number of tiles, out[7] = words per
 * bitmask row */
int pmp_grid_layout(const pmp_grid *g, int64_t *out8);

/* Grid.update(obstacles) (env.py:78-80 / env3d.py:101-103): replace the occupancy by the
 * given obstacle set.  coords_dev: n x 3 int32 on the device; coordinates outside the box are
 * ignored (the default Grid3D writes its side walls at z = -1, env3d.py:86-92). */
int pmp_grid_set_cells(pmp_grid *g, const int32_t *coords_dev, int64_t n, void *stream);
/* Same from a dense device array occ_dev[(z*Y + y)*X + x] != 0 (unpadded box). */
int pmp_grid_set_dense(pmp_grid *g, const uint8_t *occ_dev, void *stream);
/* Same from a dense device array in the caller's own index order occ_dev[(x*Y + y)*Z + z] != 0 --
 * the memory layout of a C-ordered array indexed [x, y(, z)] like the reference's coordinate
 * tuples; transposed on the fly, no host-side or extra device pass. */
int pmp_grid_set_dense_xyz(pmp_grid *g, const uint8_t *occ_dev, void *stream);
/* Occupancy already packed by the caller: rows_dev = [Z][Y][ceil(X/32)] 32-bit words, bit i of
 * word w <-> x = 32 w + i, 1 = obstacle (numpy: np.packbits(occ.transpose(2, 1, 0), axis=-1,
 * bitorder="little") viewed as uint32 when X is a multiple of 32).  1/8 of the bytes of the dense
 * forms; 2D grids: Z = 1. */
int pmp_grid_set_packed(pmp_grid *g, const uint32_t *rows_dev, void *stream);
/* The same from PINNED HOST memory, streamed: the rows are copied in z chunks of ~planes_per_chunk planes
 * on copy_stream (the copy engine) straight into the grid's bitmask, each chunk followed by a ready flag,
 * and the call returns at once.  A unit-weight field search (pmp_sssp full_field / PMP_W_U32 on the six
 * unit moves) launched on search_stream right after it starts while the upload is still running: its
 * kernel holds a tile back (re-queues it) until the chunks it needs are there.  Every other entry point
 * first waits for the whole upload.  Needs the rows to be the grid's own layout: 3D, X % 32 == 0, Y % 8 == 0, Z % 8 == 0 (else -1:
 * upload to the device yourself and call pmp_grid_set_packed).  The host buffer must stay untouched
 * until the next operation on search_stream after the search has completed. */
int pmp_grid_set_packed_host_streamed(pmp_grid *g, const uint32_t *rows_host_pinned, int planes_per_chunk,
                                      void *search_stream, void *copy_stream);
/* Toggle single cells without rebuilding (obstacle edits between plans). value: 1 block, 0 free */
int pmp_grid_edit_cells(pmp_grid *g, const int32_t *coords_dev, int64_t n, int value, void *stream);
/* Copy the device bitmask to host: words_per_row * PY * PZ uint32 (bit x%32 of word x/32). Synchronous. */
int pmp_grid_read_bits(const pmp_grid *g, uint32_t *bits_host, int64_t n_words);
/* number of free cells inside the box (synchronous) */
int pmp_grid_count_free(const pmp_grid *g, int64_t *n_free);

/* env.motions (env.py:52-55, env3d.py:56-70; planners alias it at construction,
 * global_planner/graph_search/graph_search.py:26): k <= 26 offsets in {-1,0,1}^3 (host, k x 3)
 * with positive costs (host, k).  The 2D planners accumulate motion.g (node.py:41); the 3D
 * planners use the Euclidean length of the offset (a_star3d.py:67) -- the host layer passes the
 * right numbers.  Returns <0 for offsets outside {-1,0,1}. */
int pmp_grid_set_motions(pmp_grid *g, int k, const int32_t *offsets_host, const double *cost_host);
/* Edge rule.  allow = 0 (default): a diagonal move also needs the face neighbours of its source free
 * (GraphSearcher.isCollision, graph_search.py:61-87; GraphSearcher3D.isCollision,
 * graph_search_3d.py:66-107).  allow = 1: only the target must be free -- the graph the
 * reference's JPS searches (JPS.jump / detectForceNeighbor, jps.py:85-165, never call isCollision). */
int pmp_grid_set_corner_cutting(pmp_grid *g, int allow);

/* ---- search: replaces AStar/Dijkstra(.plan) expansion loops
 *      (global_planner/graph_search/a_star.py:39-83, dijkstra.py:36-85, a_star3d.py:33-78,
 *      dijkstra3d.py:39-87) incl. getNeighbor + isCollision (graph_search.py:61-87,
 *      graph_search_3d.py:66-107) ---- */

typedef struct pmp_sssp_params {
    int32_t algo;       /* PMP_ASTAR: bucket/prune by g + h;  PMP_DIJKSTRA: by g */
    int32_t heuristic;  /* pmp_heuristic; only admissible forms are used for ordering/pruning */
    int32_t weight;     /* PMP_W_F64, or PMP_W_U32 (all motion costs must be integers) */
    int32_t full_field; /* 1: ignore goals, settle every reachable cell (cost-to-go field) */
    double window;      /* width of the global key window (delta); <= 0 picks a default */
    int32_t max_rounds; /* safety bound on relaxation rounds; <= 0 = unlimited */
    int32_t reserved;   /* tuning: in-tile sub-bucket width in units of the smallest motion cost (0/1 = 1) */
} pmp_sssp_params;

/* bytes of one query's cost field and of the scratch workspace for nq concurrent queries */
int pmp_sssp_sizes(const pmp_grid *g, int64_t nq, int weight, int64_t *g_bytes_per_query,
                   int64_t *ws_bytes);

/* Runs nq independent single-source searches that share the grid.
 *   starts_dev, goals_dev : nq x 3 int32 (goals_dev may be NULL with full_field = 1)
 *   g_dev      : nq cost fields (padded cells; 8 B or 4 B per cell), overwritten, in the internal
 *                word format (see pmp_field_finalize).
 *   ws_dev     : scratch of >= ws_bytes from pmp_sssp_sizes
 *   gcost_dev  : nq doubles, g[goal] (+inf if not reached);  status_dev : nq pmp_status
 *   stats_dev  : optional 16 x uint64: [0] rounds, [1] tile activations, [2] cells expanded,
 *                [3] edges relaxed, [4] worklist high-water, [5..7] SM cycles summed over tile
 *                activations in load / settle / publish, [8] kernel ns, [9..11] ns CTA 0 spent in
 *                scan / process / waiting at the round barrier, [12] in-tile sub-iterations.
 *                (A full unit-weight u32 field on a 3D grid with the six unit moves is computed by k_bfs,
 *                csrc/pmp_bfs.cu: [1] tile activations, [2] cells lowered, [5] tiles re-queued, [7] entries
 *                pushed into a neighbour GPU's ring, [8] kernel ns, [12] levels swept, [13] activations that
 *                found nothing new, [14] sweep passes, [15] error bits; [4] [6] [11] [3] [0] = cycles per phase,
 *                filled only by builds with -DBFS_CLOCKS=1.)
 * Fields are bit-exact to a sequential Dijkstra on the same directed graph: every value is the
 * minimum over in-edges of fl(g(u) + w). */
int pmp_sssp(pmp_grid *g, const pmp_sssp_params *params, int64_t nq,
             const int32_t *starts_dev, const int32_t *goals_dev,
             void *g_dev, void *ws_dev, int64_t ws_bytes,
             double *gcost_dev, int32_t *status_dev, uint64_t *stats_dev, void *stream);

/* Cost fields are kept in an internal word format while searching (value << 1 | clean flag).
 * This rewrites n_cells = nq * padded cells in place to plain values: IEEE doubles with +inf
 * for unreached cells (PMP_W_F64) or uint32 with 0xFFFFFFFF (PMP_W_U32).  A finalized field can
 * no longer be passed to pmp_extract_paths / pmp_count_settled. */
int pmp_field_finalize(void *g_dev, int weight, int64_t n_cells, void *stream);

/* counts cells with a finite value per query (N_settled of the GTEPS metric); if
 * gcost_dev != NULL only cells with g + h <= gcost (A-star's closed region) are counted;
 * heuristic < 0 means h = 0 (Dijkstra: cells with g <= gcost). */
int pmp_count_settled(const pmp_grid *g, int weight, int64_t nq, const void *g_dev,
                      const int32_t *goals_dev, const double *gcost_dev, int heuristic,
                      uint64_t *counts_dev, void *stream);

/* ---- path extraction: replaces extractPath (a_star.py:98-117, a_star3d.py:91-106) ----
 * Descends the cost field from each goal: at v pick the in-neighbour u (legal edge u->v under
 * the same directed rule) minimising g(u) + w(u->v), lowest motion index on ties.
 *   path_dev     : nq x path_cap uint32 padded cell indices, goal first
 *   path_len_dev : nq int32 (0 = no path, -1 = path_cap too small)
 *   cost_dev     : nq doubles, sum of Euclidean step lengths accumulated goal->start, the
 *                  reference's summation order (a_star.py:109-116) */
int pmp_extract_paths(const pmp_grid *g, int weight, int64_t nq,
                      const int32_t *starts_dev, const int32_t *goals_dev, const void *g_dev,
                      uint32_t *path_dev, int64_t path_cap, int32_t *path_len_dev,
                      double *cost_dev, void *stream);

/* ---- shortest-path certificate (no reference counterpart; tests + slab driver) ----
 * Checks, for the planes z in [z_lo, z_hi) of ONE cost field in the internal word format, that
 * g(source) = 0 and every other free cell satisfies g(v) = min over legal in-edges u->v of
 * fl(g(u) + w) (+inf when no in-neighbour is reached), which characterises the shortest-path
 * field.  out_dev: 3 x uint64 = {violating cells, reached cells, cells still flagged pending}.
 * source_dev: 3 x int32 local coordinates, or NULL when the source lies outside this grid. */
int pmp_verify_field(const pmp_grid *g, int weight, const void *g_dev, const int32_t *source_dev,
                     int z_lo, int z_hi, uint64_t *out_dev, void *stream);

/* ---- greedy best-first: replaces GBFS.plan / GBFS3D.plan (gbfs.py:36-86, gbfs3d.py:34-82) ----
 * One warp per query runs the reference's sequential loop (binary heap keyed by h, FIFO
 * counter as in gbfs3d.py:46).  family: 2 = 2D record/tie rule, 3 = 3D.
 *   parent_dev  : nq x padded cells uint32 scratch (closed set + parents); heap_dev: nq x heap_cap x 16 B
 *   outputs as pmp_extract_paths, plus n_expand_dev (len(CLOSED)). */
int pmp_gbfs(const pmp_grid *g, int heuristic, int64_t nq,
             const int32_t *starts_dev, const int32_t *goals_dev,
             uint32_t *parent_dev, void *heap_dev, int64_t heap_cap,
             uint32_t *path_dev, int64_t path_cap, int32_t *path_len_dev,
             double *cost_dev, int32_t *n_expand_dev, void *stream);

/* ---- reference-order search: AStar / Dijkstra / GBFS / ThetaStar / LazyThetaStar / JPS .plan() and
 *      their 3D classes run with the reference's own heap discipline (a_star.py:39-117,
 *      dijkstra.py:36-85, gbfs.py:36-86, theta_star.py:44-108, lazy_theta_star.py:44-118,
 *      jps.py:38-165, a_star3d.py:33-106, dijkstra3d.py:39-143, gbfs3d.py:34-110,
 *      theta_star3d.py:38-107, lazy_theta_star3d.py:41-133, jps3d.py:42-246): CPython heapq sift order, Node.__lt__ (2D handles) or
 *      (f, h, counter) tuples (3D handles), dict-ordered CLOSED, parent-pointer extractPath.
 *      Unlike pmp_sssp + pmp_extract_paths, which guarantee the cost only, this reproduces the
 *      returned path among equal-cost ones, len(expand) ("Visited Nodes", examples/benchmark.py:85)
 *      and every expand record.  One warp per query; meant for the reference's own grid sizes.
 *   heap_dev          : nq x heap_cap records of PMP_BESTFIRST_HEAP_REC_BYTES (OPEN)
 *   closed_g/h_dev    : nq x padded cells doubles, closed_parent_dev: nq x padded cells uint32
 *                       (CLOSED[cell].g / .h / .parent as padded cell index; 0xFFFFFFFF = not closed)
 *   order_dev         : nq x order_cap uint32, the keys of CLOSED in insertion order = `expand`
 *   path_dev, path_len_dev, cost_dev as pmp_extract_paths (goal first; -2 = heap or order
 *   capacity exceeded); n_expand_dev = len(CLOSED). */
#define PMP_BESTFIRST_HEAP_REC_BYTES 40
int pmp_bestfirst(const pmp_grid *g, int algo, int heuristic, int64_t nq,
                  const int32_t *starts_dev, const int32_t *goals_dev,
                  void *heap_dev, int64_t heap_cap, double *closed_g_dev, double *closed_h_dev,
                  uint32_t *closed_parent_dev, uint32_t *order_dev, int64_t order_cap,
                  uint32_t *path_dev, int64_t path_cap, int32_t *path_len_dev, double *cost_dev,
                  int32_t *n_expand_dev, void *stream);

/* ---- LPA* / D* in the reference's processing order (row f1, reference-order mode).  LPAStar /
 *      LPAStar3D (lpa_star.py:62-230, lpa_star3d.py:57-225: computeShortestPath, updateVertex,
 *      calculateKey, greedy extractPath, apply_change / OnPress) and DStar / DStar3D (d_star.py:60-291,
 *      d_star3d.py:68-280: processState, insert, delete, modify, extractPath, OnPress /
 *      apply_dynamic_obstacles) with the reference's own containers -- OPEN as a Python list searched by
 *      min() and edited by list.remove(), heapq.heappush on it -- so cost, path and len(EXPAND)
 *      ("Visited Nodes") equal the reference's, also where it is not optimal.  pmp_replan_* is the
 *      parallel repair that returns the optimum.  One thread runs the algorithm; for the
 *      reference's own grid sizes.  All state lives in caller-owned device arrays that must be kept
 *      between calls: per padded cell g, rhs, key[2] (LPA*) or h, k, parent, tag (D*), the list
 *      (list_cap cells) and ctl_dev (4 x int64: [0] list length, [1] status).
 *   variant (pmp_lpa_star): 0 = LPA*, 1 = D* Lite (DStarLite, d_star_lite.py:34-187, 2D grids: LPA* run
 *         from the goal with keys shifted by km; ctl_dev[2] keeps km).  With variant 1, op 1 is
 *         DStarLite.OnPress: the greedy walk from the start that toggles cell change_dev = {x, y, z, -}
 *         ITSELF after its first step (the kernel flips the bit of the handle's bitmask; the caller
 *         must not edit it beforehand) and repairs; path_dev then holds the walked cells, start first.
 *   op 0: initialise the state and plan.  start_goal_dev: 2 x 3 int32.
 *   op 1: LPA*: change_dev = {x, y, z, freed}: the occupancy of that cell was just edited
 *         (pmp_grid_edit_cells; freed = 1 if an obstacle was removed) -> updateVertex of the cell /
 *         its neighbours, then plan.  D*: obstacles were just added -> walk the back pointers from the
 *         start and repair (modify) where an edge collides.
 *   outputs: expand_dev = EXPAND (cells), n_expand_dev its length; path_dev: LPA* goal first, D*
 *   start first (the 2D D* walk of op 1 leaves the goal out, d_star.py:117-127); path_len_dev 0 = no
 *   path, -2 = a capacity was exceeded; cost_dev as the reference sums it (inf when a D* back
 *   pointer's reverse edge is illegal under the directed corner rule). */
int pmp_lpa_star(const pmp_grid *g, int variant, int op, int heuristic, const int32_t *start_goal_dev,
                 const int32_t *change_dev, double *g_dev, double *rhs_dev, double *key_dev,
                 uint32_t *list_dev, int64_t list_cap, int64_t *ctl_dev, uint32_t *expand_dev,
                 int64_t expand_cap, uint32_t *path_dev, int64_t path_cap, int32_t *path_len_dev,
                 double *cost_dev, int32_t *n_expand_dev, void *stream);
int pmp_d_star(const pmp_grid *g, int op, const int32_t *start_goal_dev, double *h_dev, double *k_dev,
               uint32_t *parent_dev, uint8_t *tag_dev, uint32_t *list_dev, int64_t list_cap,
               int64_t *ctl_dev, uint32_t *expand_dev, int64_t expand_cap, uint32_t *path_dev,
               int64_t path_cap, int32_t *path_len_dev, double *cost_dev, int32_t *n_expand_dev,
               void *stream);

/* ---- slab mode: one z-slab (plus one ghost plane per interior side) of a large 3D grid per
 *      GPU; the cost-to-go field of BASELINE.json config 5.  The handle describes the LOCAL slab.
 *      Host loop (python_motion_planning_b200/slab.py): begin; repeat { run(bound); exchange
 *      boundary planes with the z+-1 ranks (NCCL send/recv); merge; all-reduce(min) of the
 *      pending keys -> next bound } until no key is pending anywhere. ---- */

/* Resets field + scratch; start_dev = local coordinates of the source if it lies in this slab,
 * else NULL. */
int pmp_slab_begin(pmp_grid *g, const pmp_sssp_params *params, const int32_t *start_dev,
                   void *g_dev, void *ws_dev, int64_t ws_bytes, void *stream);
/* Processes every pending key < key_bound to local convergence (+inf: everything).
 * out_dev: 8 x uint64; [0] = smallest key still pending here (f64: bit pattern; all-ones = none),
 * [1] rounds, [2] tile activations, [3] cells expanded; [4] is used by pmp_slab_merge. */
int pmp_slab_run(pmp_grid *g, const pmp_sssp_params *params, void *g_dev, void *ws_dev,
                 int64_t ws_bytes, double key_bound, uint64_t *out_dev, void *stream);
/* Element-wise min of a received plane (PX*PY values) into local plane z_local.  mark = 1:
 * improved cells become pending (their key is min-combined into out_dev[0], their count added
 * to out_dev[4]); mark = 0: values only (refreshing a ghost plane). */
int pmp_slab_merge(pmp_grid *g, int weight, void *g_dev, void *ws_dev, int64_t ws_bytes,
                   int z_local, int mark, const void *plane_dev, uint64_t *out_dev, void *stream);

/* Full unit-weight cost-to-go field WITHOUT any allocation (no counterpart in the reference; added
 * for callers that cannot call cudaMalloc: CUDA graph capture,
 * tight serving loops).  Same result as pmp_sssp(full_field = 1, weight = PMP_W_U32) on a 3D grid whose
 * motion table is the six unit moves, but the workspace lives in a static device array of 1 MiB inside
 * the library (held by every GPU that loads it): grids up to ~81 000 tiles of 32 x 8 x 8 cells (e.g.
 * 512^3); larger grids are rejected with an error.  start_dev: int32[3]; g_dev: pcells uint32 in the
 * search's word format ((value << 1) | 1, all-ones = unreached; pmp_field_finalize decodes);
 * status_dev (1 x int32) / stats_dev (16 x uint64) may be NULL.  All calls share the array: calls that
 * overlap on one GPU (two streams, two threads) race -- serialise them. */
int pmp_cost_field_u32_static(pmp_grid *g, const int32_t *start_dev, uint32_t *g_dev, int32_t *status_dev,
                              uint64_t *stats_dev, void *stream);

/* ---- z-slabs with device-side halo exchange (round 2): the unit-weight field of BASELINE.json
 *      config 5 (6-connected, cost 1: the field Dijkstra3D.plan computes with env.motions filtered
 *      to the unit moves, dijkstra3d.py:39-126) split over up to 8 GPUs of one node.  Rank p owns
 *      planes [z0, z1) as its LOCAL grid (no ghost planes; every slab but the last a multiple of
 *      8 planes).  Field + workspace live in one arena per rank that the other ranks map through
 *      CUDA IPC; the search kernel (csrc/pmp_bfs_core.cuh) reads the z halo of a boundary tile
 *      straight from the neighbour GPU's field over NVLink and arms / queues the neighbour's
 *      tiles with remote atomics.  No host round trip and no NCCL call inside a search: the host
 *      brackets one kernel launch per rank with two barriers.
 *      Protocol (python_motion_planning_b200/slab.py, BfsSlabSolver): create; export; all-gather
 *      the (handle, info) pairs (torch.distributed); connect; then per search: begin, barrier,
 *      run, barrier.  Ordering across slabs: the slabs emulate ONE queue (ring entries carry a global
 *      sequence number; a rank starts an entry only when the entries started on all ranks come within
 *      a window of it, PMP_B200_BFS_PACE, default 1024, 0 = off) -- without it a slab with a short queue
 *      runs ahead and recomputes its cells as the bulk of the wave arrives.  SURVEY.md 8(b) sketched "pmp_slab_* taking an ncclComm_t": the exchange
 *      itself needs no communicator, only the peers' mapped arenas. ---- */
typedef struct pmp_bfs_slab pmp_bfs_slab;
int pmp_bfs_slab_create(pmp_grid *g_local, int rank, int world, pmp_bfs_slab **out);
int pmp_bfs_slab_destroy(pmp_bfs_slab *s);
/* handle64: 64 bytes (cudaIpcMemHandle_t of the arena); info8: 8 x int64 layout words */
int pmp_bfs_slab_export(pmp_bfs_slab *s, void *handle64, int64_t *info8);
/* handles: world x 64 bytes, infos: world x 8 int64, both ordered by rank (own entry ignored) */
int pmp_bfs_slab_connect(pmp_bfs_slab *s, const void *handles, const int64_t *infos);
/* several slabs held by ONE process on one device (tests, or a grid cut for other reasons) */
int pmp_bfs_slab_connect_local(pmp_bfs_slab **slabs, int n);
/* reset + queue the source tile; start_local_dev = the source in local coordinates if it lies in
 * this slab, else NULL.  All ranks must complete this (stream sync + barrier) before any runs. */
int pmp_bfs_slab_begin(pmp_bfs_slab *s, const int32_t *start_local_dev, int32_t *status_dev, void *stream);
/* one launch: search to global termination, fill never-touched tiles, counters -> stats_dev (16 x
 * uint64, layout of pmp_sssp's stats) and PMP_ST_EXHAUSTED / PMP_ST_ERROR -> status_dev */
int pmp_bfs_slab_run(pmp_bfs_slab *s, int32_t *status_dev, uint64_t *stats_dev, void *stream);
/* connect_local'ed slabs: ONE launch serves all of them (synchronous) */
int pmp_bfs_slab_run_group(pmp_bfs_slab **slabs, int n, void *stream);
/* device pointers into the arena: the field (padded local box, (value << 1 | 1) words, all-ones =
 * unreached), its size, the 16 counters and the control block */
int pmp_bfs_slab_buffers(pmp_bfs_slab *s, void **field_dev, int64_t *field_bytes, void **stats_dev, void **ctr_dev);

/* ---- incremental replanning on a resident field.  Replaces the repair loop
 *      of the reference's incremental planners: LPAStar.OnPress / LPAStar3D.apply_change +
 *      computeShortestPath + updateVertex (global_planner/graph_search/lpa_star.py:102-177,
 *      lpa_star3d.py:98-177) and DStar3D.apply_dynamic_obstacles (d_star3d.py:115-151).  One
 *      query; the field g_dev and the workspace ws_dev stay on the device between calls and
 *      must not be touched by other searches. ---- */

/* First plan: pmp_sssp for one query, run without the g[goal] pruning so that the field can be
 * repaired later.  Same outputs as pmp_sssp. */
int pmp_replan_begin(pmp_grid *g, const pmp_sssp_params *params, const int32_t *start_dev,
                     const int32_t *goal_dev, void *g_dev, void *ws_dev, int64_t ws_bytes,
                     double *gcost_dev, int32_t *status_dev, uint64_t *stats_dev, void *stream);
/* Applies n_changes occupancy changes (cells_dev: [n][3] int32 box coordinates, one entry per
 * cell; blocked_dev[i] = 1 obstacle added, 0 obstacle removed) to the handle's bitmask, resets
 * every value that lost its support, re-arms the cells around the change and resumes the search;
 * afterwards the field equals the field of a search from scratch on the new map.
 * scratch_dev: 16 control words + two candidate buffers of (scratch_words - 16) / 2 cells each;
 * after the call [3] = overflow flag, [4] = cells reset, [5] = invalidation waves.  If the
 * invalidated region does not fit, nothing is searched, status = PMP_ST_REPLAN_OVERFLOW and the
 * caller falls back to pmp_replan_begin (the bitmask is already up to date). */
int pmp_replan_update(pmp_grid *g, const pmp_sssp_params *params, const int32_t *start_dev,
                      const int32_t *goal_dev, void *g_dev, void *ws_dev, int64_t ws_bytes,
                      const int32_t *cells_dev, const uint8_t *blocked_dev, int64_t n_changes,
                      uint32_t *scratch_dev, int64_t scratch_words,
                      double *gcost_dev, int32_t *status_dev, uint64_t *stats_dev, void *stream);

/* ---- line of sight and any-angle shortcutting ---- */

/* ThetaStar.lineOfSight (global_planner/graph_search/theta_star.py:110-171; 2D handles) and
 * ThetaStar3D.lineOfSight / LazyThetaStar3D.lineOfSight (theta_star3d.py:139-215,
 * lazy_theta_star3d.py:158-235; 3D handles) for n ordered pairs a[i] -> b[i] ([n][3] int32 box
 * coordinates): out_dev[i] = 1 iff the reference returns True.  Cells outside the box are blocked. */
int pmp_line_of_sight(const pmp_grid *g, int64_t n, const int32_t *a_dev, const int32_t *b_dev,
                      uint8_t *out_dev, void *stream);
/* Any-angle shortcut of the paths written by pmp_extract_paths (same layout, goal first): from
 * each kept waypoint the grid path is followed while the next cell is still visible from it
 * (the "path 2" test of ThetaStar.updateVertex, theta_star.py:96-108, applied to the extracted
 * path).  out_path_dev: [nq][path_cap] kept waypoints (padded cell indices), out_len_dev their
 * number (0: no path; -1: input length > path_cap), out_cost_dev the summed Euclidean lengths
 * (Planner.dist, accumulated goal -> start). */
int pmp_shortcut_paths(const pmp_grid *g, int64_t nq, const uint32_t *path_dev, int64_t path_cap,
                       const int32_t *path_len_dev, uint32_t *out_path_dev, int32_t *out_len_dev,
                       double *out_cost_dev, void *stream);

#ifdef __cplusplus
}
#endif
#endif
