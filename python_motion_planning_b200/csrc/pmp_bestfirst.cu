// pmp_bestfirst.cu -- the reference's best-first planners in the reference's own pop order.
//
// pmp_sssp (pmp_sssp.cu) settles the same costs as the reference with a tiled wavefront, but which
// of several equal-cost paths comes back, and the content and order of `expand`, follow from the
// reference's heap discipline.  pmp_bestfirst runs that discipline as written --
// AStar / Dijkstra / GBFS / ThetaStar / LazyThetaStar / JPS .plan() and their 3D classes
// (global_planner/graph_search/a_star.py:39-117, dijkstra.py:36-85, gbfs.py:36-86,
// theta_star.py:44-108, lazy_theta_star.py:44-118, a_star3d.py:33-106, dijkstra3d.py:39-143,
// gbfs3d.py:34-110, theta_star3d.py:38-107, lazy_theta_star3d.py:41-133, jps.py:38-165,
// jps3d.py:42-246) -- so cost, path,
// len(expand) and every expand record equal the reference's.  One warp per query: lane 0 owns the
// OPEN heap and the CLOSED table (pmp_bestfirst_core.h), the lanes evaluate the motions of the
// popped node in parallel (edge legality, CLOSED dominance, heuristic, Theta* line of sight, a whole
// jump of JPS per lane) and
// lane 0 pushes the survivors in motion order.  Independent queries run in parallel warps.
#include "pmp_common.cuh"
#include "pmp_los.cuh"
#include "pmp_bestfirst_core.h"
#include "pmp_jps_core.h"
#include "pmp_incr_core.h"

struct BfEnvDev {
    const GridDev &G;
    const Motions &M;
    __device__ bool legal(int m, int x, int y, int z) const { return edge_legal(G, M, m, x, y, z); }
    __device__ bool los(int ax, int ay, int az, int bx, int by, int bz) const {
        return ::los(G, ax, ay, az, bx, by, bz);
    }
    __device__ bool inbox(int x, int y, int z) const {
        return x >= 0 && y >= 0 && z >= 0 && x < G.X && y < G.Y && z < G.Z;
    }
    __device__ bool obst(int x, int y, int z) const { return inbox(x, y, z) && cell_blocked(G, x, y, z); }
    // not isCollision((x,y,z), (x,y,z) + d) for any unit offset d (graph_search_3d.py:66-107)
    __device__ bool legal_dir(int dx, int dy, int dz, int x, int y, int z) const {
        if (cell_blocked(G, x, y, z) || cell_blocked(G, x + dx, y + dy, z + dz)) return false;
        if ((dx != 0) + (dy != 0) + (dz != 0) >= 2) {
            if (dx && cell_blocked(G, x + dx, y, z)) return false;
            if (dy && cell_blocked(G, x, y + dy, z)) return false;
            if (dz && cell_blocked(G, x, y, z + dz)) return false;
        }
        return true;
    }
    __device__ int k() const { return M.k; }
    __device__ void offset(int m, int &dx, int &dy, int &dz) const { dx = M.dx[m]; dy = M.dy[m]; dz = M.dz[m]; }
    __device__ double len(int m) const { return M.len[m]; }
};

template <int FAMILY>
__global__ void k_bestfirst(const GridDev G, const __grid_constant__ Motions M, int algo, int heuristic,
                            u32 nq, const int32_t *starts, const int32_t *goals, BfRec *heap_all,
                            int64_t heap_cap, double *cg_all, double *ch_all, u32 *cpar_all,
                            u32 *order_all, int64_t order_cap, u32 *path, int64_t cap, int32_t *plen,
                            double *cost, int32_t *n_expand) {
    const u32 q = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    const int lane = threadIdx.x & 31;
    if (q >= nq) return;
    const unsigned FULL = 0xFFFFFFFFu;
    const double INF_D = __longlong_as_double(0x7ff0000000000000LL);
    const int sx = starts[3 * q], sy = starts[3 * q + 1], sz = G.ndim == 3 ? starts[3 * q + 2] : 0;
    const int gx = goals[3 * q], gy = goals[3 * q + 1], gz = G.ndim == 3 ? goals[3 * q + 2] : 0;
    const bool ok = gx >= 0 && gy >= 0 && gz >= 0 && gx < G.X && gy < G.Y && gz < G.Z &&
                    sx >= 0 && sy >= 0 && sz >= 0 && sx < G.X && sy < G.Y && sz < G.Z;
    if (!ok) { // outside the box; a blocked start or goal is searched as the reference does (CLOSED = {start} / exhaustion)
        if (lane == 0) { plen[q] = 0; cost[q] = INF_D; n_expand[q] = 0; }
        return;
    }
    // every lane holds the constant part of the state; hn / counter / n_order / status are
    // lane 0's, the others only read CLOSED (global memory, ordered by __syncwarp)
    BfState S;
    S.heap = heap_all + (size_t)q * heap_cap;
    S.heap_cap = heap_cap;
    S.cg = cg_all + (size_t)q * G.pcells;
    S.ch = ch_all + (size_t)q * G.pcells;
    S.cpar = cpar_all + (size_t)q * G.pcells;
    S.order = order_all + (size_t)q * order_cap;
    S.order_cap = order_cap;
    S.algo = algo;
    S.heuristic = heuristic;
    S.PX = G.PX;
    S.PY = G.PY;
    S.gx = gx; S.gy = gy; S.gz = gz;
    S.start_cell = bf_cell(S, sx, sy, sz);
    S.goal_cell = bf_cell(S, gx, gy, gz);
    S.hn = 0; S.counter = 0; S.n_order = 0; S.status = BF_RUNNING;
    const BfEnvDev env = {G, M};
    if (lane == 0) bf_start<FAMILY>(S, sx, sy, sz);
    const double myw = lane < M.k ? (FAMILY == 2 ? M.w[lane] : M.len[lane]) : 0.0;

    for (;;) {
        BfRec node;
        node.f = node.h = node.g = 0.0;
        node.cell = node.parent = node.counter = node.pad = 0;
        int action = BF_DONE;
        if (lane == 0) action = bf_pop<FAMILY>(S, env, node);
        action = __shfl_sync(FULL, action, 0);
        if (action == BF_DONE) break;
        if (action == BF_SKIP) continue;
        node.g = __shfl_sync(FULL, node.g, 0);
        node.h = __shfl_sync(FULL, node.h, 0);
        node.cell = __shfl_sync(FULL, node.cell, 0);
        node.parent = __shfl_sync(FULL, node.parent, 0);
        __syncwarp(); // lane 0's CLOSED writes are visible to the evaluating lanes
        BfRec cand;
        cand.f = cand.h = cand.g = 0.0;
        cand.cell = cand.parent = cand.counter = cand.pad = 0;
        bool push = false;
        if (lane < M.k) {
            if (algo == BF_JPS)
                push = FAMILY == 2 ? jps2d_eval(S, env, node, lane, myw, cand) : jps3d_eval(S, env, node, lane, cand);
            else
                push = bf_eval<FAMILY>(S, env, node, lane, myw, cand);
        }
        const u32 mask = __ballot_sync(FULL, push);
        int stop = 0;
        for (int m = 0; m < M.k && !stop; ++m) { // lane 0 pushes in motion order
            if (!((mask >> m) & 1u)) continue;
            BfRec r;
            r.f = __shfl_sync(FULL, cand.f, m);
            r.h = __shfl_sync(FULL, cand.h, m);
            r.g = __shfl_sync(FULL, cand.g, m);
            r.cell = __shfl_sync(FULL, cand.cell, m);
            r.parent = __shfl_sync(FULL, cand.parent, m);
            r.counter = 0; r.pad = 0;
            if (lane == 0) stop = bf_push<FAMILY>(S, r) ? 1 : 0;
            stop = __shfl_sync(FULL, stop, 0);
        }
        if (lane == 0 && FAMILY == 2) bf_close(S, node); // a_star.py:82
        const int st = __shfl_sync(FULL, S.status, 0);
        if (st != BF_RUNNING) break;
        __syncwarp();
    }
    if (lane != 0) return;
    n_expand[q] = (int32_t)S.n_order;
    if (S.status != BF_REACHED) {
        plen[q] = S.status == BF_OVERFLOW ? -2 : 0;
        cost[q] = INF_D;
        return;
    }
    bf_extract(S, path + (size_t)q * cap, cap, &plen[q], &cost[q]);
}

extern "C" int pmp_bestfirst(const pmp_grid *g, int algo, int heuristic, int64_t nq,
                             const int32_t *starts_dev, const int32_t *goals_dev,
                             void *heap_dev, int64_t heap_cap, double *closed_g_dev,
                             double *closed_h_dev, uint32_t *closed_parent_dev, uint32_t *order_dev,
                             int64_t order_cap, uint32_t *path_dev, int64_t path_cap,
                             int32_t *path_len_dev, double *cost_dev, int32_t *n_expand_dev,
                             void *stream) {
    if (g && pmp_grid_upload_fence((pmp_grid *)g, (cudaStream_t)stream)) return -2; // a streamed occupancy upload must have landed
    if (!g || nq < 1 || !starts_dev || !goals_dev || !heap_dev || heap_cap < 1 || !closed_g_dev ||
        !closed_h_dev || !closed_parent_dev || !order_dev || order_cap < 1 || !path_dev ||
        path_cap < 1 || !path_len_dev || !cost_dev || !n_expand_dev) {
        pmp_set_error("pmp_bestfirst: bad arguments");
        return -1;
    }
    if (algo < PMP_ASTAR || algo > PMP_JPS) {
        pmp_set_error("pmp_bestfirst: unknown algorithm %d", algo);
        return -1;
    }
    if (g->d.pcells >= 0xFFFFFFFFull) {
        pmp_set_error("pmp_bestfirst: grid too large for 32-bit cell indices");
        return -1;
    }
    static_assert(sizeof(BfRec) == PMP_BESTFIRST_HEAP_REC_BYTES, "heap record size is part of the ABI");
    cudaStream_t s = (cudaStream_t)stream;
    PMP_CHECK_CUDA(cudaMemsetAsync(closed_parent_dev, 0xFF, (size_t)nq * g->d.pcells * 4, s));
    const unsigned blocks = (unsigned)((nq * 32 + 63) / 64);
    if (g->d.ndim == 2)
        k_bestfirst<2><<<blocks, 64, 0, s>>>(g->d, g->mot, algo, heuristic, (u32)nq, starts_dev, goals_dev,
                                             (BfRec *)heap_dev, heap_cap, closed_g_dev, closed_h_dev,
                                             closed_parent_dev, order_dev, order_cap, path_dev, path_cap,
                                             path_len_dev, cost_dev, n_expand_dev);
    else
        k_bestfirst<3><<<blocks, 64, 0, s>>>(g->d, g->mot, algo, heuristic, (u32)nq, starts_dev, goals_dev,
                                             (BfRec *)heap_dev, heap_cap, closed_g_dev, closed_h_dev,
                                             closed_parent_dev, order_dev, order_cap, path_dev, path_cap,
                                             path_len_dev, cost_dev, n_expand_dev);
    PMP_CHECK_CUDA(cudaGetLastError());
    return 0;
}

// ---------------------------------------------------------------------------------------------
// LPA* / D* in the reference's processing order (pmp_incr_core.h).  Sequential algorithms on Python
// lists: thread 0 of one CTA runs them, the CTA's other threads only initialise the per-cell state.
// ---------------------------------------------------------------------------------------------
// Environment of the incremental planners.  DStarLite.OnPress edits the obstacle set in the middle of
// its walk (d_star_lite.py:82-90): toggle() flips the bit from inside the kernel, and that one entry
// point reads the bitmask with volatile loads (at L2) instead of the read-only path; everything else
// keeps the cached loads (volatile everywhere made LPA* 2.2x slower).
struct IncrEnvDev {
    const GridDev &G;
    const Motions &M;
    bool coherent; // true: the bitmask may change under this kernel (D* Lite's OnPress) -- read it at L2
    __device__ bool inbox(int x, int y, int z) const {
        return x >= 0 && y >= 0 && z >= 0 && x < G.X && y < G.Y && z < G.Z;
    }
    __device__ bool blocked(int x, int y, int z) const {
        if (x < 0 || y < 0 || z < 0 || x >= G.PX || y >= G.PY || z >= G.PZ) return true;
        const u32 *w = G.bits + ((size_t)z * G.PY + y) * G.wpr + (x >> 5);
        const u32 v = coherent ? *(const volatile u32 *)w : __ldg(w);
        return (v >> (x & 31)) & 1u;
    }
    __device__ bool obst(int x, int y, int z) const { return inbox(x, y, z) && blocked(x, y, z); }
    __device__ bool legal_dir(int dx, int dy, int dz, int x, int y, int z) const {
        if (blocked(x, y, z) || blocked(x + dx, y + dy, z + dz)) return false;
        if ((dx != 0) + (dy != 0) + (dz != 0) >= 2) {
            if (dx && blocked(x + dx, y, z)) return false;
            if (dy && blocked(x, y + dy, z)) return false;
            if (dz && blocked(x, y, z + dz)) return false;
        }
        return true;
    }
    __device__ bool toggle(int x, int y, int z) const { // true: an obstacle was removed
        u32 *w = const_cast<u32 *>(G.bits) + ((size_t)z * G.PY + y) * G.wpr + (x >> 5);
        const u32 old = atomicXor(w, 1u << (x & 31));
        __threadfence();
        return (old >> (x & 31)) & 1u;
    }
    __device__ int k() const { return M.k; }
    __device__ void offset(int m, int &dx, int &dy, int &dz) const { dx = M.dx[m]; dy = M.dy[m]; dz = M.dz[m]; }
    __device__ double len(int m) const { return M.len[m]; }
};

__device__ static void incr_geometry(BfState &S, const GridDev &G, int heuristic, const int32_t *sg) {
    S.heap = nullptr; S.hn = 0; S.heap_cap = 0;
    S.cg = S.ch = nullptr; S.cpar = nullptr; S.order = nullptr; S.n_order = 0; S.order_cap = 0;
    S.counter = 0; S.algo = 0; S.heuristic = heuristic;
    S.PX = G.PX; S.PY = G.PY;
    S.gx = sg[3]; S.gy = sg[4]; S.gz = G.ndim == 3 ? sg[5] : 0;
    S.start_cell = bf_cell(S, sg[0], sg[1], G.ndim == 3 ? sg[2] : 0);
    S.goal_cell = bf_cell(S, S.gx, S.gy, S.gz);
    S.status = BF_RUNNING;
}

__device__ static bool incr_query_ok(const GridDev &G, const int32_t *sg) {
    const int sz = G.ndim == 3 ? sg[2] : 0, gz = G.ndim == 3 ? sg[5] : 0;
    return sg[0] >= 0 && sg[1] >= 0 && sz >= 0 && sg[0] < G.X && sg[1] < G.Y && sz < G.Z &&
           sg[3] >= 0 && sg[4] >= 0 && gz >= 0 && sg[3] < G.X && sg[4] < G.Y && gz < G.Z;
}

__global__ void k_lpa_star(const GridDev G, const __grid_constant__ Motions M, int variant, int op, int heuristic,
                           const int32_t *sg, const int32_t *change, double *g, double *rhs, double *key,
                           u32 *list, int64_t list_cap, int64_t *ctl, u32 *expand, int64_t expand_cap,
                           u32 *path, int64_t cap, int32_t *plen, double *cost, int32_t *n_expand) {
    const double INF_D = __longlong_as_double(0x7ff0000000000000LL);
    if (op == 0) {
        for (size_t i = threadIdx.x; i < G.pcells; i += blockDim.x) { g[i] = INF_D; rhs[i] = INF_D; }
        __syncthreads();
    }
    if (threadIdx.x != 0) return;
    if (!incr_query_ok(G, sg)) { *plen = 0; *cost = INF_D; *n_expand = 0; return; }
    LpaState L;
    L.g = g; L.rhs = rhs; L.k1 = key; L.k2 = key + G.pcells;
    L.U = list; L.ucap = list_cap; L.un = op == 0 ? 0 : ctl[0];
    L.expand = expand; L.expand_cap = expand_cap; L.n_expand = 0;
    L.family = G.ndim;
    L.status = BF_RUNNING;
    L.dlite = variant == 1;
    L.km = op == 0 ? 0.0 : __longlong_as_double(ctl[2]);
    incr_geometry(L.S, G, heuristic, sg);
    IncrEnvDev env = {G, M, variant == 1 && op == 1};
    if (L.dlite) {
        if (op == 0) {
            lpa_init(L);
            lpa_compute(L, env);
            if (L.status != BF_OVERFLOW) dlite_extract(L, env, path, cap, plen, cost);
        } else {
            dlite_press(L, env, change[0], change[1], G.ndim == 3 ? change[2] : 0, path, cap, plen, cost);
        }
    } else {
        if (op == 0) lpa_init(L);
        else lpa_apply_change(L, env, change[0], change[1], G.ndim == 3 ? change[2] : 0, change[3] != 0);
        lpa_compute(L, env);
        if (L.status != BF_OVERFLOW) lpa_extract(L, env, path, cap, plen, cost);
    }
    ctl[0] = L.un;
    ctl[1] = L.status;
    ctl[2] = __double_as_longlong(L.km);
    *n_expand = (int32_t)L.n_expand;
    if (L.status == BF_OVERFLOW) { *plen = -2; *cost = INF_D; }
}

__global__ void k_d_star(const GridDev G, const __grid_constant__ Motions M, int op, const int32_t *sg,
                         double *h, double *k, u32 *parent, uint8_t *tag, u32 *list, int64_t list_cap,
                         int64_t *ctl, u32 *expand, int64_t expand_cap, u32 *path, int64_t cap,
                         int32_t *plen, double *cost, int32_t *n_expand) {
    const double INF_D = __longlong_as_double(0x7ff0000000000000LL);
    if (op == 0) {
        for (size_t i = threadIdx.x; i < G.pcells; i += blockDim.x) {
            h[i] = INF_D; k[i] = INF_D; parent[i] = BF_NONE; tag[i] = DS_NEW;
        }
        __syncthreads();
    }
    if (threadIdx.x != 0) return;
    if (!incr_query_ok(G, sg)) { *plen = 0; *cost = INF_D; *n_expand = 0; return; }
    DsState D;
    D.h = h; D.k = k; D.parent = parent; D.t = tag;
    D.OPEN = list; D.ocap = list_cap; D.on = op == 0 ? 0 : ctl[0];
    D.expand = expand; D.expand_cap = expand_cap; D.n_expand = 0;
    D.family = G.ndim;
    D.status = BF_RUNNING;
    incr_geometry(D.S, G, PMP_H_EUCLIDEAN, sg);
    const IncrEnvDev env = {G, M, false};
    if (op == 0) {
        ds_init(D);
        ds_plan(D, env);
        if (D.status != BF_OVERFLOW) ds_extract(D, env, path, cap, plen, cost);
    } else {
        ds_replan(D, env, path, cap, plen, cost);
    }
    ctl[0] = D.on;
    ctl[1] = D.status;
    *n_expand = (int32_t)D.n_expand;
    if (D.status == BF_OVERFLOW) { *plen = -2; *cost = INF_D; }
}

extern "C" int pmp_lpa_star(const pmp_grid *g, int variant, int op, int heuristic, const int32_t *start_goal_dev,
                            const int32_t *change_dev, double *g_dev, double *rhs_dev, double *key_dev,
                            uint32_t *list_dev, int64_t list_cap, int64_t *ctl_dev, uint32_t *expand_dev,
                            int64_t expand_cap, uint32_t *path_dev, int64_t path_cap, int32_t *path_len_dev,
                            double *cost_dev, int32_t *n_expand_dev, void *stream) {
    if (g && pmp_grid_upload_fence((pmp_grid *)g, (cudaStream_t)stream)) return -2; // a streamed occupancy upload must have landed
    if (!g || (variant != 0 && variant != 1) || (op != 0 && op != 1) || !start_goal_dev || (op == 1 && !change_dev) || !g_dev || !rhs_dev ||
        !key_dev || !list_dev || list_cap < 1 || !ctl_dev || !expand_dev || expand_cap < 1 || !path_dev ||
        path_cap < 1 || !path_len_dev || !cost_dev || !n_expand_dev) {
        pmp_set_error("pmp_lpa_star: bad arguments");
        return -1;
    }
    if (g->d.pcells >= 0xFFFFFFFFull) { pmp_set_error("pmp_lpa_star: grid too large for 32-bit cell indices"); return -1; }
    if (variant == 1 && g->d.ndim != 2) { pmp_set_error("pmp_lpa_star: D* Lite is a 2D planner (d_star_lite.py)"); return -1; }
    k_lpa_star<<<1, 256, 0, (cudaStream_t)stream>>>(g->d, g->mot, variant, op, heuristic, start_goal_dev, change_dev, g_dev,
                                                    rhs_dev, key_dev, list_dev, list_cap, ctl_dev, expand_dev,
                                                    expand_cap, path_dev, path_cap, path_len_dev, cost_dev,
                                                    n_expand_dev);
    PMP_CHECK_CUDA(cudaGetLastError());
    return 0;
}

extern "C" int pmp_d_star(const pmp_grid *g, int op, const int32_t *start_goal_dev, double *h_dev, double *k_dev,
                          uint32_t *parent_dev, uint8_t *tag_dev, uint32_t *list_dev, int64_t list_cap,
                          int64_t *ctl_dev, uint32_t *expand_dev, int64_t expand_cap, uint32_t *path_dev,
                          int64_t path_cap, int32_t *path_len_dev, double *cost_dev, int32_t *n_expand_dev,
                          void *stream) {
    if (g && pmp_grid_upload_fence((pmp_grid *)g, (cudaStream_t)stream)) return -2; // a streamed occupancy upload must have landed
    if (!g || (op != 0 && op != 1) || !start_goal_dev || !h_dev || !k_dev || !parent_dev || !tag_dev ||
        !list_dev || list_cap < 1 || !ctl_dev || !expand_dev || expand_cap < 1 || !path_dev || path_cap < 1 ||
        !path_len_dev || !cost_dev || !n_expand_dev) {
        pmp_set_error("pmp_d_star: bad arguments");
        return -1;
    }
    if (g->d.pcells >= 0xFFFFFFFFull) { pmp_set_error("pmp_d_star: grid too large for 32-bit cell indices"); return -1; }
    k_d_star<<<1, 256, 0, (cudaStream_t)stream>>>(g->d, g->mot, op, start_goal_dev, h_dev, k_dev, parent_dev, tag_dev,
                                                  list_dev, list_cap, ctl_dev, expand_dev, expand_cap, path_dev,
                                                  path_cap, path_len_dev, cost_dev, n_expand_dev);
    PMP_CHECK_CUDA(cudaGetLastError());
    return 0;
}
