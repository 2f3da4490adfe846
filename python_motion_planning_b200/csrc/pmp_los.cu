// pmp_los.cu -- line of sight on the occupancy bitmask and any-angle shortcutting of extracted
// paths.
//
// pmp_line_of_sight replaces ThetaStar.lineOfSight (global_planner/graph_search/
// theta_star.py:110-171) and ThetaStar3D.lineOfSight / LazyThetaStar3D.lineOfSight
// (theta_star3d.py:139-215, lazy_theta_star3d.py:158-235): the same Bresenham walks, one thread
// per ordered pair, obstacle tests as bit reads.  pmp_shortcut_paths applies them to the paths
// written by pmp_extract_paths: from each kept waypoint the walk along the grid path continues
// while the next cell is still visible (the "path 2" test of updateVertex, theta_star.py:96-108,
// applied after the search instead of inside it), one warp per query, 32 candidates per step.
#include "pmp_common.cuh"
#include "pmp_los.cuh"

__global__ void k_line_of_sight(const GridDev G, size_t n, const int32_t *a, const int32_t *b, uint8_t *out) {
    for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x) {
        const int32_t *p = a + 3 * i, *q = b + 3 * i;
        out[i] = los(G, p[0], p[1], G.ndim == 3 ? p[2] : 0, q[0], q[1], G.ndim == 3 ? q[2] : 0) ? 1 : 0;
    }
}

extern "C" int pmp_line_of_sight(const pmp_grid *g, int64_t n, const int32_t *a_dev, const int32_t *b_dev,
                                 uint8_t *out_dev, void *stream) {
    if (g && pmp_grid_upload_fence((pmp_grid *)g, (cudaStream_t)stream)) return -2; // a streamed occupancy upload must have landed
    if (!g || n < 0 || (n && (!a_dev || !b_dev || !out_dev))) {
        pmp_set_error("pmp_line_of_sight: bad arguments");
        return -1;
    }
    if (n == 0) return 0;
    size_t blocks = ((size_t)n + 127) / 128;
    if (blocks > 148 * 16) blocks = 148 * 16;
    k_line_of_sight<<<(unsigned)blocks, 128, 0, (cudaStream_t)stream>>>(g->d, (size_t)n, a_dev, b_dev, out_dev);
    PMP_CHECK_CUDA(cudaGetLastError());
    return 0;
}

// one warp per query; path = padded cell indices in the order pmp_extract_paths wrote them
// (goal first), so a kept waypoint plays the child and the candidate the ancestor, as in
// updateVertex(node_p, node_c) -> lineOfSight(node_c, node_p) (theta_star.py:103).
__global__ void k_shortcut_paths(const GridDev G, u32 nq, const u32 *path, int64_t cap, const int32_t *plen,
                                 u32 *out_path, int32_t *out_len, double *out_cost) {
    const u32 q = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    const int lane = threadIdx.x & 31;
    if (q >= nq) return;
    const u32 *p = path + (size_t)q * cap;
    u32 *o = out_path + (size_t)q * cap;
    const int n = plen[q];
    if (n <= 0 || n > cap) {
        if (lane == 0) { out_len[q] = n <= 0 ? 0 : -1; out_cost[q] = __longlong_as_double(0x7ff0000000000000LL); }
        return;
    }
    const size_t plane = (size_t)G.PX * G.PY;
    int i = 0, m = 0;
    double c = 0.0;
    if (lane == 0) o[m] = p[0];
    ++m;
    while (i < n - 1) {
        const u32 ci = p[i];
        const int ax = (int)(ci % G.PX), ay = (int)((ci / G.PX) % G.PY), az = (int)(ci / plane);
        int j = i + 1; // the next grid cell is always visible (one legal step)
        for (;;) {
            const int cand = j + 1 + lane;
            bool ok = false;
            if (cand < n) {
                const u32 cc = p[cand];
                ok = los(G, ax, ay, az, (int)(cc % G.PX), (int)((cc / G.PX) % G.PY), (int)(cc / plane));
            }
            const u32 fail = ~__ballot_sync(0xFFFFFFFFu, ok);
            const int adv = fail ? __ffs(fail) - 1 : 32;
            j += adv;
            if (adv < 32) break;
        }
        const u32 cj = p[j];
        const int dx = (int)(cj % G.PX) - ax, dy = (int)((cj / G.PX) % G.PY) - ay, dz = (int)(cj / plane) - az;
        // Planner.dist / Planner3D.dist (planner.py:22-23, planner3d.py:22-27): exact integer sum, IEEE sqrt
        c += sqrt((double)((long long)dx * dx + (long long)dy * dy + (long long)dz * dz));
        if (lane == 0) o[m] = cj;
        ++m;
        i = j;
    }
    if (lane == 0) { out_len[q] = m; out_cost[q] = c; }
}

extern "C" int pmp_shortcut_paths(const pmp_grid *g, int64_t nq, const uint32_t *path_dev, int64_t path_cap,
                                  const int32_t *path_len_dev, uint32_t *out_path_dev, int32_t *out_len_dev,
                                  double *out_cost_dev, void *stream) {
    if (g && pmp_grid_upload_fence((pmp_grid *)g, (cudaStream_t)stream)) return -2; // a streamed occupancy upload must have landed
    if (!g || nq < 1 || !path_dev || path_cap < 1 || !path_len_dev || !out_path_dev || !out_len_dev || !out_cost_dev) {
        pmp_set_error("pmp_shortcut_paths: bad arguments");
        return -1;
    }
    const unsigned blocks = (unsigned)((nq * 32 + 127) / 128);
    k_shortcut_paths<<<blocks, 128, 0, (cudaStream_t)stream>>>(g->d, (u32)nq, path_dev, path_cap, path_len_dev,
                                                                out_path_dev, out_len_dev, out_cost_dev);
    PMP_CHECK_CUDA(cudaGetLastError());
    return 0;
}
