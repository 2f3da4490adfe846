// pmp_bfs.cu -- kernels + C ABI of the unit-weight cost-to-go field solver (one warp per tile,
// pmp_bfs_core.cuh has the algorithm) and of its multi-GPU z-slab form with device-side halo pulls
// over peer memory.
//
// Replaces Dijkstra3D.plan / getNeighbor (global_planner/graph_search/dijkstra3d.py:39-126) run
// to exhaustion on the six unit moves of env3d.py:56-70 -- BASELINE.json config 5.
#include <stdlib.h>
#include <string.h>

#include "pmp_common.cuh"
#include "pmp_bfs_core.cuh"

#ifndef BFS_NT
#define BFS_NT 128 // threads per CTA: four independent warps, no shared memory, no CTA barrier
#endif
#ifndef BFS_MINB
#define BFS_MINB 4 // resident CTAs per SM the register budget is sized for
#endif
#define BFS_RING_SLACK 8192

__global__ void __launch_bounds__(BFS_NT, BFS_MINB) k_bfs(const __grid_constant__ BfsArgs A) {
    bfs_warp_main(A, (int)(threadIdx.x & 31));
    if ((threadIdx.x & 31) == 0 && A.stats) atomicMax((unsigned long long *)&A.stats[10], (unsigned long long)bfs_now_ns());
}

// One thread: validate the source, queue its tile.  `start` = local coordinates (x, y, z) of the
// source inside this box, or nullptr when the source lies in another slab.
__global__ void k_bfs_init(BfsArgs A, const int32_t *start, int32_t *status) {
    const GridDev &Gd = A.G;
    A.stats[9] = bfs_now_ns();
    A.stats[0] = 1;
    if (A.rank == 0) A.ctr[BFS_C_ACTIVE] = 1u;
    int st = PMP_ST_RUNNING;
    if (start) {
        const int sx = start[0], sy = start[1], sz = start[2];
        const bool ok = sx >= 0 && sy >= 0 && sz >= 0 && sx < Gd.X && sy < Gd.Y && sz < Gd.Z && !cell_blocked(Gd, sx, sy, sz);
        if (ok) {
            const u32 t = (u32)(((sz / 8) * A.nby + sy / 8) * A.nbx + sx / 32);
            A.state[t] = BFS_SRC | BFS_Q;
            A.ring[0] = t + 1u;
            A.ctr[BFS_C_TAIL] = 1u;
            A.ctr[BFS_C_INFLIGHT] = 1u;
            A.ctr[BFS_C_SRC] = (u32)sx; A.ctr[BFS_C_SRC + 1] = (u32)sy; A.ctr[BFS_C_SRC + 2] = (u32)sz;
        } else {
            st = PMP_ST_BAD_QUERY;
            for (int r = 0; r < A.world; ++r) A.peer_ctr[r][BFS_C_DONE] = 1u; // nothing to search
        }
    }
    if (status) *status = st;
}

// untouched tiles -> all-ones; one warp per tile
__global__ void k_bfs_fill(BfsArgs A) {
    const int lane = (int)(threadIdx.x & 31);
    const u32 warp = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, nwarp = (gridDim.x * blockDim.x) >> 5;
    for (u32 t = warp; t < A.ntiles; t += nwarp)
        if (!(A.state[t] & BFS_T)) bfs_fill_tile(A, t, lane);
}

__global__ void k_bfs_finish(BfsArgs A, double *gcost, int32_t *status, u64 *stats_out) {
    A.stats[8] = A.stats[10] > A.stats[9] ? A.stats[10] - A.stats[9] : 0;
    A.stats[15] = A.ctr[BFS_C_ERR];
    if (gcost) *gcost = __longlong_as_double(0x7ff0000000000000LL); // a full field has no goal
    if (status && *status != PMP_ST_BAD_QUERY) *status = A.ctr[BFS_C_ERR] ? PMP_ST_ERROR : PMP_ST_EXHAUSTED;
    if (stats_out)
        for (int i = 0; i < 16; ++i) stats_out[i] = A.stats[i];
}

// ---------------------------------------------------------------------------------------------
// host side
// ---------------------------------------------------------------------------------------------
struct BfsLayout {
    size_t state, ring, ctr, stats, total;
    u32 ntiles, ring_cap;
    int nbx, nby, nbz;
};

static size_t bfs_align(size_t v) { return (v + 255) / 256 * 256; }

static BfsLayout bfs_layout(const pmp_grid *g) {
    BfsLayout L;
    L.nbx = g->d.wpr; L.nby = g->d.PY / 8; L.nbz = g->d.PZ / 8;
    L.ntiles = (u32)((size_t)L.nbx * L.nby * L.nbz);
    L.ring_cap = L.ntiles + BFS_RING_SLACK;
    size_t o = 0;
    L.state = o; o = bfs_align(o + (size_t)L.ntiles * 4);
    L.ring = o; o = bfs_align(o + (size_t)L.ring_cap * 4);
    L.ctr = o; o = bfs_align(o + BFS_C_WORDS * 4);
    L.stats = o; o = bfs_align(o + 16 * 8);
    L.total = o;
    return L;
}

// the motion table is exactly the six unit axis moves?
bool pmp_bfs_applicable(const pmp_grid *g) {
    if (!g || g->d.ndim != 3 || g->mot.k != 6 || g->d.PY % 8 || g->d.PZ % 8) return false;
    u32 seen = 0;
    for (int m = 0; m < 6; ++m) {
        const int dx = g->mot.dx[m], dy = g->mot.dy[m], dz = g->mot.dz[m];
        if ((dx != 0) + (dy != 0) + (dz != 0) != 1 || g->mot.w[m] != 1.0) return false;
        seen |= 1u << ((dx ? 0 : (dy ? 2 : 4)) + ((dx + dy + dz) > 0 ? 1 : 0));
    }
    return seen == 0x3Fu;
}

size_t pmp_bfs_ws_bytes(const pmp_grid *g) { return bfs_layout(g).total; }

static void bfs_fill_args(const pmp_grid *g, void *g_dev, void *ws_dev, const BfsLayout &L, BfsArgs &A) {
    char *ws = (char *)ws_dev;
    memset(&A, 0, sizeof A);
    A.G = g->d;
    A.nbx = L.nbx; A.nby = L.nby; A.nbz = L.nbz;
    A.ntiles = L.ntiles;
    A.g = (u32 *)g_dev;
    A.state = (u32 *)(ws + L.state);
    A.ring = (u32 *)(ws + L.ring);
    A.ring_cap = L.ring_cap;
    A.ctr = (u32 *)(ws + L.ctr);
    A.stats = (u64 *)(ws + L.stats);
    A.rank = 0; A.world = 1;
    A.peer_ctr[0] = A.ctr;
    A.active = A.ctr + BFS_C_ACTIVE;
    const char *e = getenv("PMP_B200_BFS_TIMEOUT_MS");
    A.timeout_ns = (u64)(e ? atoll(e) : 20000) * 1000000ull;
}

static int bfs_grid_blocks(int *out) {
    static int blocks = 0;
    if (!blocks) {
        int per_sm = 0, dev = 0, n_sm = 0;
        PMP_CHECK_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, k_bfs, BFS_NT, 0));
        PMP_CHECK_CUDA(cudaGetDevice(&dev));
        PMP_CHECK_CUDA(cudaDeviceGetAttribute(&n_sm, cudaDevAttrMultiProcessorCount, dev));
        if (per_sm < 1) { pmp_set_error("k_bfs does not fit on an SM"); return -2; }
        const char *e = getenv("PMP_B200_BFS_CTAS_PER_SM");
        if (e && atoi(e) >= 1 && atoi(e) < per_sm) per_sm = atoi(e);
        blocks = per_sm * n_sm; // persistent: every warp resident, a multiple of the SM count
    }
    *out = blocks;
    return 0;
}

static int bfs_reset_ws(const BfsArgs &A, const BfsLayout &L, cudaStream_t s) {
    PMP_CHECK_CUDA(cudaMemsetAsync(A.state, 0, (size_t)L.ntiles * 4, s));
    PMP_CHECK_CUDA(cudaMemsetAsync(A.ring, 0, (size_t)L.ring_cap * 4, s));
    PMP_CHECK_CUDA(cudaMemsetAsync(A.ctr, 0, BFS_C_WORDS * 4, s));
    PMP_CHECK_CUDA(cudaMemsetAsync(A.stats, 0, 16 * 8, s));
    return 0;
}

static int bfs_launch_search(const BfsArgs &A, cudaStream_t s) {
    int blocks = 0;
    if (bfs_grid_blocks(&blocks)) return -2;
    k_bfs<<<blocks, BFS_NT, 0, s>>>(A);
    PMP_CHECK_CUDA(cudaGetLastError());
    return 0;
}

static int bfs_launch_fill(const BfsArgs &A, cudaStream_t s) {
    unsigned b = (A.ntiles + 7) / 8;
    if (b > 148 * 16) b = 148 * 16;
    k_bfs_fill<<<b, 256, 0, s>>>(A);
    PMP_CHECK_CUDA(cudaGetLastError());
    return 0;
}

// Single-GPU entry used by pmp_sssp for full fields on 6-connected unit-weight grids.
int pmp_bfs_field(pmp_grid *g, const int32_t *start_dev, void *g_dev, void *ws_dev, int64_t ws_bytes,
                  double *gcost_dev, int32_t *status_dev, uint64_t *stats_dev, void *stream) {
    const BfsLayout L = bfs_layout(g);
    if ((size_t)ws_bytes < L.total) {
        pmp_set_error("pmp_bfs_field: workspace %lld B < required %zu B", (long long)ws_bytes, L.total);
        return -1;
    }
    cudaStream_t s = (cudaStream_t)stream;
    BfsArgs A;
    bfs_fill_args(g, g_dev, ws_dev, L, A);
    if (bfs_reset_ws(A, L, s)) return -2;
    k_bfs_init<<<1, 1, 0, s>>>(A, start_dev, status_dev);
    PMP_CHECK_CUDA(cudaGetLastError());
    if (bfs_launch_search(A, s)) return -2;
    if (bfs_launch_fill(A, s)) return -2;
    k_bfs_finish<<<1, 1, 0, s>>>(A, gcost_dev, status_dev, (u64 *)stats_dev);
    PMP_CHECK_CUDA(cudaGetLastError());
    return 0;
}
