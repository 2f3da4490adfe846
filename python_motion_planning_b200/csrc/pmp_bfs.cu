// pmp_bfs.cu -- kernels + C ABI of the unit-weight cost-to-go field solver (one warp per tile,
// pmp_bfs_core.cuh has the algorithm) and of its multi-GPU z-slab form with device-side halo pulls
// over peer memory.
//
// Replaces Dijkstra3D.plan / getNeighbor (global_planner/graph_search/dijkstra3d.py:39-126) run
// to exhaustion on the six unit moves of env3d.py:56-70 -- BASELINE.json config 5.
#include <stdlib.h>
#include <string.h>

#include "pmp_common.cuh"
// the warps of a CTA enter every phase of a tile activation together (k_bfs): see the kernel
#define BFS_PHASE_SYNC() __syncthreads()
#include "pmp_bfs_core.cuh"

#ifndef BFS_NT
#define BFS_NT 128 // threads per CTA: four independent warps, no shared memory, no CTA barrier
#endif
#ifndef BFS_MINB
#define BFS_MINB 4 // resident CTAs per SM the register budget is sized for
#endif
#ifndef BFS_TRY_POLLS
#define BFS_TRY_POLLS 3 // polls of the ring slot (32, 64, 128 ns apart) before a warp joins its CTA's round without a tile
#endif
#ifndef BFS_SYNC_NT
#define BFS_SYNC_NT 128 // threads per CTA of the phase-locked kernel (4 warps in step; measured best of 128 / 256 / 512)
#endif
#ifndef BFS_SYNC_MINB
#define BFS_SYNC_MINB 4 // resident CTAs per SM of the phase-locked kernel (122 registers, no spills)
#endif
#ifndef BFS_PACE_DEFAULT
#define BFS_PACE_DEFAULT 1024 // measured: 256 starves (48 ms at 2 GPUs), 512-1024 best (22.1 ms, 1.003x cells lowered), 8192 ~ unpaced
#endif
#define BFS_RING_SLACK 8192

// The search kernel.  Persistent; one warp per tile activation.  The kernel's code is far larger
// than the instruction caches (a setup, a sweep, an extract and a publish phase of 0.3-1.5 K
// instructions each); free-running warps sit in different phases and thrash them (ncu: 29 cycles
// of "no instruction" stall per issued instruction, profiles/r02_notes.md).  So the warps of a CTA
// move through the phases in step: CTA barriers at the phase boundaries (BFS_PHASE_SYNC inside
// bfs_process_tile), warps without a tile take the same barriers.  A warp may therefore not block in
// the queue while its CTA waits for it: it polls the ring a few times (bfs_pop_warp keeps the
// ticket) and joins the round empty-handed otherwise.
// MULTI = false: a whole grid on one GPU (no peer, pacing or remote-queue code in the kernel at all);
// MULTI = true: one z-slab of several (pmp_bfs_slab_*).
template <bool MULTI>
__global__ void __launch_bounds__(BFS_SYNC_NT, BFS_SYNC_MINB) k_bfs(const __grid_constant__ BfsArgs A) {
    const int lane = (int)(threadIdx.x & 31);
    bfs_reject_check(A, lane, blockIdx.x == 0 && threadIdx.x < 32);
    BfsCount cnt;
    bfs_count_init(cnt);
    BfsPop P;
    P.ticket = 0; P.have = 0; P.t0 = 0; P.held = 0; P.held_polls = 0;
    for (;;) {
        const long long tp = BFS_NOW();
        bool done = false;
        const u64 item = bfs_pop_warp<MULTI>(A, P, lane, BFS_TRY_POLLS, done);
        if (!__syncthreads_or(item != 0ull)) {          // nobody in the CTA has work
            if (__syncthreads_and(done)) break;       // ... and the search is over for everyone
            continue;
        }
        if (item) {
            cnt.c_pop += (u64)(BFS_NOW() - tp);
            bfs_process_tile<true, MULTI>(A, item, lane, cnt);
        } else {
            if (BFS_SYNC_MASK & 1) BFS_PHASE_SYNC(); // the phase boundaries of bfs_process_tile
            if (BFS_SYNC_MASK & 2) BFS_PHASE_SYNC();
            if (BFS_SYNC_MASK & 4) BFS_PHASE_SYNC();
        }
    }
    bfs_count_flush(A, cnt, lane);
    if (lane == 0 && A.stats) atomicMax((unsigned long long *)&A.stats[10], (unsigned long long)bfs_now_ns());
}

// free-running form of the same search (every warp on its own; PMP_B200_BFS_FREE=1): the A/B partner
// of k_bfs in profiles/r02_notes.md
__global__ void __launch_bounds__(BFS_NT, BFS_MINB) k_bfs_free(const __grid_constant__ BfsArgs A) {
    bfs_warp_main<false>(A, (int)(threadIdx.x & 31), blockIdx.x == 0 && threadIdx.x < 32);
    if ((threadIdx.x & 31) == 0 && A.stats) atomicMax((unsigned long long *)&A.stats[10], (unsigned long long)bfs_now_ns());
}

// One thread: validate the source, queue its tile.  `start` = local coordinates (x, y, z) of the
// source inside this box, or nullptr when the source lies in another slab.
__global__ void k_bfs_init(BfsArgs A, const int32_t *start, int32_t *status) {
    const GridDev &Gd = A.G;
    A.stats[9] = bfs_now_ns();
    if (A.rank == 0) A.ctr[BFS_C_ACTIVE] = 1u;
    int st = PMP_ST_RUNNING;
    if (start) {
        const int sx = start[0], sy = start[1], sz = start[2];
        bool ok = sx >= 0 && sy >= 0 && sz >= 0 && sx < Gd.X && sy < Gd.Y && sz < Gd.Z;
        if (ok && Gd.ready) { // streamed upload: the source's planes must have arrived (one thread: poll here)
            const u32 *f = Gd.ready + (sz >> Gd.ready_shift);
            for (unsigned w = 0; w < (1u << 24) && bfs_ld_acquire_sys(f) == 0u; ++w) __nanosleep(200);
        }
        ok = ok && !cell_blocked(Gd, sx, sy, sz);
        if (ok) {
            const u32 t = (u32)(((sz / 8) * A.nby + sy / 8) * A.nbx + sx / 32);
            A.state[t] = BFS_SRC | BFS_Q;
            A.ring[0] = (u64)(t + 1u); // sequence number 0
            A.ctr[BFS_C_TAIL] = 1u;
            A.ctr[BFS_C_INFLIGHT] = 1u;
            A.ctr[BFS_C_SRC] = (u32)sx; A.ctr[BFS_C_SRC + 1] = (u32)sy; A.ctr[BFS_C_SRC + 2] = (u32)sz;
        } else {
            st = PMP_ST_BAD_QUERY;
            A.ctr[BFS_C_SRC + 3] = 1u; // nothing to search: the first warp of k_bfs ends the search on every rank
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

// ring capacity: tiles + slack, rounded up to a power of two where the caller's memory allows it (slot = ticket &
// (cap - 1) instead of a 32-bit division on every poll of every warp); the static-workspace entry keeps the exact size
static u32 bfs_ring_cap(u32 ntiles, bool pow2) {
    u32 cap = ntiles + BFS_RING_SLACK;
    if (pow2) { u32 p = 1u; while (p < cap) p <<= 1; cap = p; }
    return cap;
}

static BfsLayout bfs_layout(const pmp_grid *g, bool pow2 = true) {
    BfsLayout L;
    L.nbx = g->d.wpr; L.nby = g->d.PY / 8; L.nbz = g->d.PZ / 8;
    L.ntiles = (u32)((size_t)L.nbx * L.nby * L.nbz);
    L.ring_cap = bfs_ring_cap(L.ntiles, pow2);
    size_t o = 0;
    L.state = o; o = bfs_align(o + (size_t)L.ntiles * 4);
    L.ring = o; o = bfs_align(o + (size_t)L.ring_cap * 8);
    L.ctr = o; o = bfs_align(o + BFS_C_WORDS * 4);
    L.stats = o; o = bfs_align(o + 32 * 8);
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
    A.ring = (u64 *)(ws + L.ring);
    A.ring_cap = L.ring_cap;
    A.ctr = (u32 *)(ws + L.ctr);
    A.stats = (u64 *)(ws + L.stats);
    A.rank = 0; A.world = 1;
    A.peer_ctr[0] = A.ctr;
    A.active = A.ctr + BFS_C_ACTIVE;
    const char *e = getenv("PMP_B200_BFS_TIMEOUT_MS");
    A.timeout_ns = (u64)(e ? atoll(e) : 20000) * 1000000ull;
    // a tile is activated a few times; a search that pops 64x its tiles is stuck (abort with PMP_ST_ERROR)
    const u64 budget = 64ull * L.ntiles + (1ull << 20);
    A.max_pops = budget > 0xFFFFFFF0ull ? 0u : (u32)budget;
    A.pace_window = 0; // single grid: no pacing (pmp_bfs_slab_create turns it on for world > 1)
    A.pace_hold_polls = 8000; // ~2-4 ms at most per held entry
}

__global__ void k_bfs_group(const BfsArgs *args, int n);

// launch shape: persistent grid, every warp resident, a multiple of the SM count.
// PMP_B200_BFS_FREE=1: the free-running kernel; PMP_B200_BFS_NT: threads per CTA of the phase-locked
// kernel (default BFS_SYNC_NT); PMP_B200_BFS_CTAS_PER_SM: fewer CTAs than fit.
struct BfsLaunch { int blocks, nt, free_running; };
static int bfs_launch_shape(BfsLaunch *out) {
    static BfsLaunch cache[64]; // per device (zero-initialised)
    int dev = 0;
    PMP_CHECK_CUDA(cudaGetDevice(&dev));
    BfsLaunch &L = cache[dev & 63];
    if (!L.blocks) {
        int per_sm = 0, n_sm = 0;
        const char *f = getenv("PMP_B200_BFS_FREE");
        L.free_running = f && atoi(f) != 0;
        const char *t = getenv("PMP_B200_BFS_NT");
        L.nt = L.free_running ? BFS_NT : (t && atoi(t) >= 32 && atoi(t) <= BFS_SYNC_NT && atoi(t) % 32 == 0 ? atoi(t) : BFS_SYNC_NT);
        if (L.free_running) PMP_CHECK_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, k_bfs_free, L.nt, 0));
        else PMP_CHECK_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, k_bfs<true>, L.nt, 0)); // (the larger of the two)
        PMP_CHECK_CUDA(cudaDeviceGetAttribute(&n_sm, cudaDevAttrMultiProcessorCount, dev));
        if (per_sm < 1) { pmp_set_error("k_bfs does not fit on an SM"); return -2; }
        const char *e = getenv("PMP_B200_BFS_CTAS_PER_SM");
        if (e && atoi(e) >= 1 && atoi(e) < per_sm) per_sm = atoi(e);
        L.blocks = per_sm * n_sm;
    }
    *out = L;
    return 0;
}

static int bfs_grid_blocks(int *out) { // the free-running group kernel
    static int cache[64];
    int dev = 0;
    PMP_CHECK_CUDA(cudaGetDevice(&dev));
    int &blocks = cache[dev & 63];
    if (!blocks) {
        int per_sm = 0, n_sm = 0;
        PMP_CHECK_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, k_bfs_group, BFS_NT, 0));
        PMP_CHECK_CUDA(cudaDeviceGetAttribute(&n_sm, cudaDevAttrMultiProcessorCount, dev));
        if (per_sm < 1) { pmp_set_error("k_bfs_group does not fit on an SM"); return -2; }
        blocks = per_sm * n_sm;
    }
    *out = blocks;
    return 0;
}

static int bfs_reset_ws(const BfsArgs &A, const BfsLayout &L, cudaStream_t s) {
    PMP_CHECK_CUDA(cudaMemsetAsync(A.state, 0, (size_t)L.ntiles * 4, s));
    PMP_CHECK_CUDA(cudaMemsetAsync(A.ring, 0, (size_t)L.ring_cap * 8, s));
    PMP_CHECK_CUDA(cudaMemsetAsync(A.ctr, 0, BFS_C_WORDS * 4, s));
    PMP_CHECK_CUDA(cudaMemsetAsync(A.stats, 0, 32 * 8, s));
    return 0;
}

static int bfs_launch_search(const BfsArgs &A, cudaStream_t s) {
    BfsLaunch L;
    if (bfs_launch_shape(&L)) return -2;
    if (L.free_running && A.world == 1) k_bfs_free<<<L.blocks, L.nt, 0, s>>>(A);
    else if (A.world > 1) k_bfs<true><<<L.blocks, L.nt, 0, s>>>(A);
    else k_bfs<false><<<L.blocks, L.nt, 0, s>>>(A);
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
    // a streamed occupancy upload was consumed chunk by chunk by k_bfs (GridDev.ready); whatever comes next on
    // this stream sees all of it
    if (pmp_grid_upload_fence(g, s)) return -2;
    return 0;
}

// ---------------------------------------------------------------------------------------------
// pmp_cost_field_u32_static: the same search with its workspace (tile states, ring, counters) in a
// file-scope device array instead of caller memory -- for callers that cannot allocate (CUDA graph
// capture, tight serving loops, harnesses that forbid cudaMalloc).  No counterpart in the reference.
// Allocates nothing: no cudaMalloc / cudaFree, no
// allocating library call.  Limit: the workspace of the grid must fit PMP_STATIC_WS_BYTES (1 MiB:
// grids up to ~81 000 tiles of 32 x 8 x 8 cells, e.g. 512^3).  Every GPU that loads the library holds
// the array until exit.  All calls share it: it keeps the last call's contents (every word a search
// reads is cleared at the start of the call), and calls that overlap on one GPU race.
// ---------------------------------------------------------------------------------------------
#define PMP_STATIC_WS_BYTES (1u << 20)
static __device__ __align__(256) unsigned char g_bfs_static_ws[PMP_STATIC_WS_BYTES];

// In host code `g_bfs_static_ws` is a host-side stand-in, not a device address: resolve it once per
// device, with that device current.
static int bfs_static_ws_ptr(void **out) {
    static void *ptr[64] = {nullptr};
    int dev = 0;
    PMP_CHECK_CUDA(cudaGetDevice(&dev));
    if (dev < 0 || dev >= 64) { pmp_set_error("pmp_cost_field_u32_static: device index %d not supported", dev); return -1; }
    if (!ptr[dev]) {
        void *p = nullptr;
        PMP_CHECK_CUDA(cudaGetSymbolAddress(&p, g_bfs_static_ws));
        ptr[dev] = p;
    }
    *out = ptr[dev];
    return 0;
}

extern "C" int pmp_cost_field_u32_static(pmp_grid *g, const int32_t *start_dev, uint32_t *g_dev, int32_t *status_dev,
                                         uint64_t *stats_dev, void *stream) {
    if (!g || !start_dev || !g_dev) { pmp_set_error("pmp_cost_field_u32_static: bad arguments"); return -1; }
    if (!pmp_bfs_applicable(g)) {
        pmp_set_error("pmp_cost_field_u32_static: needs a 3D grid whose motion table is the six unit moves");
        return -1;
    }
    int dev = 0;
    PMP_CHECK_CUDA(cudaGetDevice(&dev));
    if (dev != g->device) { pmp_set_error("pmp_cost_field_u32_static: the grid lives on device %d, current device is %d", g->device, dev); return -1; }
    if (pmp_grid_upload_fence(g, (cudaStream_t)stream)) return -2;
    const BfsLayout L = bfs_layout(g, false);
    if (L.total > (size_t)PMP_STATIC_WS_BYTES) {
        pmp_set_error("pmp_cost_field_u32_static: grid of %u tiles needs %zu B of workspace, the static array holds %u B "
                      "(use pmp_sssp with caller memory)", L.ntiles, L.total, PMP_STATIC_WS_BYTES);
        return -1;
    }
    void *ws = nullptr;
    if (bfs_static_ws_ptr(&ws)) return -2;
    cudaStream_t s = (cudaStream_t)stream;
    BfsArgs A;
    bfs_fill_args(g, g_dev, ws, L, A);
    if (bfs_reset_ws(A, L, s)) return -2; // the array keeps the last call's contents: clear what a search reads
    k_bfs_init<<<1, 1, 0, s>>>(A, start_dev, status_dev);
    PMP_CHECK_CUDA(cudaGetLastError());
    if (bfs_launch_search(A, s)) return -2;
    if (bfs_launch_fill(A, s)) return -2;
    k_bfs_finish<<<1, 1, 0, s>>>(A, nullptr, status_dev, (u64 *)stats_dev);
    PMP_CHECK_CUDA(cudaGetLastError());
    return 0;
}

// ---------------------------------------------------------------------------------------------
// z-slabs over several GPUs (BASELINE.json config 5).
//
// Rank p owns the planes [z0, z1) of the grid (z0, z1 multiples of 8 except the last z1) as its
// LOCAL box -- no ghost planes.  Field and workspace live in one cudaMalloc'ed arena per rank,
// exported with cudaIpcGetMemHandle and mapped by the other ranks: the tiles of the lowest /
// highest tile row read their z halo straight out of the neighbour GPU's field (NVLink loads)
// and arm / queue the neighbour's tiles with remote atomics (pmp_bfs_core.cuh).  One kernel
// launch per rank per search; the host only brackets it with two barriers (every rank has
// reset its workspace before any rank searches; every rank has finished before any resets).
// Several slabs can also be held by ONE process on one GPU (pmp_bfs_slab_connect_local +
// pmp_bfs_slab_run_group): one launch serves all of them, so no kernel ever waits for another
// launch -- the single-GPU test of the same protocol.
// ---------------------------------------------------------------------------------------------
#define BFS_INFO_WORDS 8
#define BFS_MAX_RANKS 8

struct pmp_bfs_slab {
    pmp_grid *grid;
    int rank, world;
    char *arena;
    size_t arena_bytes, field_bytes;
    BfsLayout L;
    BfsArgs A;
    void *peer_base[BFS_MAX_RANKS];
    bool connected;
};

static BfsLayout bfs_layout_tiles(int nbx, int nby, int nbz) {
    BfsLayout L;
    L.nbx = nbx; L.nby = nby; L.nbz = nbz;
    L.ntiles = (u32)((size_t)nbx * nby * nbz);
    L.ring_cap = bfs_ring_cap(L.ntiles, true);
    size_t o = 0;
    L.state = o; o = bfs_align(o + (size_t)L.ntiles * 4);
    L.ring = o; o = bfs_align(o + (size_t)L.ring_cap * 8);
    L.ctr = o; o = bfs_align(o + BFS_C_WORDS * 4);
    L.stats = o; o = bfs_align(o + 32 * 8);
    L.total = o;
    return L;
}

extern "C" int pmp_bfs_slab_create(pmp_grid *g, int rank, int world, pmp_bfs_slab **out) {
    if (!g || !out || world < 1 || world > BFS_MAX_RANKS || rank < 0 || rank >= world) {
        pmp_set_error("pmp_bfs_slab_create: bad arguments (1 <= world <= %d)", BFS_MAX_RANKS);
        return -1;
    }
    if (!pmp_bfs_applicable(g)) {
        pmp_set_error("pmp_bfs_slab_create: needs a 3D grid whose motion table is the six unit moves");
        return -1;
    }
    if (rank < world - 1 && g->d.Z % 8) {
        pmp_set_error("pmp_bfs_slab_create: every slab but the last must hold a multiple of 8 planes (got %d)", g->d.Z);
        return -1;
    }
    pmp_bfs_slab *s = new pmp_bfs_slab();
    memset(s, 0, sizeof *s);
    s->grid = g; s->rank = rank; s->world = world;
    s->L = bfs_layout(g);
    s->field_bytes = bfs_align((size_t)g->d.pcells * 4);
    s->arena_bytes = s->field_bytes + s->L.total;
    cudaError_t e = cudaMalloc(&s->arena, s->arena_bytes);
    if (e != cudaSuccess) {
        pmp_set_error("pmp_bfs_slab_create: cudaMalloc(%zu B): %s", s->arena_bytes, cudaGetErrorString(e));
        delete s;
        return -2;
    }
    bfs_fill_args(g, s->arena, s->arena + s->field_bytes, s->L, s->A);
    s->A.rank = rank; s->A.world = world;
    {
        const char *pw = getenv("PMP_B200_BFS_PACE"); // entries a slab may run ahead of the global FIFO position (0: off)
        s->A.pace_window = world > 1 ? (u32)(pw ? atoi(pw) : BFS_PACE_DEFAULT) : 0u;
    }
    s->A.peer_ctr[rank] = s->A.ctr;
    if (world == 1) s->connected = true;
    *out = s;
    return 0;
}

extern "C" int pmp_bfs_slab_destroy(pmp_bfs_slab *s) {
    if (!s) return 0;
    for (int r = 0; r < BFS_MAX_RANKS; ++r)
        if (s->peer_base[r]) cudaIpcCloseMemHandle(s->peer_base[r]);
    cudaFree(s->arena);
    delete s;
    return 0;
}

// 64-byte IPC handle of the arena + 8 int64 describing its layout, for the other ranks
extern "C" int pmp_bfs_slab_export(pmp_bfs_slab *s, void *handle64, int64_t *info8) {
    if (!s || !handle64 || !info8) { pmp_set_error("pmp_bfs_slab_export: bad arguments"); return -1; }
    cudaIpcMemHandle_t h;
    PMP_CHECK_CUDA(cudaIpcGetMemHandle(&h, s->arena));
    static_assert(sizeof h == 64, "IPC handle size");
    memcpy(handle64, &h, 64);
    info8[0] = s->L.nbx; info8[1] = s->L.nby; info8[2] = s->L.nbz;
    info8[3] = s->grid->d.Z; info8[4] = (int64_t)s->field_bytes; info8[5] = (int64_t)s->arena_bytes;
    info8[6] = s->grid->d.PX; info8[7] = s->grid->d.PY;
    return 0;
}

static int bfs_set_peer(pmp_bfs_slab *s, int r, char *base, const int64_t *info) {
    if (info[0] != s->L.nbx || info[1] != s->L.nby || info[6] != s->grid->d.PX || info[7] != s->grid->d.PY) {
        pmp_set_error("pmp_bfs_slab_connect: rank %d has a different x / y extent", r);
        return -1;
    }
    const BfsLayout PL = bfs_layout_tiles((int)info[0], (int)info[1], (int)info[2]);
    char *ws = base + info[4];
    s->A.peer_ctr[r] = (u32 *)(ws + PL.ctr);
    if (r == 0) s->A.active = (u32 *)(ws + PL.ctr) + BFS_C_ACTIVE;
    if (r == s->rank - 1 || r == s->rank + 1) {
        BfsPeer &P = r < s->rank ? s->A.lo : s->A.hi;
        P.g = (u32 *)base;
        P.state = (u32 *)(ws + PL.state);
        P.ring = (u64 *)(ws + PL.ring);
        P.ctr = (u32 *)(ws + PL.ctr);
        P.ring_cap = PL.ring_cap;
        P.nbz = PL.nbz;
        P.zadj = r < s->rank ? (int)info[3] - 1 : 0;
        P.bzadj = r < s->rank ? PL.nbz - 1 : 0;
    }
    return 0;
}

// handles: world x 64 bytes, infos: world x 8 int64 (all ranks' pmp_bfs_slab_export outputs, by rank)
extern "C" int pmp_bfs_slab_connect(pmp_bfs_slab *s, const void *handles, const int64_t *infos) {
    if (!s || !handles || !infos) { pmp_set_error("pmp_bfs_slab_connect: bad arguments"); return -1; }
    for (int r = 0; r < s->world; ++r) {
        if (r == s->rank) continue;
        cudaIpcMemHandle_t h;
        memcpy(&h, (const char *)handles + 64 * r, 64);
        void *base = nullptr;
        PMP_CHECK_CUDA(cudaIpcOpenMemHandle(&base, h, cudaIpcMemLazyEnablePeerAccess));
        s->peer_base[r] = base;
        if (bfs_set_peer(s, r, (char *)base, infos + BFS_INFO_WORDS * r)) return -1;
    }
    if (s->rank == 0) s->A.active = s->A.ctr + BFS_C_ACTIVE;
    s->connected = true;
    return 0;
}

// all slabs in this process, on this device: plain pointers instead of IPC mappings
extern "C" int pmp_bfs_slab_connect_local(pmp_bfs_slab **slabs, int n) {
    if (!slabs || n < 1 || n > BFS_MAX_RANKS) { pmp_set_error("pmp_bfs_slab_connect_local: bad arguments"); return -1; }
    for (int i = 0; i < n; ++i)
        if (!slabs[i] || slabs[i]->rank != i || slabs[i]->world != n) { pmp_set_error("pmp_bfs_slab_connect_local: slabs must be given by rank"); return -1; }
    for (int i = 0; i < n; ++i) {
        pmp_bfs_slab *s = slabs[i];
        for (int r = 0; r < n; ++r) {
            if (r == i) continue;
            int64_t info[BFS_INFO_WORDS];
            char dummy[64];
            pmp_bfs_slab *p = slabs[r];
            (void)dummy;
            info[0] = p->L.nbx; info[1] = p->L.nby; info[2] = p->L.nbz; info[3] = p->grid->d.Z;
            info[4] = (int64_t)p->field_bytes; info[5] = (int64_t)p->arena_bytes; info[6] = p->grid->d.PX; info[7] = p->grid->d.PY;
            if (bfs_set_peer(s, r, p->arena, info)) return -1;
        }
        if (i == 0) s->A.active = s->A.ctr + BFS_C_ACTIVE;
        s->connected = true;
    }
    return 0;
}

// reset the workspace and queue the source tile (start_local_dev: the source in LOCAL coordinates
// if it lies in this slab, else NULL).  Every rank must have finished this call (stream
// synchronised + a barrier across ranks) before any rank calls pmp_bfs_slab_run.
extern "C" int pmp_bfs_slab_begin(pmp_bfs_slab *s, const int32_t *start_local_dev, int32_t *status_dev, void *stream) {
    if (!s || !s->connected) { pmp_set_error("pmp_bfs_slab_begin: slab not connected"); return -1; }
    cudaStream_t st = (cudaStream_t)stream;
    if (pmp_grid_upload_fence(s->grid, st)) return -2; // (the slab's copy of the grid description does not follow uploads)
    if (bfs_reset_ws(s->A, s->L, st)) return -2;
    k_bfs_init<<<1, 1, 0, st>>>(s->A, start_local_dev, status_dev);
    PMP_CHECK_CUDA(cudaGetLastError());
    return 0;
}

extern "C" int pmp_bfs_slab_run(pmp_bfs_slab *s, int32_t *status_dev, uint64_t *stats_dev, void *stream) {
    if (!s || !s->connected) { pmp_set_error("pmp_bfs_slab_run: slab not connected"); return -1; }
    cudaStream_t st = (cudaStream_t)stream;
    if (bfs_launch_search(s->A, st)) return -2;
    if (bfs_launch_fill(s->A, st)) return -2;
    k_bfs_finish<<<1, 1, 0, st>>>(s->A, nullptr, status_dev, (u64 *)stats_dev);
    PMP_CHECK_CUDA(cudaGetLastError());
    return 0;
}

// ---- several slabs, one launch (single-GPU form of the multi-GPU protocol) ----
__global__ void __launch_bounds__(BFS_NT, BFS_MINB) k_bfs_group(const BfsArgs *args, int n) {
    const u32 gwarp = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    const BfsArgs &A = args[gwarp % (u32)n];
    bfs_warp_main<true>(A, (int)(threadIdx.x & 31), gwarp < (u32)n);
    if ((threadIdx.x & 31) == 0 && A.stats) atomicMax((unsigned long long *)&A.stats[10], (unsigned long long)bfs_now_ns());
}

extern "C" int pmp_bfs_slab_run_group(pmp_bfs_slab **slabs, int n, void *stream) {
    if (!slabs || n < 1 || n > BFS_MAX_RANKS) { pmp_set_error("pmp_bfs_slab_run_group: bad arguments"); return -1; }
    cudaStream_t st = (cudaStream_t)stream;
    BfsArgs *dargs = nullptr;
    PMP_CHECK_CUDA(cudaMalloc(&dargs, sizeof(BfsArgs) * n));
    for (int i = 0; i < n; ++i) {
        if (!slabs[i] || !slabs[i]->connected) { cudaFree(dargs); pmp_set_error("pmp_bfs_slab_run_group: slab %d not connected", i); return -1; }
        PMP_CHECK_CUDA(cudaMemcpyAsync(dargs + i, &slabs[i]->A, sizeof(BfsArgs), cudaMemcpyHostToDevice, st));
    }
    int blocks = 0;
    if (bfs_grid_blocks(&blocks)) { cudaFree(dargs); return -2; }
    k_bfs_group<<<blocks, BFS_NT, 0, st>>>(dargs, n);
    cudaError_t e = cudaGetLastError();
    for (int i = 0; i < n && e == cudaSuccess; ++i) {
        if (bfs_launch_fill(slabs[i]->A, st)) { e = cudaErrorUnknown; break; }
        k_bfs_finish<<<1, 1, 0, st>>>(slabs[i]->A, nullptr, nullptr, nullptr);
    }
    cudaStreamSynchronize(st);
    cudaFree(dargs);
    if (e != cudaSuccess) { pmp_set_error("pmp_bfs_slab_run_group: %s", cudaGetErrorString(e)); return -2; }
    PMP_CHECK_CUDA(cudaGetLastError());
    return 0;
}

// field pointer (padded local box, the library's (value << 1 | 1) words), its size in bytes, and
// the 16 counters of the last run (device pointers into the arena)
extern "C" int pmp_bfs_slab_buffers(pmp_bfs_slab *s, void **field_dev, int64_t *field_bytes, void **stats_dev, void **ctr_dev) {
    if (!s) { pmp_set_error("pmp_bfs_slab_buffers: null"); return -1; }
    if (field_dev) *field_dev = s->arena;
    if (field_bytes) *field_bytes = (int64_t)s->grid->d.pcells * 4;
    if (stats_dev) *stats_dev = s->A.stats;
    if (ctr_dev) *ctr_dev = s->A.ctr;
    return 0;
}
