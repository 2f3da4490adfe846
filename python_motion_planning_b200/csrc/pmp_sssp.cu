// pmp_sssp.cu -- tiled, windowed label-correcting shortest paths on the occupancy bitmask.
//
// Replaces the expansion loops of the reference planners (global_planner/graph_search/
// a_star.py:39-83, dijkstra.py:36-85, a_star3d.py:33-78, dijkstra3d.py:39-87) together with
// getNeighbor (a_star.py:85-96, a_star3d.py:80-89, dijkstra3d.py:89-126) and isCollision
// (graph_search.py:61-87, graph_search_3d.py:66-107).
//
// Scheme (DESIGN.md section 3.1 has the long form):
//   * the grid is cut into tiles (32x32 cells in 2D, 16x8x8 in 3D); a cost field g is a dense
//     array per query whose words carry the value and a "still has to propagate" flag
//     ((value << 1) | clean); per (query, tile) there is a `minkey`, the smallest key pushed into
//     the tile since it was last claimed;
//   * one persistent cooperative kernel (k_sssp) runs rounds: SCAN every (query, tile) and queue
//     those whose minkey lies inside the query's key window; then an asynchronous cascade: CTAs
//     pop tiles from a ring, process_tile stages g (tile + 1-cell halo) and the occupancy rows in
//     shared memory, settles cells in key order (owner threads keep the per-cell state in
//     registers, frontier records are expanded with one lane per direction; unit-weight grids
//     use the bit-parallel level sweep solve_bfs), publishes with atomic min and queues the
//     neighbour tiles it lowered -- no grid barrier inside a round.  Single A-star queries use
//     the banded ring (ring_push / banded_pop): the window is cut into key bands served lowest
//     first.  Global g only ever decreases, every value is fl(g(u) + w) for a legal edge, and a
//     cell keeps its flag until it has propagated, so the fixed point is the same field a
//     sequential Dijkstra produces -- bit for bit, also for the sqrt(2)/sqrt(3) weights;
//   * A-star = the same with key g + h (h an admissible round-down heuristic), pruning by the best
//     g[goal] seen, and termination once no pending key is below g[goal];
//   * slab mode (pmp_slab_*: one z-slab per GPU, planes exchanged by the host) and incremental
//     replanning (pmp_replan.cuh) resume the same kernel from a stored field.
#include <cooperative_groups.h>
#include <stdlib.h>
#include <string.h>
#include <math.h>

#include "pmp_common.cuh"

namespace cg = cooperative_groups;

#define NT 256
#ifndef PMP_TP_MINB
#define PMP_TP_MINB 4 // CTAs per SM of the throughput build of k_sssp (64 registers; measured best for batches and fields)
#endif
#ifndef PMP_MINB
#define PMP_MINB 2 // resident CTAs per SM the register budget is sized for (128 registers: no spills)
#endif

template <int TX_, int TY_, int TZ_, typename W_>
struct TileCfg {
    typedef W_ W;
    typedef typename W_::G G;
    static constexpr int TX = TX_, TY = TY_, TZ = TZ_;
    static constexpr int ZH = TZ_ > 1 ? 1 : 0; // halo width in z (0 for 2D)
    static constexpr int HX = TX_ + 2, HY = TY_ + 2, HZ = TZ_ + 2 * ZH;
    static constexpr int NCELL = TX_ * TY_ * TZ_;
    static constexpr int CPT = NCELL / NT;     // cells owned by one thread
    static constexpr int HC = HX * HY * HZ;
    static constexpr int NCW = (HC + 31) / 32; // words of the changed mask
    static constexpr int NROWS = HY * HZ;
    static_assert(NCELL == 1024 && CPT == 4, "tiles hold 1024 cells, 4 per thread");
    static_assert((TZ_ == 1 && TX_ == 32 && TY_ == 32) || (TX_ == 16 && TY_ == 8 && TZ_ == 8), "ownership map");
    // Cells owned by thread `tid` (j = 0..3).  The map is skewed so that the cells of any
    // axis-aligned line / plane of the tile are spread evenly over the 8 warps: every warp
    // expands the frontier cells its own lanes own, so a wavefront sweeping the tile keeps all
    // warps equally busy.
    __device__ static inline void owned(int tid, int j, int &lx, int &ly, int &lz) {
        if (TZ_ == 1) {
            const int tlx = tid & 31, w = tid >> 5;
            lx = (tlx + 8 * j) & 31;
            ly = ((w + (tlx >> 2)) & 7) + 8 * j;
            lz = 0;
        } else {
            const int tlx = tid & 15, tly = (tid >> 4) & 7, tlz = tid >> 7;
            lx = (tlx + 4 * j) & 15;
            ly = (tly + 2 * j) & 7;
            lz = 2 * j + (tlz ^ (tly & 1));
        }
    }
    // lanes of a warp that work on one cell (one lane per canonical direction) and cells
    // expanded per warp step
    static constexpr int DPL = TZ_ == 1 ? 16 : 32;
    static constexpr int NG = 32 / DPL;
    __device__ static inline int own_hi(int tid, int j) {
        int lx, ly, lz;
        owned(tid, j, lx, ly, lz);
        return ((lz + ZH) * HY + (ly + 1)) * HX + (lx + 1);
    }
    __device__ static inline int halo_index(int lx, int ly, int lz) {
        return ((lz + ZH) * HY + (ly + 1)) * HX + (lx + 1);
    }
};

#define BFS_LV 63 // levels one pass of the unit-weight solver sweeps (2 x 64 x 32 words = the record queues' 16 KB)

template <class C>
struct TileSmem {
    typename C::G g[C::HC]; // cost words of tile + halo in the global (value << 1 | clean) format
    u64 rows[C::NROWS];     // occupancy bits of the TX+2 cells of each (y,z) row
    u64 nbrmin[27];         // smallest key pushed into each neighbour tile
    u64 kmin[2];            // min key of the pending cells (computed on empty sub-iterations)
    u64 T, U, k0;
    u64 Tlim;               // banded mode: the popped band's upper key bound (KEY_INF otherwise)
    u32 chg[C::NCW];        // cells (tile + halo, by halo'd index) lowered during this activation
    union {
        ulonglong2 fq[NT / 32][C::NCELL / (NT / 32)]; // per-warp frontier records: .x = value, .y = free mask | halo'd index << 32
        struct {                  // unit-weight fast path (solve_bfs): one bit per cell, by level
            u32 fin[BFS_LV + 1][32];  // cells whose value at load is Lmin + i
            u32 seed[BFS_LV][32];     // ... and that are pending
        } b;
    };
    u32 bw[4][32];          // solve_bfs: free cells, cells below Lmin, cells lowered, cells expanded
    u32 bmin, bmax;         // solve_bfs: smallest / largest pending level
    u64 stat[16];           // per-CTA counters (flushed at kernel end)
    u64 tm[6];              // CTA 0's phase timers: t0, scan, proc, wait, ta/tb scratch
    u32 item, n_edge, n_exp;
    u32 fqn[2];             // records in the CTA-wide queue (double-buffered by sub-iteration parity)
    u32 s_live, s_wl_base;
};

struct SsspArgs {
    GridDev G;
    // env.motions by canonical direction d = (dz+1)*9 + (dy+1)*3 + (dx+1): edge cost (f64 bits or
    // integer) and the cells of the source's 3x3x3 neighbourhood that must be free (isCollision);
    // directions not in the motion table have need = all-ones (never satisfied).
    u64 wdir[27];
    u32 need[27];
    int bfs;                  // unit-cost axis motions + integer field: bit-parallel in-tile solver
    int ndir, dpl_log2;       // number of motions; lanes per cell in the expansion = 1 << dpl_log2 (>= ndir)
    unsigned char dirs[32];   // canonical direction handled by lane index i of a lane group (13 = none)
    u32 nq, ring_cap;
    const int32_t *starts, *goals;
    void *g;
    u64 *minkey;
    u32 *wl[2];    // [0] ring of queued (query, tile) ids + 1 (0 = empty slot), [1] unused
    u64 *qmin[2];
    u64 *T[2];
    int *done;
    u32 *goalcell; // padded cell index of each goal (0xFFFFFFFF = none)
    u32 *ctr;      // per round parity p: [p] ring tail, [2+p] ring head, [4+p] live tiles, [6+p] tiles in flight;
                   // [8] abort flag (a ring wait timed out: the result is invalid), [9] the search ran out of live tiles
    u64 *stats;    // 16 counters, see include/pmp_b200.h (pmp_sssp)
    double window, sub, he, hm;
    int nbands;   // > 1: the ring is split into key bands served lowest first (single A-star queries)
    double inv_band_w; // 1 / band width
    u32 *rb;      // banded mode: [parity][avail | head | tail][32] counters
    int no_prune; // replanning: keep relaxations that cannot beat the current g[goal] (pmp_replan.cuh)
    u64 key_cap; // slab mode: only keys below this bound are processed (KEY_INF otherwise)
    int use_h, full_field, max_rounds;
};

// shared-memory atomic min that reports whether it lowered the word
__device__ static inline bool smem_min(u64 *p, u64 v) { return atomicMin((unsigned long long *)p, (unsigned long long)v) > v; }
__device__ static inline bool smem_min(u32 *p, u32 v) { return atomicMin(p, v) > v; }

__device__ static inline u64 globaltimer_ns() {
    u64 t;
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
    return t;
}

__device__ static inline u64 warp_min_u64(u64 v) {
#pragma unroll
    for (int o = 16; o; o >>= 1) {
        const u64 other = __shfl_xor_sync(0xFFFFFFFFu, v, o);
        v = other < v ? other : v;
    }
    return v;
}

// One frontier record against this lane's motion.
template <class C>
__device__ __forceinline__ void relax_record(TileSmem<C> &S, const SsspArgs &A, const ulonglong2 rec,
                                             const u32 my_need, const u64 my_w, const int my_off,
                                             const u64 U, const int use_h, const int rx, const int ry,
                                             const int rz, u32 &my_edges) {
    typedef typename C::W W;
    typedef typename C::G G;
    const u32 fm = (u32)rec.y;
    if ((fm & my_need) != my_need) return; // isCollision: target / face cells occupied
    ++my_edges;
    const int vhi = (int)(rec.y >> 32) + my_off;
    const G tvr = S.g[vhi];
    const G cand = W::addw((G)rec.x, my_w);
    if (cand >= W::dec(tvr)) return;
    if (U != KEY_INF) { // cannot be on a better path to the goal?
        double hv = 0.0;
        if (use_h) {
            const int hx = vhi % C::HX, hy = (vhi / C::HX) % C::HY, hz = vhi / (C::HX * C::HY);
            hv = heuristic_lo(rx + hx - 1, ry + hy - 1, rz + hz - C::ZH, A.he, A.hm);
        }
        if (W::key(cand, hv) > U) return;
    }
    const G want = W::enc(cand, true);
    const G old = atomicCAS(&S.g[vhi], tvr, want);
    bool won = old == tvr;
    if (!won && want < old) won = smem_min(&S.g[vhi], want); // raced with a worse value
    if (won) atomicOr(&S.chg[vhi >> 5], 1u << (vhi & 31));
}

// ---------------------------------------------------------------------------------------------
// Unit-weight in-tile solver (integer fields, axis motions of cost 1: the 4/6-connected grids).
// With unit weights the cost is a BFS level, so a whole tile is one bit per cell in the 32
// lanes of ONE warp (3D: word = two x-rows of 16, lane = lz*4 + ly/2; 2D: word = one row) and a
// level step is a handful of shifts and shuffles instead of a round of shared-memory atomics:
//     frontier(L)  = cells assigned L by the previous step | pending cells whose value is L
//     next         = neighbours(frontier) & free & ~(cells whose value is already <= L + 1)
// Cells are first bucketed by their value at load (fin / seed masks per level, by all 256
// threads); afterwards every thread pushes its expanded boundary cells into the halo.  The
// result is the same fixed point as the generic solver's.
// ---------------------------------------------------------------------------------------------
template <class C>
__device__ __forceinline__ void bfs_cellbit(int lx, int ly, int lz, int &w, u32 &bit) {
    if (C::TZ == 1) { w = ly; bit = 1u << lx; }
    else { w = lz * 4 + (ly >> 1); bit = 1u << ((ly & 1) * 16 + lx); }
}

// The per-thread state goes in and out by value.  Measured on the 768^3 field (4 CTAs/SM, 64
// registers): inlined 20.9 ms; out of line (__noinline__) the sweep itself is twice as fast (the
// loop no longer runs on spilled values) but the caller's spills around the call cost more than
// that: 22.8 ms.  PMP_BFS_ATTR selects the variant.
struct BfsIo {
    uint4 F, gprop;
    u64 km_final;
    u32 n_exp, n_iter, my_edges;
};

#ifndef PMP_BFS_ATTR
#define PMP_BFS_ATTR __forceinline__
#endif
template <class C>
__device__ PMP_BFS_ATTR BfsIo solve_bfs(TileSmem<C> &S, const SsspArgs &A, const u64 T, BfsIo io) {
    typedef typename C::W W;
    typedef u32 G;
    constexpr int CPT = C::CPT;
    static_assert(CPT == 4, "solve_bfs passes the four owned cells in a uint4");
    const int tid = threadIdx.x, lane = tid & 31;
    const u32 F[CPT] = {io.F.x, io.F.y, io.F.z, io.F.w};
    G gprop[CPT] = {io.gprop.x, io.gprop.y, io.gprop.z, io.gprop.w};
    u64 km_final = KEY_INF;
    u32 n_exp = 0, n_iter = 0, my_edges = 0;
    // canonical directions of the axis moves: d = (dz+1)*9 + (dy+1)*3 + (dx+1)
    const bool xp = A.need[14] != 0xFFFFFFFFu, xm = A.need[12] != 0xFFFFFFFFu;
    const bool yp = A.need[16] != 0xFFFFFFFFu, ym = A.need[10] != 0xFFFFFFFFu;
    const bool zp = A.need[22] != 0xFFFFFFFFu, zm = A.need[4] != 0xFFFFFFFFu;
    int cw[CPT];
    u32 cb[CPT];
#pragma unroll
    for (int j = 0; j < CPT; ++j) {
        int lx, ly, lz;
        C::owned(tid, j, lx, ly, lz);
        bfs_cellbit<C>(lx, ly, lz, cw[j], cb[j]);
    }
    for (;;) {
#ifdef PMP_BFSSTAT
        const long long bt0 = clock64();
#endif
        // ---- smallest pending level; clear the level masks ----
        u32 lminp = 0xFFFFFFFFu, lmaxp = 0;
        G gv[CPT];
#pragma unroll
        for (int j = 0; j < CPT; ++j) {
            gv[j] = W::dec(S.g[C::own_hi(tid, j)]);
            if (gv[j] < gprop[j]) {
                lminp = gv[j] < lminp ? (u32)gv[j] : lminp;
                lmaxp = gv[j] > lmaxp ? (u32)gv[j] : lmaxp;
            }
        }
        lminp = __reduce_min_sync(0xFFFFFFFFu, lminp);
        lmaxp = __reduce_max_sync(0xFFFFFFFFu, lmaxp);
        if (tid == 0) { S.bmin = 0xFFFFFFFFu; S.bmax = 0; }
        for (int i = tid; i < (2 * BFS_LV + 1) * 32; i += NT) (&S.b.fin[0][0])[i] = 0;
        if (tid < 128) S.bw[tid >> 5][tid & 31] = 0;
        __syncthreads();
        if (lane == 0 && lminp != 0xFFFFFFFFu) { atomicMin(&S.bmin, lminp); atomicMax(&S.bmax, lmaxp); }
        __syncthreads();
        const u32 Lmin = S.bmin;
        if (Lmin == 0xFFFFFFFFu) { km_final = KEY_INF; break; }
        if ((u64)Lmin >= T) { km_final = (u64)Lmin; break; }
        const u64 lend64 = (u64)Lmin + BFS_LV;
        const u32 Lend = (u32)(lend64 < T ? lend64 : T);
        // ---- bucket the owned cells by their value ----
#pragma unroll
        for (int j = 0; j < CPT; ++j) {
            if (!((F[j] >> 13) & 1u)) continue; // the cell itself is blocked
            atomicOr(&S.bw[0][cw[j]], cb[j]);
            const G v = gv[j];
            if (v == W::inf()) continue;
            if (v < Lmin) { atomicOr(&S.bw[1][cw[j]], cb[j]); continue; }
            const u32 i = (u32)v - Lmin;
            if (i > BFS_LV) continue;
            atomicOr(&S.b.fin[i][cw[j]], cb[j]);
            if (v < gprop[j] && i < BFS_LV) atomicOr(&S.b.seed[i][cw[j]], cb[j]);
        }
        __syncthreads();
#ifdef PMP_BFSSTAT
        const long long bt1 = clock64();
#endif
        // ---- the sweep: warp 0, lane = word ----
        if (tid < 32) {
#ifdef PMP_BFSSTAT
            const long long st0 = clock64();
#endif
            const u32 freew = S.bw[0][lane];
            u32 vis = S.bw[1][lane], lowered = 0, expd = 0, carry = 0, nx = 0, iters = 0;
            const u32 Lmaxp = S.bmax;
            u32 L = Lmin;
            // the rows of the next level are loaded one step ahead of their use
            u32 fin_cur = S.b.fin[0][lane], seed_cur = S.b.seed[0][lane];
            for (; L < Lend; ++L) {
                const int i = (int)(L - Lmin);
                const u32 fin_nxt = S.b.fin[i + 1][lane];
                const u32 seed_nxt = i + 1 < BFS_LV ? S.b.seed[i + 1][lane] : 0u;
                vis |= fin_cur;
                const u32 Fr = carry | (seed_cur & ~lowered);
                carry = 0;
                fin_cur = fin_nxt;
                seed_cur = seed_nxt;
                if (!__any_sync(0xFFFFFFFFu, Fr != 0)) {
                    S.b.seed[i][lane] = lowered;
                    if (L >= Lmaxp) { ++L; break; } // no frontier and no pending cell further up
                    continue;
                }
                ++iters;
                expd |= Fr;
                nx += __popc(Fr);
                const u32 up1 = __shfl_up_sync(0xFFFFFFFFu, Fr, 1), dn1 = __shfl_down_sync(0xFFFFFFFFu, Fr, 1);
                u32 N = 0;
                if (C::TZ == 1) {
                    if (xp) N |= Fr << 1;
                    if (xm) N |= Fr >> 1;
                    if (yp && lane >= 1) N |= up1;
                    if (ym && lane < 31) N |= dn1;
                } else {
                    const u32 up4 = __shfl_up_sync(0xFFFFFFFFu, Fr, 4), dn4 = __shfl_down_sync(0xFFFFFFFFu, Fr, 4);
                    if (xp) N |= (Fr & 0x7FFF7FFFu) << 1;
                    if (xm) N |= (Fr & 0xFFFEFFFEu) >> 1;
                    if (yp) { N |= (Fr & 0xFFFFu) << 16; if (lane & 3) N |= up1 >> 16; }
                    if (ym) { N |= Fr >> 16; if ((lane & 3) != 3) N |= (dn1 & 0xFFFFu) << 16; }
                    if (zp && lane >= 4) N |= up4;
                    if (zm && lane < 28) N |= dn4;
                }
                N &= freew & ~(vis | fin_nxt);
                lowered |= N;
                vis |= N;
                carry = N;
                // seed[i] is consumed: reuse the row for the cells lowered up to and including this
                // step (value Lmin + i + 1 or less); owners binary-search their level afterwards
                S.b.seed[i][lane] = lowered;
            }
            if (lane == 0) S.bmax = L - Lmin; // number of level rows written
            S.bw[2][lane] = lowered;
            S.bw[3][lane] = expd;
            n_exp += nx;
            if (lane == 0) n_iter += iters;
#ifdef PMP_BFSSTAT
            if (lane == 0) { S.stat[14] += (u64)(clock64() - st0); S.stat[12] += iters; }
#endif
        }
        __syncthreads();
#ifdef PMP_BFSSTAT
        const long long bt2 = clock64();
#endif
        // ---- owners: expanded cells are clean now; push them into the halo ----
#pragma unroll
        for (int j = 0; j < CPT; ++j) {
            const int hj = C::own_hi(tid, j);
            G v = W::dec(S.g[hj]);
            if (S.bw[2][cw[j]] & cb[j]) { // lowered: first level row whose cumulative mask has the bit
                int lo = 0, hi = (int)S.bmax - 1;
                while (lo < hi) {
                    const int mid = (lo + hi) >> 1;
                    if (S.b.seed[mid][cw[j]] & cb[j]) hi = mid; else lo = mid + 1;
                }
                v = (G)(Lmin + (u32)lo + 1u);
                S.g[hj] = W::enc(v, true);
                atomicOr(&S.chg[hj >> 5], 1u << (hj & 31));
            }
            if (!(S.bw[3][cw[j]] & cb[j])) continue;
            gprop[j] = v;
            int lx, ly, lz;
            C::owned(tid, j, lx, ly, lz);
            const G want = W::enc(v + 1, true);
#pragma unroll
            for (int d = 0; d < 6; ++d) {
                const int dx = d == 0 ? 1 : (d == 1 ? -1 : 0), dy = d == 2 ? 1 : (d == 3 ? -1 : 0);
                const int dz = d == 4 ? 1 : (d == 5 ? -1 : 0);
                if (C::TZ == 1 && dz != 0) continue;
                const int dcan = (dz + 1) * 9 + (dy + 1) * 3 + (dx + 1);
                if (A.need[dcan] == 0xFFFFFFFFu || !((F[j] >> dcan) & 1u)) continue;
                ++my_edges;
                if ((unsigned)(lx + dx) < (unsigned)C::TX && (unsigned)(ly + dy) < (unsigned)C::TY &&
                    (unsigned)(lz + dz) < (unsigned)C::TZ)
                    continue; // interior target: done by the sweep
                const int vhi = hj + (dz * C::HY + dy) * C::HX + dx;
                if (v + 1 < W::dec(S.g[vhi]) && W::dec(atomicMin(&S.g[vhi], want)) > v + 1)
                    atomicOr(&S.chg[vhi >> 5], 1u << (vhi & 31));
            }
        }
        __syncthreads();
#ifdef PMP_BFSSTAT
        if (tid == 32) {
            S.stat[13] += (u64)(bt1 - bt0);
            S.stat[15] += (u64)(clock64() - bt2);
        }
#endif
    }
    io.gprop = make_uint4(gprop[0], gprop[1], gprop[2], gprop[3]);
    io.km_final = km_final;
    io.n_exp = n_exp; io.n_iter = n_iter; io.my_edges = my_edges;
    return io;
}

// ---------------------------------------------------------------------------------------------
// Banded ring (single A-star queries).  Inside a round the window [T - W, T) is cut into `nbands`
// key bands, each with its own FIFO ring; CTAs always serve the lowest non-empty band and a tile
// popped from band b only works up to that band's upper bound, re-queueing itself for the rest.
// The cascade therefore runs in approximate key order (little re-relaxation, like a narrow
// window) while every band of the wide window is available as parallel work (no barrier per band).
// Counters per band: avail (a semaphore: completed pushes minus reservations), head (pop tickets),
// tail (push tickets).  A pop reserves an item on the semaphore first and only then takes a
// ticket, so a ticket never runs ahead of the pushes; the slot of a ticket may still be in flight
// (a later push completed first) and is then awaited for a moment.
// ---------------------------------------------------------------------------------------------
#define PMP_MAXBANDS 16
#ifndef PMP_DEFAULT_BANDS
#define PMP_DEFAULT_BANDS 12 // measured on c4 (512^3 A-star): FIFO ring + window 6: 6.6 ms; 12 bands x 3: 5.6 ms
#endif
#ifndef PMP_BANDED_WINDOW
#define PMP_BANDED_WINDOW 36.0 // key window of a round in banded mode (x smallest edge cost)
#endif
__device__ static inline u64 band_top(const SsspArgs &A, int b, u64 Tw) {
    if (b >= A.nbands - 1) return Tw;
    const double t = __longlong_as_double((long long)Tw) - (double)(A.nbands - 1 - b) / A.inv_band_w;
    return t > 0.0 ? (u64)__double_as_longlong(t) : 0ull;
}
// the band of key k (< Tw) = the first band whose top lies above k.  Defined through band_top so
// that a tile served up to band_top(b) always continues in a band > b, also for keys that sit
// exactly on a boundary (an independent rounding here could send it back to band b for ever).
__device__ static inline int band_of(const SsspArgs &A, u64 k, u64 Tw) {
    const double d = __longlong_as_double((long long)Tw) - __longlong_as_double((long long)k); // > 0
    int b = A.nbands - 1 - (int)(d * A.inv_band_w);
    b = b < 0 ? 0 : (b > A.nbands - 1 ? A.nbands - 1 : b);
    while (b < A.nbands - 1 && k >= band_top(A, b, Tw)) ++b;
    while (b > 0 && k < band_top(A, b - 1, Tw)) --b;
    return b;
}
// append (query, tile) gt to the ring of round parity `cur`; k = the key that brings it in
__device__ static inline void ring_push(const SsspArgs &A, int cur, u32 gt, u64 k, u64 Tw) {
    atomicAdd(&A.ctr[6 + cur], 1u); // in flight (queued or being processed)
    if (A.nbands <= 1) {
        const u32 pos = atomicAdd(&A.ctr[cur], 1u);
        *(volatile u32 *)&A.wl[0][pos % A.ring_cap] = gt + 1u;
        return;
    }
    const int b = band_of(A, k, Tw);
    u32 *c = A.rb + cur * 96;
    const u32 pos = atomicAdd(&c[64 + b], 1u);
    *(volatile u32 *)&A.wl[0][(size_t)b * A.ring_cap + pos % A.ring_cap] = gt + 1u;
    fence_gpu();
    atomicAdd(&c[b], 1u);
}

// ---------------------------------------------------------------------------------------------
// one tile activation
//
// Owner-computes push.  Every thread owns 4 cells of the tile and keeps their search state in
// registers: the halo'd index, the free-mask of the 3x3(x3) neighbourhood, the heuristic and
// `gprop`, the value it last propagated (a cell is pending iff its value is below gprop).  One
// sub-iteration: read the own cells, expand those that are pending with key < B (loop over the
// motions: read the target, compare, shared-memory atomic min for the few that improve), one
// barrier-with-vote.  B advances by the smallest edge cost per sub-iteration, or jumps to the
// smallest pending key when a sub-iteration expanded nothing.  A cell lowered by another thread
// after its owner read it is simply seen in the next sub-iteration: no update can be lost, and
// the fixed point does not depend on the order.
// ---------------------------------------------------------------------------------------------
template <class C>
__device__ void process_tile(const SsspArgs &A, const u32 gt, const int cur, TileSmem<C> &S) {
    typedef typename C::W W;
    typedef typename C::G G;
    constexpr int CPT = C::CPT;
    constexpr int NLD = (C::HC + NT - 1) / NT;
    const GridDev &Gd = A.G;
    const int tid = threadIdx.x, lane = tid & 31;
    const u32 q = gt / Gd.ntiles, t = gt - q * Gd.ntiles;
    const int ttx = t % Gd.ntx, tty = (t / Gd.ntx) % Gd.nty, ttz = t / (Gd.ntx * Gd.nty);
    const int ox = ttx * C::TX, oy = tty * C::TY, oz = ttz * C::TZ;
    G *gq = (G *)A.g + (size_t)q * Gd.pcells;
    const int use_h = A.use_h;
    int gx = 0, gy = 0, gz = 0;
    if (use_h) {
        gx = __ldg(A.goals + 3 * q); gy = __ldg(A.goals + 3 * q + 1); gz = Gd.ndim == 3 ? __ldg(A.goals + 3 * q + 2) : 0;
    }

    const u32 t_begin = (u32)clock64();
    // ---- prologue: the dependent global round trips are spread over different warps ----
    if (tid == 0) {
        // claim.  Invariant: a tile is in the ring iff its minkey is below the window bound T, so
        // resetting minkey also "leaves the queue": a later push below T appends the tile again.
        S.k0 = atomicExch(&A.minkey[gt], KEY_INF);
        fence_gpu();
    } else if (tid == 32) {
        S.T = ld_cg(&A.T[cur][q]);
    } else if (tid == 64) {
        u64 U = KEY_INF;
        const u32 gc = (A.full_field || A.no_prune) ? 0xFFFFFFFFu : __ldg(A.goalcell + q);
        if (gc != 0xFFFFFFFFu) {
            const G gg = W::dec(ld_cg(gq + gc));
            if (gg != W::inf()) U = W::key(gg, 0.0);
        }
        S.U = U;
    } else if (tid == 96) {
        S.kmin[0] = S.kmin[1] = KEY_INF;
        S.fqn[0] = S.fqn[1] = 0;
    }
    if (tid >= 128 && tid < 128 + 27) S.nbrmin[tid - 128] = KEY_INF;
    if (tid >= NT - C::NROWS) { // occupancy never changes during a search: no ordering needed
        const int r = tid - (NT - C::NROWS);
        const int hy = r % C::HY, hz = r / C::HY;
        S.rows[r] = row_bits(Gd, ox - 1, oy + hy - 1, oz + hz - C::ZH, C::HX);
    }
    for (int i = tid; i < C::NCW; i += NT) S.chg[i] = 0;
    __syncthreads(); // the claim is ordered before every load of g below
    if (A.nbands > 1) {
        const u64 k0 = S.k0, tw = S.T, tl = S.Tlim < tw ? S.Tlim : tw;
        if (k0 == KEY_INF) return; // a duplicate entry of a tile that was served already
        if (k0 >= tl) {            // queued for an earlier band than its keys: hand it on unopened
            if (tid == 0) {
                const u64 was = atomicMin(&A.minkey[gt], k0);
                if (k0 < tw && was >= tw) ring_push(A, cur, gt, k0, tw);
            }
            return;
        }
    }

    // ---- stage the raw cost words (tile + halo); while the loads are in flight, set up the
    // per-thread state of the owned cells that depends only on the occupancy and the position ----
    u32 F[CPT];
    G gprop[CPT];
    double hh[CPT];
    u32 ldp = 0; // owned cells that were flagged pending in global memory
    {
        G raw[NLD];
#pragma unroll
        for (int k = 0; k < NLD; ++k) {
            const int i = tid + k * NT;
            const int hx = i % C::HX, hy = (i / C::HX) % C::HY, hz = i / (C::HX * C::HY);
            const int x = ox + hx - 1, y = oy + hy - 1, z = oz + hz - C::ZH;
            raw[k] = W::ginf();
            if (i < C::HC && (unsigned)x < (unsigned)Gd.PX && (unsigned)y < (unsigned)Gd.PY && (unsigned)z < (unsigned)Gd.PZ)
                raw[k] = ld_cg(gq + ((size_t)z * Gd.PY + y) * Gd.PX + x);
        }
#pragma unroll
        for (int j = 0; j < CPT; ++j) {
            int lx, ly, lz;
            C::owned(tid, j, lx, ly, lz);
            u32 occ27 = 0; // occupancy of the 3x3(x3) neighbourhood from the staged rows
#pragma unroll
            for (int dzz = -C::ZH; dzz <= C::ZH; ++dzz)
#pragma unroll
                for (int dyy = -1; dyy <= 1; ++dyy) {
                    const u64 r = S.rows[(lz + C::ZH + dzz) * C::HY + (ly + 1 + dyy)];
                    occ27 |= (u32)((r >> lx) & 7ull) << ((dzz + 1) * 9 + (dyy + 1) * 3);
                }
            F[j] = ~occ27 & 0x7FFFFFFu;
            hh[j] = use_h ? heuristic_lo(ox + lx - gx, oy + ly - gy, oz + lz - gz, A.he, A.hm) : 0.0;
        }
#pragma unroll
        for (int k = 0; k < NLD; ++k) {
            const int i = tid + k * NT;
            if (i < C::HC) S.g[i] = raw[k];
        }
    }
    __syncthreads();
#pragma unroll
    for (int j = 0; j < CPT; ++j) {
        const G sw = S.g[C::own_hi(tid, j)];
        const bool pend = !(sw & 1);
        gprop[j] = pend ? W::inf() : W::dec(sw);
        if (pend) ldp |= 1u << j;
    }

    // direction handled by this lane in the warp-cooperative expansion: a group of 2^dpl_log2 lanes
    // (8 for up to 8 motions, 16, or 32) works on one frontier cell, one lane per motion
    const int dpl_log2 = A.dpl_log2;
    const int grp = lane >> dpl_log2, ngrp = 32 >> dpl_log2;
    const int dcan = A.dirs[lane & ((1 << dpl_log2) - 1)];
    const u32 my_need = A.need[dcan]; // all-ones for unused lanes
    const u64 my_w = A.wdir[dcan];
    const int my_off = ((dcan / 9 - 1) * C::HY + ((dcan / 3) % 3 - 1)) * C::HX + (dcan % 3 - 1);
    ulonglong2 *wq = S.fq[tid >> 5];

    const u32 t_loaded = (u32)clock64();
    const u64 Tw = S.T, U = S.U;                       // Tw: top of the round's window
    const u64 T = S.Tlim < Tw ? S.Tlim : Tw;          // this activation works on keys below T
    // exclusive upper key of the current sub-bucket.  The claimed minkey is a lower bound of the
    // smallest pending key, so the first bucket is usually non-empty.
    u64 B = 0;
    if (S.k0 != KEY_INF) {
        const u64 nb = W::key_add(S.k0, A.sub);
        B = nb < T ? nb : T;
    }
    u64 km_final = KEY_INF;
    u32 n_exp = 0, n_iter = 0, my_edges = 0, n_empty = 0;
    const bool bfs = W::kind == PMP_W_U32 && A.bfs;
    if (bfs) {
        if constexpr (W::kind == PMP_W_U32) { // A.bfs is only ever set for integer fields
            BfsIo io;
            io.F = make_uint4(F[0], F[1], F[2], F[3]);
            io.gprop = make_uint4(gprop[0], gprop[1], gprop[2], gprop[3]);
            io = solve_bfs<C>(S, A, T, io);
            gprop[0] = io.gprop.x; gprop[1] = io.gprop.y; gprop[2] = io.gprop.z; gprop[3] = io.gprop.w;
            km_final = io.km_final;
            n_exp = io.n_exp; n_iter = io.n_iter; my_edges = io.my_edges;
        }
    }
    else for (;;) {
        G gv[CPT];
#pragma unroll
        for (int j = 0; j < CPT; ++j) gv[j] = W::dec(S.g[C::own_hi(tid, j)]);
        u64 lmin = KEY_INF;
        bool act[CPT];
        u32 am[CPT], tot = 0;
#pragma unroll
        for (int j = 0; j < CPT; ++j) {
            act[j] = false;
            if (gv[j] < gprop[j]) { // pending
                const u64 k = use_h ? W::key(gv[j], hh[j]) : (u64)gv[j];
                if (k < B) { act[j] = true; gprop[j] = gv[j]; ++n_exp; }
                else lmin = k < lmin ? k : lmin;
            }
            am[j] = __ballot_sync(0xFFFFFFFFu, act[j]);
            tot += __popc(am[j]);
        }
        const int anyact = tot != 0;
        // ---- warp-cooperative expansion: frontier cells become records in shared memory and are
        // expanded one lane per motion: test the edge, read the target, one compare-and-swap if it
        // improves.  With more than 16 motions a cell takes a whole warp, so the records of all
        // warps go into one CTA-wide queue and are dealt round-robin to the 8 warps (one extra
        // barrier, but a skewed frontier no longer serialises on one warp); with fewer motions
        // each warp expands the cells its own lanes own, 2-4 at a time.
        int vote;
        if (dpl_log2 == 5) {
            const int pb = n_iter & 1;
            ulonglong2 *cq = &S.fq[0][0];
            if (tot) {
                u32 base = 0;
                if (lane == 0) base = atomicAdd(&S.fqn[pb], tot);
                base = __shfl_sync(0xFFFFFFFFu, base, 0);
#pragma unroll
                for (int j = 0; j < CPT; ++j) {
                    if (act[j])
                        cq[base + __popc(am[j] & ((1u << lane) - 1u))] =
                            make_ulonglong2((u64)gv[j], (u64)F[j] | ((u64)(u32)C::own_hi(tid, j) << 32));
                    base += __popc(am[j]);
                }
            }
            __syncthreads();
            const u32 n = S.fqn[pb];
            if (tid == 0) S.fqn[pb ^ 1] = 0;
            for (u32 r = tid >> 5; r < n; r += NT / 32)
                relax_record<C>(S, A, cq[r], my_need, my_w, my_off, U, use_h, ox - gx, oy - gy, oz - gz, my_edges);
            vote = n != 0;
            if (vote) __syncthreads(); // every relaxation lands before the next scan
        } else {
            if (tot) {
                u32 base = 0;
#pragma unroll
                for (int j = 0; j < CPT; ++j) {
                    if (act[j])
                        wq[base + __popc(am[j] & ((1u << lane) - 1u))] =
                            make_ulonglong2((u64)gv[j], (u64)F[j] | ((u64)(u32)C::own_hi(tid, j) << 32));
                    base += __popc(am[j]);
                }
                __syncwarp();
                for (u32 r = grp; r < tot; r += ngrp)
                    relax_record<C>(S, A, wq[r], my_need, my_w, my_off, U, use_h, ox - gx, oy - gy, oz - gz, my_edges);
            }
            vote = __syncthreads_or(anyact);
        }
        ++n_iter;
        if (vote) {
            const u64 nb = W::key_add(B, A.sub);
            B = nb < T ? nb : T;
            continue;
        }
        // nothing was expanded anywhere: the values are stable, take the exact min pending key
        const int e = n_empty & 1;
        ++n_empty;
        if (__any_sync(0xFFFFFFFFu, lmin != KEY_INF)) {
            lmin = warp_min_u64(lmin);
            if (lane == 0) atomicMin(&S.kmin[e], lmin);
        }
        __syncthreads();
        const u64 km = S.kmin[e];
        if (tid == 0) S.kmin[e ^ 1] = KEY_INF;
        if (km >= T) { km_final = km; break; }
        const u64 nb = W::key_add(km, A.sub);
        B = nb < T ? nb : T;
    }

    const u32 t_looped = (u32)clock64();
    // ---- publish.  Values and pending flags travel in one word, so one atomic per cell:
    //  * own cells that were lowered: atomicMin(value, flag) -- a neighbour tile may have lowered
    //    a border cell meanwhile, then its (lower, pending) word stays;
    //  * own cells that were pending at load, have propagated now and kept their value: CAS
    //    pending -> clean (fails harmlessly if somebody lowered the cell in between);
    //  * lowered halo cells: atomicMin(value, pending) into the neighbour tile.
#pragma unroll
    for (int j = 0; j < CPT; ++j) {
        const int hj = C::own_hi(tid, j);
        const G v = W::dec(S.g[hj]);
        const bool pending = v < gprop[j];
        const bool changed = (S.chg[hj >> 5] >> (hj & 31)) & 1u;
        if (changed || ((ldp >> j) & 1u)) {
            int lx, ly, lz;
            C::owned(tid, j, lx, ly, lz);
            G *addr = gq + ((size_t)(oz + lz) * Gd.PY + (oy + ly)) * Gd.PX + (ox + lx);
            if (changed) red_min(addr, W::enc(v, pending));
            else if (!pending) cas_word(addr, W::enc(v, true), W::enc(v, false));
        }
    }
#pragma unroll
    for (int k = 0; k < NLD; ++k) {
        const int i = tid + k * NT;
        if (i < C::HC && ((S.chg[i >> 5] >> (i & 31)) & 1u)) {
            const int hx = i % C::HX, hy = (i / C::HX) % C::HY, hz = i / (C::HX * C::HY);
            const int sx = hx == 0 ? -1 : (hx == C::HX - 1 ? 1 : 0);
            const int sy = hy == 0 ? -1 : (hy == C::HY - 1 ? 1 : 0);
            const int sz = C::ZH ? (hz == 0 ? -1 : (hz == C::HZ - 1 ? 1 : 0)) : 0;
            if (sx | sy | sz) { // a halo cell
                const int x = ox + hx - 1, y = oy + hy - 1, z = oz + hz - C::ZH;
                const G v = W::dec(S.g[i]);
                red_min(gq + ((size_t)z * Gd.PY + y) * Gd.PX + x, W::enc(v, true));
                const double hv = use_h ? heuristic_lo(x - gx, y - gy, z - gz, A.he, A.hm) : 0.0;
                atomicMin(&S.nbrmin[(sz + 1) * 9 + (sy + 1) * 3 + (sx + 1)], W::key(v, hv));
            }
        }
    }
    for (int o = 16; o; o >>= 1) {
        my_edges += __shfl_down_sync(0xFFFFFFFFu, my_edges, o);
        n_exp += __shfl_down_sync(0xFFFFFFFFu, n_exp, o);
    }
    if (lane == 0) {
        if (my_edges) atomicAdd(&S.n_edge, my_edges);
        if (n_exp) atomicAdd(&S.n_exp, n_exp);
    }
    __syncthreads();
    if (tid == 0) fence_gpu(); // cumulative: every word above is visible before a tile is armed
    __syncthreads();
    // arm the tiles that received work
    if (tid < 27) {
        const u64 k = S.nbrmin[tid];
        if (k != KEY_INF) {
            const int sx = tid % 3 - 1, sy = (tid / 3) % 3 - 1, sz = tid / 9 - 1;
            const u32 ngt = q * Gd.ntiles + (u32)(((ttz + sz) * Gd.nty + (tty + sy)) * Gd.ntx + (ttx + sx));
            const u64 was = atomicMin(&A.minkey[ngt], k);
            if (k < Tw) {
                if (was >= Tw) ring_push(A, cur, ngt, k, Tw); // enters the window now: not queued yet (see the claim)
                else if (A.nbands > 1 && band_of(A, k, Tw) < band_of(A, was, Tw))
                    ring_push(A, cur, ngt, k, Tw);            // queued in a later band: a second, earlier entry
            }
        }
    }
    if (tid == 32) {
        if (km_final != KEY_INF) {
            const u64 was = atomicMin(&A.minkey[gt], km_final);
            // banded mode: the rest of this tile's window keys belongs to a later band
            if (km_final < Tw && was >= Tw) ring_push(A, cur, gt, km_final, Tw);
        }
        const u32 t_end = (u32)clock64();
        S.stat[1] += 1; S.stat[2] += S.n_exp; S.n_exp = 0;   // per-CTA counters, flushed once at kernel end
        S.stat[3] += S.n_edge; S.n_edge = 0;
        S.stat[5] += (u64)(t_loaded - t_begin);
        S.stat[6] += (u64)(t_looped - t_loaded);
        S.stat[7] += (u64)(t_end - t_looped);
        S.stat[12] += n_iter;
    }
    __syncthreads();
}

// warp 0 of the CTA: take the next tile from the lowest non-empty band (S.item = 0: round over)
template <class C>
__device__ void banded_pop(const SsspArgs &A, const int cur, TileSmem<C> &S) {
    const int lane = threadIdx.x & 31;
    u32 *c = A.rb + cur * 96;
    u32 v = 0;
    int bsel = 0;
    unsigned ns = 20, spins = 0;
    for (;;) {
        const int av = lane < A.nbands ? (int)ld_cg(&c[lane]) : 0;
        const u32 m = __ballot_sync(0xFFFFFFFFu, av > 0);
        if (m) {
            const int b = __ffs(m) - 1;
            u32 ticket = 0xFFFFFFFFu;
            if (lane == b) {
                if ((int)atomicSub(&c[b], 1u) > 0) ticket = atomicAdd(&c[32 + b], 1u); // reserved, then numbered
                else atomicAdd(&c[b], 1u);                                             // lost the race: give it back
            }
            ticket = __shfl_sync(0xFFFFFFFFu, ticket, b);
            if (ticket == 0xFFFFFFFFu) continue;
            if (lane == 0) {
                u32 *slot = &A.wl[0][(size_t)b * A.ring_cap + ticket % A.ring_cap];
                unsigned w = 0;
                while ((v = ld_cg(slot)) == 0 && ++w < (1u << 24) && !ld_cg(&A.ctr[8])) { } // the push of this ticket is in flight
                if (v) { fence_gpu(); *(volatile u32 *)slot = 0; }
                else atomicOr(&A.ctr[8], 1u); // timed out: abort the search (k_sssp_finish reports PMP_ST_ERROR)
            }
            bsel = b;
            break;
        }
        int fl = 0;
        if (lane == 0) fl = ld_cg(&A.ctr[6 + cur]) == 0 || ld_cg(&A.ctr[8]) != 0;
        if (__shfl_sync(0xFFFFFFFFu, fl, 0)) break; // nothing queued or being processed (or aborted): nobody can append
        __nanosleep(ns);
        if (ns < 160) ns *= 2;
        if (++spins > (1u << 22)) { if (lane == 0) atomicOr(&A.ctr[8], 1u); break; } // safety net (seconds): abort
    }
    if (lane == 0) {
        S.item = v;
        const u32 q = v ? (v - 1u) / A.G.ntiles : 0u;
        S.Tlim = v ? band_top(A, bsel, ld_cg(&A.T[cur][q])) : KEY_INF;
    }
}

// ---------------------------------------------------------------------------------------------
// persistent round loop
// ---------------------------------------------------------------------------------------------
// MINB = resident CTAs per SM the register budget is sized for: 2 (128 registers, no spills,
// fastest single activation -- latency-bound single queries) or 3-4 (80 / 64 registers, more tiles in
// flight -- batches and full fields, which are throughput-bound; the launcher uses 4 there).
template <class C, int MINB>
__global__ void __launch_bounds__(NT, MINB) k_sssp(const __grid_constant__ SsspArgs A) {
    typedef typename C::W W;
    typedef typename C::G G;
    cg::grid_group grid = cg::this_grid();
    __shared__ TileSmem<C> S;
    const int tid = threadIdx.x, lane = tid & 31;
    const u32 ntiles = A.G.ntiles;
    const u32 total = A.nq * ntiles;
    const u32 gstride = gridDim.x * NT;
    if (tid < 16) S.stat[tid] = 0;
    if (tid == 0) { S.n_edge = 0; S.n_exp = 0; }
    const bool timer = blockIdx.x == 0 && tid == 0;
    if (timer) { S.tm[0] = globaltimer_ns(); S.tm[1] = S.tm[2] = S.tm[3] = 0; }

    for (int r = 0;; ++r) {
        const int cur = r & 1, nxt = cur ^ 1;
        if (timer) S.tm[4] = globaltimer_ns();
        // ---- SCAN: which (query, tile) pairs fall inside their query's key window ----
        if (tid == 0) S.s_live = 0;
        __syncthreads();
        u32 my_live = 0;
        for (u32 base = blockIdx.x * NT; base < total; base += gstride) {
            const u32 gt = base + tid;
            bool live = false, take = false;
            u64 kk = KEY_INF, tw = 0;
            if (gt < total) {
                const u64 k = ld_cg(&A.minkey[gt]);
                if (k != KEY_INF) {
                    const u32 q = gt / ntiles;
                    if (ld_cg(&A.done[q]) == 0) {
                        if (k < ld_cg(&A.qmin[cur][q])) atomicMin(&A.qmin[cur][q], k);
                        if (k < A.key_cap) {
                            live = true;
                            kk = k;
                            tw = ld_cg(&A.T[cur][q]);
                            take = k < tw;
                        }
                    }
                }
            }
            my_live += live;
            if (A.nbands > 1) { // banded ring: every tile goes to the band of its key
                if (take) ring_push(A, cur, gt, kk, tw);
                continue;
            }
            const u32 m = __ballot_sync(0xFFFFFFFFu, take);
            if (m) {
                u32 pos = 0;
                if (lane == 0) {
                    atomicAdd(&A.ctr[6 + cur], (u32)__popc(m)); // tiles in flight
                    pos = atomicAdd(&A.ctr[cur], (u32)__popc(m));
                }
                pos = __shfl_sync(0xFFFFFFFFu, pos, 0);
                if (take) A.wl[0][(pos + __popc(m & ((1u << lane) - 1u))) % A.ring_cap] = gt + 1u;
            }
        }
        if (my_live) atomicAdd(&S.s_live, my_live);
        __syncthreads();
        if (tid == 0 && S.s_live) atomicAdd(&A.ctr[4 + cur], S.s_live);
        grid.sync();

        const u32 live = ld_cg(&A.ctr[4 + cur]);
        if (timer) { const u64 tb = globaltimer_ns(); S.tm[1] += tb - S.tm[4]; S.tm[4] = tb; }
        if (ld_cg(&A.ctr[8])) break; // a ring wait timed out in the last round (uniform: read after the grid barrier)
        if (live == 0) {
            // nothing pending below the key cap.  Without a cap that means every query still
            // running has settled all it can reach (k_sssp_finish turns it into EXHAUSTED / REACHED).
            if (blockIdx.x == 0 && tid == 0 && A.key_cap == KEY_INF) A.ctr[9] = 1u;
            break;
        }
        if (A.max_rounds > 0 && r >= A.max_rounds) break;

        // ---- per-query bookkeeping: termination, next window ----
        for (u32 q = blockIdx.x * NT + tid; q < A.nq; q += gstride) {
            const u64 k = ld_cg(&A.qmin[cur][q]);
            if (ld_cg(&A.done[q]) == 0) {
                int st = 0;
                if (k == KEY_INF) {
                    st = PMP_ST_EXHAUSTED;
                } else if (!A.full_field) {
                    const u32 gc = A.goalcell[q];
                    if (gc != 0xFFFFFFFFu) {
                        const G gg = W::dec(ld_cg((const G *)A.g + (size_t)q * A.G.pcells + gc));
                        // every pending key is >= k, keys never decrease along an edge: g[goal] is final
                        if (gg != W::inf() && W::key(gg, 0.0) <= k) st = PMP_ST_REACHED;
                    }
                }
                if (st) A.done[q] = st;
            }
            u64 tn = (k == KEY_INF) ? KEY_INF : W::key_add(k, A.window);
            A.T[nxt][q] = tn < A.key_cap ? tn : A.key_cap;
            A.qmin[nxt][q] = KEY_INF;
        }
        if (blockIdx.x == 0 && tid == 0) {
            A.ctr[nxt] = 0; A.ctr[2 + nxt] = 0; A.ctr[4 + nxt] = 0; A.ctr[6 + nxt] = 0;
            A.stats[0] = (u64)(r + 1);
        }
        if (A.nbands > 1 && blockIdx.x == 0 && tid < 96) A.rb[nxt * 96 + tid] = 0;

        // ---- PROCESS: asynchronous cascade.  The ring holds the tiles whose smallest pending key
        // lies inside the window; a tile that pushes such a key into a neighbour appends the
        // neighbour itself (process_tile), so the wave crosses many tiles per round without a
        // grid barrier.  The round ends when no tile is queued or being processed.
        for (;;) {
            if (A.nbands > 1) {
                if (tid < 32) banded_pop<C>(A, cur, S);
            } else if (tid == 0) {
                S.Tlim = KEY_INF;
                const u32 ticket = atomicAdd(&A.ctr[2 + cur], 1u);
                u32 *slot = &A.wl[0][ticket % A.ring_cap];
                u32 v;
                unsigned ns = 20, spins = 0;
#ifdef PMP_POPSTAT
                const long long pt0 = clock64();
#endif
                for (;;) {
                    v = ld_cg(slot);
                    if (v) break;
                    if (ld_cg(&A.ctr[6 + cur]) == 0 || ld_cg(&A.ctr[8])) break; // nothing in flight (or aborted): nobody can append
                    __nanosleep(ns);
                    if (ns < 160) ns *= 2;
                    if (++spins > (1u << 22)) { atomicOr(&A.ctr[8], 1u); break; } // safety net (seconds): abort
                }
#ifdef PMP_POPSTAT
                if (v) { S.stat[13] += (u64)(clock64() - pt0); S.stat[14] += spins == 0; S.stat[15] += 1; }
#endif
                if (v) { fence_gpu(); *(volatile u32 *)slot = 0; } // acquire, then recycle the slot
                S.item = v;
            }
            __syncthreads();
            const u32 it = S.item;
            __syncthreads();
            if (it == 0) break;
            process_tile<C>(A, it - 1u, cur, S);
            if (tid == 0) atomicSub(&A.ctr[6 + cur], 1u); // after this tile's own appends
        }
        if (timer) { const u64 tc = globaltimer_ns(); S.tm[2] += tc - S.tm[4]; S.tm[4] = tc; }
        grid.sync();
        if (timer) {
            S.tm[3] += globaltimer_ns() - S.tm[4];
            const u64 nwl = ld_cg(&A.ctr[cur]);
            if (nwl > A.stats[4]) A.stats[4] = nwl;
        }
    }
    __syncthreads();
    if (tid < 15 && tid != 0 && tid != 4 && (tid < 8 || tid >= 12) && S.stat[tid]) atomicAdd(&A.stats[tid], S.stat[tid]); // [15] = error flag
    if (timer) {
        A.stats[8] = globaltimer_ns() - S.tm[0];
        A.stats[9] = S.tm[1]; A.stats[10] = S.tm[2]; A.stats[11] = S.tm[3];
    }
}

// ---------------------------------------------------------------------------------------------
// init / finalize / statistics kernels
// ---------------------------------------------------------------------------------------------
template <class C>
__global__ void k_sssp_init(SsspArgs A) {
    typedef typename C::W W;
    typedef typename C::G G;
    const GridDev &Gd = A.G;
    for (u32 q = blockIdx.x * blockDim.x + threadIdx.x; q < A.nq; q += gridDim.x * blockDim.x) {
        const int sx = A.starts[3 * q], sy = A.starts[3 * q + 1], sz = Gd.ndim == 3 ? A.starts[3 * q + 2] : 0;
        bool ok = sx >= 0 && sy >= 0 && sz >= 0 && sx < Gd.X && sy < Gd.Y && sz < Gd.Z && !cell_blocked(Gd, sx, sy, sz);
        u32 gc = 0xFFFFFFFFu;
        int gx = 0, gy = 0, gz = 0;
        if (A.goals && !A.full_field) {
            gx = A.goals[3 * q]; gy = A.goals[3 * q + 1]; gz = Gd.ndim == 3 ? A.goals[3 * q + 2] : 0;
            if (gx >= 0 && gy >= 0 && gz >= 0 && gx < Gd.X && gy < Gd.Y && gz < Gd.Z)
                gc = (u32)(((size_t)gz * Gd.PY + gy) * Gd.PX + gx);
            else
                ok = false;
        }
        A.goalcell[q] = gc;
        A.qmin[0][q] = KEY_INF; A.qmin[1][q] = KEY_INF;
        A.T[0][q] = KEY_INF; A.T[1][q] = KEY_INF;
        if (!ok) { A.done[q] = PMP_ST_BAD_QUERY; continue; }
        A.done[q] = 0;
        G *gq = (G *)A.g + (size_t)q * Gd.pcells;
        gq[((size_t)sz * Gd.PY + sy) * Gd.PX + sx] = W::enc((G)0, true);
        const int tx = sx / C::TX, ty = sy / C::TY, tz = sz / C::TZ;
        const u32 t = (u32)((tz * Gd.nty + ty) * Gd.ntx + tx);
        const u32 gt = q * Gd.ntiles + t;
        const double h = A.use_h ? heuristic_lo(sx - gx, sy - gy, sz - gz, A.he, A.hm) : 0.0;
        const u64 k0 = W::key((G)0, h);
        A.minkey[gt] = k0;
        A.T[0][q] = W::key_add(k0, A.window);
    }
}

template <class W>
__global__ void k_sssp_finish(SsspArgs A, double *gcost, int32_t *status) {
    typedef typename W::G G;
    for (u32 q = blockIdx.x * blockDim.x + threadIdx.x; q < A.nq; q += gridDim.x * blockDim.x) {
        int st = A.done[q];
        if (st == PMP_ST_RUNNING && A.ctr[9]) st = PMP_ST_EXHAUSTED; // ran out of pending keys (REACHED below if the goal has a value)
        double c = __longlong_as_double(0x7ff0000000000000LL);
        const u32 gc = A.goalcell[q];
        if (st != PMP_ST_BAD_QUERY && gc != 0xFFFFFFFFu) {
            const G gg = W::dec(((const G *)A.g)[(size_t)q * A.G.pcells + gc]);
            if (gg != W::inf()) {
                c = W::to_double(gg);
                if (st == PMP_ST_EXHAUSTED) st = PMP_ST_REACHED;
            } else if (st == PMP_ST_REACHED) {
                st = PMP_ST_EXHAUSTED;
            }
        }
        if (A.ctr[8] && st != PMP_ST_BAD_QUERY) { // aborted: whatever g[goal] holds is not trustworthy
            st = PMP_ST_ERROR;
            c = __longlong_as_double(0x7ff0000000000000LL);
        }
        if (gcost) gcost[q] = c;
        if (status) status[q] = st;
    }
    if (blockIdx.x == 0 && threadIdx.x == 0) A.stats[15] = A.ctr[8];
}

__global__ void k_field_finalize(u64 *g, size_t n) {
    for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x) {
        const u64 s = g[i];
        g[i] = s == WF64::ginf() ? 0x7ff0000000000000ull : WF64::dec(s);
    }
}
__global__ void k_field_finalize_u32(u32 *g, size_t n) {
    for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x) {
        const u32 s = g[i];
        g[i] = s == WU32::ginf() ? 0xFFFFFFFFu : WU32::dec(s);
    }
}

template <class W>
__global__ void k_count_settled(GridDev Gd, u32 nq, const typename W::G *g, const int32_t *goals,
                                const double *gcost, double he, double hm, u64 *counts) {
    typedef typename W::G G;
    const u32 q = blockIdx.y;
    if (q >= nq) return;
    const G *gq = g + (size_t)q * Gd.pcells;
    double bound = 0.0;
    int gx = 0, gy = 0, gz = 0;
    const bool use_bound = gcost != nullptr && goals != nullptr;
    if (use_bound) {
        bound = gcost[q];
        gx = goals[3 * q]; gy = goals[3 * q + 1]; gz = Gd.ndim == 3 ? goals[3 * q + 2] : 0;
    }
    u64 local = 0;
    for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < Gd.pcells; i += (size_t)gridDim.x * blockDim.x) {
        const G v = W::dec(gq[i]);
        if (v == W::inf()) continue;
        if (use_bound) {
            const int x = (int)(i % Gd.PX), y = (int)((i / Gd.PX) % Gd.PY), z = (int)(i / ((size_t)Gd.PX * Gd.PY));
            const double f = W::to_double(v) + heuristic_lo(x - gx, y - gy, z - gz, he, hm);
            if (f > bound) continue;
        }
        ++local;
    }
    for (int o = 16; o; o >>= 1) local += __shfl_down_sync(0xFFFFFFFFu, local, o);
    if ((threadIdx.x & 31) == 0 && local) atomicAdd(&counts[q], local);
}

// ---------------------------------------------------------------------------------------------
// host side
// ---------------------------------------------------------------------------------------------
typedef TileCfg<32, 32, 1, WF64> C2F;
typedef TileCfg<32, 32, 1, WU32> C2U;
typedef TileCfg<16, 8, 8, WF64> C3F;
typedef TileCfg<16, 8, 8, WU32> C3U;

static inline size_t align_up(size_t v, size_t a) { return (v + a - 1) / a * a; }

// The ring has more slots than tiles + resident CTAs, so a waiting consumer ticket can never alias
// the slot of a queued tile (at most one ticket per CTA is outstanding beyond the tail).
#define RING_SLACK 1024
struct WsLayout {
    size_t minkey, wl0, wl1, qmin0, qmin1, T0, T1, done, goalcell, ctr, stats, rb, total;
    size_t ring_words; // u32 words of the ring area (all bands)
};

static WsLayout ws_layout(const pmp_grid *g, int64_t nq) {
    WsLayout L;
    const size_t nt = (size_t)nq * g->d.ntiles;
    size_t o = 0;
    L.minkey = o; o = align_up(o + nt * 8, 256);
    // a single query may run the banded ring: one ring per band
    L.ring_words = (nt + RING_SLACK) * (size_t)(nq == 1 ? PMP_MAXBANDS : 1);
    L.wl0 = o; o = align_up(o + L.ring_words * 4, 256);
    L.wl1 = o; o = align_up(o + nt * 4, 256);
    L.qmin0 = o; o = align_up(o + (size_t)nq * 8, 256);
    L.qmin1 = o; o = align_up(o + (size_t)nq * 8, 256);
    L.T0 = o; o = align_up(o + (size_t)nq * 8, 256);
    L.T1 = o; o = align_up(o + (size_t)nq * 8, 256);
    L.done = o; o = align_up(o + (size_t)nq * 4, 256);
    L.goalcell = o; o = align_up(o + (size_t)nq * 4, 256);
    L.ctr = o; o = align_up(o + 16 * 4, 256);
    L.stats = o; o = align_up(o + 16 * 8, 256);
    L.rb = o; o = align_up(o + 2 * 96 * 4, 256);
    L.total = o;
    return L;
}

size_t pmp_bfs_ws_bytes(const pmp_grid *g);

extern "C" int pmp_sssp_sizes(const pmp_grid *g, int64_t nq, int weight, int64_t *g_bytes_per_query,
                              int64_t *ws_bytes) {
    if (!g || nq < 1 || (weight != PMP_W_F64 && weight != PMP_W_U32)) {
        pmp_set_error("pmp_sssp_sizes: bad arguments");
        return -1;
    }
    if ((u64)nq * g->d.ntiles >= 0xFFFFFFFFull) {
        pmp_set_error("nq * tiles = %llu exceeds 2^32-2; split the batch", (unsigned long long)((u64)nq * g->d.ntiles));
        return -1;
    }
    if (g_bytes_per_query) *g_bytes_per_query = (int64_t)(g->d.pcells * (weight == PMP_W_F64 ? 8 : 4));
    if (ws_bytes) {
        size_t need = ws_layout(g, nq).total;
        if (g->d.ndim == 3 && weight == PMP_W_U32) { // a full unit-weight field is routed to k_bfs, which lays the workspace out itself
            const size_t b = pmp_bfs_ws_bytes(g);
            need = b > need ? b : need;
        }
        *ws_bytes = (int64_t)need;
    }
    return 0;
}

template <class C, int MINB>
static int launch_coop(SsspArgs &A, cudaStream_t s) {
    static int blocks_per_sm = 0, n_sm = 0;
    if (!blocks_per_sm) {
        PMP_CHECK_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&blocks_per_sm, k_sssp<C, MINB>, NT, 0));
        int dev = 0;
        PMP_CHECK_CUDA(cudaGetDevice(&dev));
        PMP_CHECK_CUDA(cudaDeviceGetAttribute(&n_sm, cudaDevAttrMultiProcessorCount, dev));
        if (blocks_per_sm < 1) { pmp_set_error("k_sssp does not fit on an SM"); return -2; }
        const char *e = getenv("PMP_B200_CTAS_PER_SM"); // tuning: fewer resident CTAs than fit
        if (e && atoi(e) >= 1 && atoi(e) < blocks_per_sm) blocks_per_sm = atoi(e);
    }
    // persistent grid: every CTA resident (cooperative launch), a multiple of the SM count
    dim3 grid((unsigned)(blocks_per_sm * n_sm)), block(NT);
    void *args[] = {(void *)&A};
    PMP_CHECK_CUDA(cudaLaunchCooperativeKernel((void *)k_sssp<C, MINB>, grid, block, args, 0, s));
    return 0;
}

template <class C>
static int launch_sssp(pmp_grid *g, SsspArgs &A, double *gcost_dev, int32_t *status_dev, cudaStream_t s, int mode) {
    typedef typename C::W W;
    const unsigned ib = (unsigned)((A.nq + 255) / 256);
    if (mode & 1) {
        k_sssp_init<C><<<ib > 1024 ? 1024 : ib, 256, 0, s>>>(A);
        PMP_CHECK_CUDA(cudaGetLastError());
    }
    if (mode & 2) {
        const char *e = getenv("PMP_B200_MINB");
        const bool throughput = e ? atoi(e) >= 3 : (A.full_field || A.nq >= 8);
        const int rc = throughput ? launch_coop<C, PMP_TP_MINB>(A, s) : launch_coop<C, 2>(A, s);
        if (rc) return rc;
    }
    if (mode & 4) {
        k_sssp_finish<W><<<ib > 1024 ? 1024 : ib, 256, 0, s>>>(A, gcost_dev, status_dev);
        PMP_CHECK_CUDA(cudaGetLastError());
    }
    return 0;
}

static int dispatch_sssp(pmp_grid *g, int weight, SsspArgs &A, double *gcost_dev, int32_t *status_dev,
                         cudaStream_t s, int mode) {
    if (g->d.ndim == 2)
        return weight == PMP_W_F64 ? launch_sssp<C2F>(g, A, gcost_dev, status_dev, s, mode)
                                   : launch_sssp<C2U>(g, A, gcost_dev, status_dev, s, mode);
    return weight == PMP_W_F64 ? launch_sssp<C3F>(g, A, gcost_dev, status_dev, s, mode)
                               : launch_sssp<C3U>(g, A, gcost_dev, status_dev, s, mode);
}

static void fill_args(pmp_grid *g, const pmp_sssp_params *P, int64_t nq, const int32_t *starts_dev,
                      const int32_t *goals_dev, void *g_dev, void *ws_dev, const WsLayout &L, SsspArgs &A,
                      int no_prune = 0) {
    char *ws = (char *)ws_dev;
    memset(&A, 0, sizeof A);
    A.key_cap = KEY_INF;
    A.no_prune = no_prune;

    A.G = g->d;
    for (int d = 0; d < 27; ++d) { A.wdir[d] = ~0ull; A.need[d] = 0xFFFFFFFFu; }
    for (int m = 0; m < g->mot.k; ++m) {
        const int d = (g->mot.dz[m] + 1) * 9 + (g->mot.dy[m] + 1) * 3 + (g->mot.dx[m] + 1);
        u64 w;
        if (P->weight == PMP_W_F64) memcpy(&w, &g->mot.w[m], 8);
        else w = g->mot.wi[m];
        if (A.need[d] == 0xFFFFFFFFu || w < A.wdir[d]) A.wdir[d] = w; // duplicate offsets: cheapest
        A.need[d] = g->mot.need[m];
    }
    A.ndir = 0;
    for (int i = 0; i < 32; ++i) A.dirs[i] = 13; // centre: need = all-ones
    for (int d = 0; d < 27; ++d)
        if (A.need[d] != 0xFFFFFFFFu) A.dirs[A.ndir++] = (unsigned char)d;
    A.dpl_log2 = A.ndir <= 8 ? 3 : (A.ndir <= 16 ? 4 : 5);
    A.bfs = P->weight == PMP_W_U32 && !getenv("PMP_B200_NO_BFS");
    for (int m = 0; m < g->mot.k; ++m) {
        const int nax = (g->mot.dx[m] != 0) + (g->mot.dy[m] != 0) + (g->mot.dz[m] != 0);
        if (nax != 1 || g->mot.wi[m] != 1) A.bfs = 0;
    }
    A.nq = (u32)nq;
    A.ring_cap = (u32)((size_t)nq * g->d.ntiles + RING_SLACK);
    A.starts = starts_dev;
    A.goals = goals_dev;
    A.g = g_dev;
    A.minkey = (u64 *)(ws + L.minkey);
    A.wl[0] = (u32 *)(ws + L.wl0); A.wl[1] = (u32 *)(ws + L.wl1);
    A.qmin[0] = (u64 *)(ws + L.qmin0); A.qmin[1] = (u64 *)(ws + L.qmin1);
    A.T[0] = (u64 *)(ws + L.T0); A.T[1] = (u64 *)(ws + L.T1);
    A.done = (int *)(ws + L.done);
    A.goalcell = (u32 *)(ws + L.goalcell);
    A.ctr = (u32 *)(ws + L.ctr);
    A.stats = (u64 *)(ws + L.stats);
    A.rb = (u32 *)(ws + L.rb);
    A.full_field = P->full_field ? 1 : 0;
    A.max_rounds = P->max_rounds;
    A.sub = g->wmin;
    // The heuristic only orders and prunes; it must be admissible for the motion table, so the
    // requested type is scaled by the largest admissible factor (manhattan on 8/26-connected
    // grids over-estimates in the reference, graph_search.py:41-42 -- see DESIGN.md).
    A.use_h = (P->algo == PMP_ASTAR && !A.full_field) ? 1 : 0;
    if (A.use_h) A.bfs = 0; // the unit-weight solver orders by g alone
    A.he = g->he_scale;
    A.hm = g->hm_scale;
    if (P->weight == PMP_W_U32) { // integer keys: floor the factors so h stays an integer lower bound
        A.he = 0.0;
        A.hm = floor(g->hm_scale);
        A.sub = floor(g->wmin) < 1.0 ? 1.0 : floor(g->wmin);
    }
    if (P->reserved > 1) A.sub *= (double)P->reserved; // wider in-tile sub-buckets (tuning knob)
    // 2D fp64 batches / fields: a 32 x 32 tile spans ~45 sub-buckets of width wmin with ~9 frontier cells
    // each, so the in-tile loop is bound by its barriers; 4 x wmin trades 30 % more expansions for
    // 16 % fewer sub-iterations (c3, 1 024 queries: 137.6 -> 129.0 ms; 8: 128.1, 16: 127.6)
    else if (P->reserved <= 0 && !A.use_h && P->weight == PMP_W_F64 && g->d.ndim == 2) A.sub *= 4.0;
    double window = P->window;
    // Default key window per round.  Inside a round tiles cascade asynchronously, so the window
    // only bounds how far ahead of the smallest pending key work may run: wide for full sweeps
    // (few rounds), narrow for A-star, whose keys g + h span a small range and where everything
    // above the final cost is wasted.  Integer (unit) weights re-relax least when run far ahead.
    // The bit-parallel unit-weight solver settles whole tiles per activation and the ring is FIFO, so
    // the cascade already runs in near-BFS order: no window at all is fastest there (768^3: 22 ms
    // against 31 ms with a 256-wide window).
    if (!(window > 0.0))
        window = A.bfs ? 4.0e9 : (A.use_h ? 6.0 : (P->weight == PMP_W_U32 ? 256.0 : 96.0)) * g->wmin;
    if (P->weight == PMP_W_U32 && window < 1.0) window = 1.0;
    // Single A-star queries on fp64 fields: banded ring (see ring_push).  PMP_B200_BANDS overrides.
    A.nbands = 1;
    // (not for repairable searches: without the g[goal] pruning a wide window is all wasted work)
    if (A.use_h && nq == 1 && P->weight == PMP_W_F64 && !no_prune) {
        const char *e = getenv("PMP_B200_BANDS");
        int nb = e ? atoi(e) : (g->d.ntiles >= 256 ? PMP_DEFAULT_BANDS : 1); // tiny grids: launch-bound, keep the plain ring
        if (nb > PMP_MAXBANDS) nb = PMP_MAXBANDS;
        if (nb > 1) {
            A.nbands = nb;
            if (!(P->window > 0.0) && !getenv("PMP_B200_WINDOW")) window = PMP_BANDED_WINDOW * g->wmin;
        }
    }
    A.inv_band_w = (double)A.nbands / window;
    A.window = window;

}

// pmp_bfs.cu: warp-per-tile solver for full fields on 6-connected unit-weight 3D grids
bool pmp_bfs_applicable(const pmp_grid *g);
int pmp_bfs_field(pmp_grid *g, const int32_t *start_dev, void *g_dev, void *ws_dev, int64_t ws_bytes,
                  double *gcost_dev, int32_t *status_dev, uint64_t *stats_dev, void *stream);

static int sssp_from_scratch(pmp_grid *g, const pmp_sssp_params *P, int64_t nq, const int32_t *starts_dev,
                             const int32_t *goals_dev, void *g_dev, void *ws_dev, int64_t ws_bytes,
                             double *gcost_dev, int32_t *status_dev, uint64_t *stats_dev, void *stream,
                             int no_prune) {
    if (!g || !P || nq < 1 || !starts_dev || !g_dev || !ws_dev) {
        pmp_set_error("pmp_sssp: bad arguments");
        return -1;
    }
    if (!goals_dev && !P->full_field) {
        pmp_set_error("pmp_sssp: goals_dev is NULL but full_field is 0");
        return -1;
    }
    if (P->weight != PMP_W_F64 && P->weight != PMP_W_U32) { pmp_set_error("pmp_sssp: bad weight type"); return -1; }
    if (P->weight == PMP_W_U32 && !g->integer_costs) {
        pmp_set_error("pmp_sssp: PMP_W_U32 needs integer motion costs");
        return -1;
    }
    if (P->algo != PMP_ASTAR && P->algo != PMP_DIJKSTRA) {
        pmp_set_error("pmp_sssp: algo must be PMP_ASTAR or PMP_DIJKSTRA (GBFS: pmp_gbfs)");
        return -1;
    }
    if ((u64)nq * g->d.ntiles >= 0xFFFFFFFFull) { pmp_set_error("pmp_sssp: nq * tiles too large"); return -1; }
    const WsLayout L = ws_layout(g, nq);
    if ((size_t)ws_bytes < L.total) {
        pmp_set_error("pmp_sssp: workspace %lld B < required %zu B", (long long)ws_bytes, L.total);
        return -1;
    }
    cudaStream_t s = (cudaStream_t)stream;
    // One full field, integer words, exactly the six unit moves: the warp-per-tile level solver
    // (pmp_bfs.cu) -- no field fill, no key windows.  PMP_B200_OLD_BFS=1 keeps the tiled kernel.
    if (nq == 1 && P->full_field && P->weight == PMP_W_U32 && !no_prune && pmp_bfs_applicable(g) &&
        !getenv("PMP_B200_OLD_BFS"))
        return pmp_bfs_field(g, starts_dev, g_dev, ws_dev, ws_bytes, gcost_dev, status_dev, stats_dev, stream);
    if (pmp_grid_upload_fence(g, s)) return -2; // k_sssp cannot wait chunk by chunk for a streamed occupancy upload
    SsspArgs A;
    fill_args(g, P, nq, starts_dev, goals_dev, g_dev, ws_dev, L, A, no_prune);
    const size_t gsz = (size_t)nq * g->d.pcells * (P->weight == PMP_W_F64 ? 8 : 4);
    PMP_CHECK_CUDA(cudaMemsetAsync(g_dev, 0xFF, gsz, s));
    const size_t nt = (size_t)nq * g->d.ntiles;
    PMP_CHECK_CUDA(cudaMemsetAsync(A.minkey, 0xFF, nt * 8, s));
    PMP_CHECK_CUDA(cudaMemsetAsync(A.wl[0], 0, L.ring_words * 4, s));
    PMP_CHECK_CUDA(cudaMemsetAsync(A.rb, 0, 2 * 96 * 4, s));
    PMP_CHECK_CUDA(cudaMemsetAsync(A.ctr, 0, 16 * 4, s));
    PMP_CHECK_CUDA(cudaMemsetAsync(A.stats, 0, 16 * 8, s));

    const int rc = dispatch_sssp(g, P->weight, A, gcost_dev, status_dev, s, 7);
    if (rc) return rc;
    if (stats_dev)
        PMP_CHECK_CUDA(cudaMemcpyAsync(stats_dev, A.stats, 16 * 8, cudaMemcpyDeviceToDevice, s));
    return 0;
}

extern "C" int pmp_sssp(pmp_grid *g, const pmp_sssp_params *P, int64_t nq,
                        const int32_t *starts_dev, const int32_t *goals_dev,
                        void *g_dev, void *ws_dev, int64_t ws_bytes,
                        double *gcost_dev, int32_t *status_dev, uint64_t *stats_dev, void *stream) {
    return sssp_from_scratch(g, P, nq, starts_dev, goals_dev, g_dev, ws_dev, ws_bytes, gcost_dev, status_dev,
                             stats_dev, stream, 0);
}

extern "C" int pmp_field_finalize(void *g_dev, int weight, int64_t n_cells, void *stream) {
    if (!g_dev || n_cells < 0 || (weight != PMP_W_F64 && weight != PMP_W_U32)) { pmp_set_error("pmp_field_finalize: bad arguments"); return -1; }
    if (n_cells == 0) return 0;
    size_t b = ((size_t)n_cells + 255) / 256;
    if (b > 148 * 32) b = 148 * 32;
    if (weight == PMP_W_F64) k_field_finalize<<<(unsigned)b, 256, 0, (cudaStream_t)stream>>>((u64 *)g_dev, (size_t)n_cells);
    else k_field_finalize_u32<<<(unsigned)b, 256, 0, (cudaStream_t)stream>>>((u32 *)g_dev, (size_t)n_cells);
    PMP_CHECK_CUDA(cudaGetLastError());
    return 0;
}

extern "C" int pmp_count_settled(const pmp_grid *g, int weight, int64_t nq, const void *g_dev,
                                 const int32_t *goals_dev, const double *gcost_dev, int heuristic,
                                 uint64_t *counts_dev, void *stream) {
    const double he = heuristic < 0 ? 0.0 : g->he_scale, hm = heuristic < 0 ? 0.0 : g->hm_scale;
    if (!g || !g_dev || !counts_dev || nq < 1 || nq > 65535) { pmp_set_error("pmp_count_settled: bad arguments (nq <= 65535)"); return -1; }
    cudaStream_t s = (cudaStream_t)stream;
    PMP_CHECK_CUDA(cudaMemsetAsync(counts_dev, 0, (size_t)nq * 8, s));
    size_t bx = (g->d.pcells + 255) / 256;
    const size_t cap = nq >= 148 ? 8 : (148 * 8) / (size_t)nq;
    if (bx > cap) bx = cap;
    dim3 grid((unsigned)bx, (unsigned)nq);
    if (weight == PMP_W_F64)
        k_count_settled<WF64><<<grid, 256, 0, s>>>(g->d, (u32)nq, (const u64 *)g_dev, goals_dev, gcost_dev, he, hm, (u64 *)counts_dev);
    else
        k_count_settled<WU32><<<grid, 256, 0, s>>>(g->d, (u32)nq, (const u32 *)g_dev, goals_dev, gcost_dev, 0.0, floor(hm), (u64 *)counts_dev);
    PMP_CHECK_CUDA(cudaGetLastError());
    return 0;
}

// ---------------------------------------------------------------------------------------------
// slab mode (one z-slab of a large grid per GPU; the host exchanges boundary planes between
// runs -- python_motion_planning_b200/slab.py).  The handle describes the LOCAL slab including
// its ghost planes; a run processes every key below `key_bound` to local convergence.
// ---------------------------------------------------------------------------------------------
__global__ void k_slab_arm(SsspArgs A, u64 cap) {
    A.qmin[0][0] = KEY_INF; A.qmin[1][0] = KEY_INF;
    A.T[0][0] = cap; A.T[1][0] = cap;
    A.done[0] = 0;
    A.goalcell[0] = 0xFFFFFFFFu;
    for (int i = 0; i < 16; ++i) A.ctr[i] = 0;
    A.stats[0] = 0; // k_slab_export picks the qmin parity from the rounds of THIS run
}

__global__ void k_slab_export(SsspArgs A, u64 *out) {
    // the scan of the last round (parity = rounds completed & 1) saw every key still pending
    out[0] = A.qmin[A.stats[0] & 1][0]; // smallest pending key left in this slab (KEY_INF: none)
    out[1] = A.stats[0];
    out[2] = A.stats[1];
    out[3] = A.stats[2];
    out[5] = A.ctr[8]; // non-zero: a ring wait timed out, the slab state is invalid
}

template <class C>
__global__ void k_slab_merge(SsspArgs A, int z, const typename C::G *plane, int mark, u64 *out) {
    typedef typename C::W W;
    typedef typename C::G G;
    const GridDev &Gd = A.G;
    const size_t n = (size_t)Gd.PX * Gd.PY;
    G *gz = (G *)A.g + (size_t)z * n;
    u64 local_min = KEY_INF;
    u32 improved = 0;
    for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x) {
        const G sw = plane[i];
        if (sw == W::ginf()) continue;
        const G v = W::dec(sw);
        if (v >= W::dec(gz[i])) continue;
        const int x = (int)(i % Gd.PX), y = (int)(i / Gd.PX);
        if (cell_blocked(Gd, x, y, z)) continue;
        gz[i] = W::enc(v, mark != 0);
        ++improved;
        if (!mark) continue;
        const int tx = x / C::TX, ty = y / C::TY, tz = z / C::TZ;
        const u32 t = (u32)((tz * Gd.nty + ty) * Gd.ntx + tx);
        const u64 k = W::key(v, 0.0);
        atomicMin(&A.minkey[t], k);
        local_min = k < local_min ? k : local_min;
    }
    local_min = warp_min_u64(local_min);
    for (int o = 16; o; o >>= 1) improved += __shfl_down_sync(0xFFFFFFFFu, improved, o);
    if ((threadIdx.x & 31) == 0) {
        if (local_min != KEY_INF) atomicMin(&out[0], local_min);
        if (improved) atomicAdd(&out[4], (u64)improved);
    }
}

static int slab_check(pmp_grid *g, const pmp_sssp_params *P, void *g_dev, void *ws_dev, int64_t ws_bytes,
                      WsLayout &L) {
    if (!g || !P || !g_dev || !ws_dev) { pmp_set_error("pmp_slab: bad arguments"); return -1; }
    if (P->weight != PMP_W_F64 && P->weight != PMP_W_U32) { pmp_set_error("pmp_slab: bad weight type"); return -1; }
    if (P->weight == PMP_W_U32 && !g->integer_costs) { pmp_set_error("pmp_slab: PMP_W_U32 needs integer motion costs"); return -1; }
    L = ws_layout(g, 1);
    if ((size_t)ws_bytes < L.total) { pmp_set_error("pmp_slab: workspace too small"); return -1; }
    return 0;
}

extern "C" int pmp_slab_begin(pmp_grid *g, const pmp_sssp_params *P, const int32_t *start_dev,
                              void *g_dev, void *ws_dev, int64_t ws_bytes, void *stream) {
    if (g && pmp_grid_upload_fence((pmp_grid *)g, (cudaStream_t)stream)) return -2; // a streamed occupancy upload must have landed
    WsLayout L;
    if (slab_check(g, P, g_dev, ws_dev, ws_bytes, L)) return -1;
    cudaStream_t s = (cudaStream_t)stream;
    pmp_sssp_params Q = *P;
    Q.full_field = 1;
    Q.algo = PMP_DIJKSTRA;
    SsspArgs A;
    fill_args(g, &Q, 1, start_dev, nullptr, g_dev, ws_dev, L, A);
    PMP_CHECK_CUDA(cudaMemsetAsync(g_dev, 0xFF, (size_t)g->d.pcells * (Q.weight == PMP_W_F64 ? 8 : 4), s));
    PMP_CHECK_CUDA(cudaMemsetAsync(A.minkey, 0xFF, (size_t)g->d.ntiles * 8, s));
    PMP_CHECK_CUDA(cudaMemsetAsync(A.wl[0], 0, L.ring_words * 4, s));
    PMP_CHECK_CUDA(cudaMemsetAsync(A.rb, 0, 2 * 96 * 4, s));
    PMP_CHECK_CUDA(cudaMemsetAsync(A.ctr, 0, 16 * 4, s));
    PMP_CHECK_CUDA(cudaMemsetAsync(A.stats, 0, 16 * 8, s));
    if (start_dev) return dispatch_sssp(g, Q.weight, A, nullptr, nullptr, s, 1);
    k_slab_arm<<<1, 1, 0, s>>>(A, KEY_INF);
    PMP_CHECK_CUDA(cudaGetLastError());
    return 0;
}

extern "C" int pmp_slab_run(pmp_grid *g, const pmp_sssp_params *P, void *g_dev, void *ws_dev,
                            int64_t ws_bytes, double key_bound, uint64_t *out_dev, void *stream) {
    if (g && pmp_grid_upload_fence((pmp_grid *)g, (cudaStream_t)stream)) return -2; // a streamed occupancy upload must have landed
    WsLayout L;
    if (slab_check(g, P, g_dev, ws_dev, ws_bytes, L)) return -1;
    if (!out_dev) { pmp_set_error("pmp_slab_run: out_dev is NULL"); return -1; }
    cudaStream_t s = (cudaStream_t)stream;
    pmp_sssp_params Q = *P;
    Q.full_field = 1;
    Q.algo = PMP_DIJKSTRA;
    SsspArgs A;
    fill_args(g, &Q, 1, nullptr, nullptr, g_dev, ws_dev, L, A);
    u64 cap = KEY_INF;
    if (isfinite(key_bound)) {
        if (Q.weight == PMP_W_F64) memcpy(&cap, &key_bound, 8);
        else cap = (u64)key_bound;
    }
    A.key_cap = cap;
    k_slab_arm<<<1, 1, 0, s>>>(A, cap);
    PMP_CHECK_CUDA(cudaGetLastError());
    const int rc = dispatch_sssp(g, Q.weight, A, nullptr, nullptr, s, 2);
    if (rc) return rc;
    k_slab_export<<<1, 1, 0, s>>>(A, (u64 *)out_dev);
    PMP_CHECK_CUDA(cudaGetLastError());
    return 0;
}

extern "C" int pmp_slab_merge(pmp_grid *g, int weight, void *g_dev, void *ws_dev, int64_t ws_bytes,
                              int z_local, int mark, const void *plane_dev, uint64_t *out_dev, void *stream) {
    if (g && pmp_grid_upload_fence((pmp_grid *)g, (cudaStream_t)stream)) return -2; // a streamed occupancy upload must have landed
    pmp_sssp_params Q;
    memset(&Q, 0, sizeof Q);
    Q.weight = weight; Q.full_field = 1; Q.algo = PMP_DIJKSTRA;
    WsLayout L;
    if (slab_check(g, &Q, g_dev, ws_dev, ws_bytes, L)) return -1;
    if (!plane_dev || !out_dev || z_local < 0 || z_local >= g->d.PZ || g->d.ndim != 3) {
        pmp_set_error("pmp_slab_merge: bad arguments");
        return -1;
    }
    cudaStream_t s = (cudaStream_t)stream;
    SsspArgs A;
    fill_args(g, &Q, 1, nullptr, nullptr, g_dev, ws_dev, L, A);
    const size_t n = (size_t)g->d.PX * g->d.PY;
    unsigned blocks = (unsigned)((n + 255) / 256);
    if (blocks > 148 * 8) blocks = 148 * 8;
    if (weight == PMP_W_F64) k_slab_merge<C3F><<<blocks, 256, 0, s>>>(A, z_local, (const u64 *)plane_dev, mark, (u64 *)out_dev);
    else k_slab_merge<C3U><<<blocks, 256, 0, s>>>(A, z_local, (const u32 *)plane_dev, mark, (u64 *)out_dev);
    PMP_CHECK_CUDA(cudaGetLastError());
    return 0;
}

#include "pmp_replan.cuh"
