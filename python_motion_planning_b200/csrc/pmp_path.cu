// pmp_path.cu -- path extraction by descending the cost field, and the greedy best-first search.
//
// pmp_extract_paths replaces extractPath (global_planner/graph_search/a_star.py:98-117,
// a_star3d.py:91-106): the reference walks parent pointers kept in CLOSED; here the parent of v
// is recomputed as the in-neighbour u minimising g(u) + w(u->v) under the same directed
// no-corner-cutting rule (graph_search.py:61-87, graph_search_3d.py:66-107).  One warp per
// query, one lane per motion.
//
// pmp_gbfs replaces GBFS.plan / GBFS3D.plan (gbfs.py:36-86, gbfs3d.py:34-82).  Greedy
// best-first search is defined by its sequential pop order, so it is run as written: one warp per
// query, lanes evaluate the motions of the popped cell in parallel, lane 0 keeps the binary
// heap with CPython heapq's sift order (the 2D heap is ordered by Node.__lt__,
// utils/environment/node.py:51-54, which is not a total order).
#include <stdlib.h>

#include "pmp_common.cuh"

// 64-bit / 32-bit warp minimum with the hardware reductions (REDUX works on 32-bit words)
__device__ static inline u64 warp_min(u64 v) {
    const u32 hi = (u32)(v >> 32);
    const u32 mh = __reduce_min_sync(0xFFFFFFFFu, hi);
    const u32 lo = hi == mh ? (u32)v : 0xFFFFFFFFu;
    const u32 ml = __reduce_min_sync(0xFFFFFFFFu, lo);
    return ((u64)mh << 32) | ml;
}
__device__ static inline u32 warp_min(u32 v) { return __reduce_min_sync(0xFFFFFFFFu, v); }

template <class W>
__global__ void k_extract_paths(const GridDev G, const __grid_constant__ Motions M, u32 nq,
                                const int32_t *starts, const int32_t *goals,
                                const typename W::G *g, u32 *path, int64_t cap, int32_t *plen,
                                double *cost) {
    typedef typename W::G GT;
    const u32 q = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    const int lane = threadIdx.x & 31;
    if (q >= nq) return;
    const GT *gq = g + (size_t)q * G.pcells;
    const double INF_D = __longlong_as_double(0x7ff0000000000000LL);
    int vx = goals[3 * q], vy = goals[3 * q + 1], vz = G.ndim == 3 ? goals[3 * q + 2] : 0;
    const int sx = starts[3 * q], sy = starts[3 * q + 1], sz = G.ndim == 3 ? starts[3 * q + 2] : 0;
    const bool ok = vx >= 0 && vy >= 0 && vz >= 0 && vx < G.X && vy < G.Y && vz < G.Z &&
                    sx >= 0 && sy >= 0 && sz >= 0 && sx < G.X && sy < G.Y && sz < G.Z;
    if (!ok || gq[((size_t)vz * G.PY + vy) * G.PX + vx] == W::ginf()) {
        if (lane == 0) { plen[q] = 0; cost[q] = INF_D; }
        return;
    }
    // this lane's motion: offset, cost, and the face cells of the SOURCE u = v - offset that
    // the directed rule needs free (graph_search.py:78-86, graph_search_3d.py:97-105); u itself
    // is free iff it has a finite value, v is on the path
    const bool has = lane < M.k;
    const int dx = has ? M.dx[lane] : 0, dy = has ? M.dy[lane] : 0, dz = has ? M.dz[lane] : 0;
    const int nax = (dx != 0) + (dy != 0) + (dz != 0);
    const double mylen = has ? M.len[lane] : 0.0;
    int64_t n = 0;
    double c = 0.0;
    int32_t result = 0;
    for (;;) {
        if (lane == 0 && n < cap) path[(size_t)q * cap + n] = (u32)(((size_t)vz * G.PY + vy) * G.PX + vx);
        ++n;
        if (vx == sx && vy == sy && vz == sz) { result = n <= cap ? (int32_t)n : -1; break; }
        GT cand = W::inf();
        const int ux = vx - dx, uy = vy - dy, uz = vz - dz;
        if (has && (unsigned)ux < (unsigned)G.PX && (unsigned)uy < (unsigned)G.PY && (unsigned)uz < (unsigned)G.PZ) {
            // independent loads: the value of u and the (0, 2 or 3) face cells
            const GT gu = W::dec(gq[((size_t)uz * G.PY + uy) * G.PX + ux]);
            int blocked = 0;
            if (nax >= 2 && !M.cut) {
                if (dx) blocked |= cell_blocked(G, ux + dx, uy, uz);
                if (dy) blocked |= cell_blocked(G, ux, uy + dy, uz);
                if (dz) blocked |= cell_blocked(G, ux, uy, uz + dz);
            }
            if (!blocked && gu != W::inf()) cand = W::add(gu, M, lane);
        }
        const GT best = warp_min(cand);
        if (best == W::inf()) { result = 0; c = INF_D; break; } // cannot happen on a converged field
        const u32 who = __ballot_sync(0xFFFFFFFFu, cand == best);
        const int m = __ffs(who) - 1;
        c += __shfl_sync(0xFFFFFFFFu, mylen, m); // Planner.dist (planner.py:22-23): hypot of the step, summed goal -> start
        vx -= __shfl_sync(0xFFFFFFFFu, dx, m); vy -= __shfl_sync(0xFFFFFFFFu, dy, m); vz -= __shfl_sync(0xFFFFFFFFu, dz, m);
        if (n > (int64_t)G.pcells) { result = 0; c = INF_D; break; }
        // The walk is one dependent L2 round trip per step unless the lines are already in L1:
        // prefetch the rows within distance 3 of the new cell (needed two steps from now).
        const int zr = G.ndim == 3 ? 3 : 0, nrow = 7 * (2 * zr + 1);
        for (int r = lane; r < nrow; r += 32) {
            const int py = vy + r % 7 - 3, pz = vz + r / 7 - zr;
            if ((unsigned)py < (unsigned)G.PY && (unsigned)pz < (unsigned)G.PZ) {
                const size_t row = (size_t)pz * G.PY + py;
                const int x0 = vx - 4 < 0 ? 0 : vx - 4, x1 = vx + 4 >= G.PX ? G.PX - 1 : vx + 4;
                asm volatile("prefetch.global.L1 [%0];" ::"l"(gq + row * G.PX + x0));
                asm volatile("prefetch.global.L1 [%0];" ::"l"(gq + row * G.PX + x1));
                asm volatile("prefetch.global.L1 [%0];" ::"l"(G.bits + row * G.wpr + (vx >> 5)));
            }
        }
    }
    if (lane == 0) { plen[q] = result; cost[q] = c; }
}

// Few queries (one CTA each): the walk is a chain of dependent loads, ~1.2 us per step from
// L2.  Here the CTA stages a 16^3 (2D: 16^2) box of the field and its occupancy rows around the
// current cell in shared memory, warp 0 walks inside it -- same candidates, same tie-break, same
// summation as k_extract_paths, so the path is identical -- and the box is re-centred when the walk
// reaches its border (every ~7 steps).
#define XB 16
template <class W>
__global__ void __launch_bounds__(256) k_extract_paths_box(const GridDev G, const __grid_constant__ Motions M, u32 nq,
                                                           const int32_t *starts, const int32_t *goals,
                                                           const typename W::G *g, u32 *path, int64_t cap,
                                                           int32_t *plen, double *cost) {
    typedef typename W::G GT;
    __shared__ GT box[XB * XB * XB];
    __shared__ u32 occ[XB * XB];   // bit i of row (y, z): cell (bx + i, by + y, bz + z) blocked
    __shared__ int st[8];          // vx, vy, vz, finished
    const u32 q = blockIdx.x;
    const int tid = threadIdx.x, lane = tid & 31;
    const GT *gq = g + (size_t)q * G.pcells;
    const double INF_D = __longlong_as_double(0x7ff0000000000000LL);
    const int BZ = G.ndim == 3 ? XB : 1;
    int vx = goals[3 * q], vy = goals[3 * q + 1], vz = G.ndim == 3 ? goals[3 * q + 2] : 0;
    const int sx = starts[3 * q], sy = starts[3 * q + 1], sz = G.ndim == 3 ? starts[3 * q + 2] : 0;
    const bool ok = vx >= 0 && vy >= 0 && vz >= 0 && vx < G.X && vy < G.Y && vz < G.Z &&
                    sx >= 0 && sy >= 0 && sz >= 0 && sx < G.X && sy < G.Y && sz < G.Z;
    if (!ok || gq[((size_t)vz * G.PY + vy) * G.PX + vx] == W::ginf()) {
        if (tid == 0) { plen[q] = 0; cost[q] = INF_D; }
        return;
    }
    const bool has = lane < M.k;
    const int dx = has ? M.dx[lane] : 0, dy = has ? M.dy[lane] : 0, dz = has ? M.dz[lane] : 0;
    const int nax = (dx != 0) + (dy != 0) + (dz != 0);
    const double mylen = has ? M.len[lane] : 0.0;
    int64_t n = 0;     // warp 0's copy is the authoritative one
    double c = 0.0;
    int32_t result = 0;
    for (;;) {
        // the walk heads for the start: put the current cell near the far side of the box
        const int bx = vx - (sx < vx ? XB - 4 : (sx > vx ? 3 : XB / 2));
        const int by = vy - (sy < vy ? XB - 4 : (sy > vy ? 3 : XB / 2));
        const int bz = G.ndim == 3 ? vz - (sz < vz ? XB - 4 : (sz > vz ? 3 : XB / 2)) : 0;
        for (int i = tid; i < XB * XB * BZ; i += 256) {
            const int x = bx + (i % XB), y = by + (i / XB) % XB, z = bz + i / (XB * XB);
            GT w = W::ginf();
            if ((unsigned)x < (unsigned)G.PX && (unsigned)y < (unsigned)G.PY && (unsigned)z < (unsigned)G.PZ)
                w = gq[((size_t)z * G.PY + y) * G.PX + x];
            box[i] = w;
        }
        for (int r = tid; r < XB * BZ; r += 256) occ[r] = (u32)row_bits(G, bx, by + r % XB, bz + r / XB, XB);
        __syncthreads();
        if (tid < 32) {
            int finished = 0;
            for (;;) {
                const int lx = vx - bx, ly = vy - by, lz = vz - bz;
                const bool inside = lx >= 1 && lx <= XB - 2 && ly >= 1 && ly <= XB - 2 && (G.ndim == 2 || (lz >= 1 && lz <= XB - 2));
                if (!inside) break; // re-centre the box (the first cell of a fresh box is always inside)
                if (lane == 0 && n < cap) path[(size_t)q * cap + n] = (u32)(((size_t)vz * G.PY + vy) * G.PX + vx);
                ++n;
                if (vx == sx && vy == sy && vz == sz) { result = n <= cap ? (int32_t)n : -1; finished = 1; break; }
                GT cand = W::inf();
                if (has) {
                    const int ux = lx - dx, uy = ly - dy, uz = lz - dz;
                    const GT gu = W::dec(box[(uz * XB + uy) * XB + ux]);
                    int blocked = 0;
                    if (nax >= 2 && !M.cut) {
                        if (dx) blocked |= (occ[uz * XB + uy] >> (ux + dx)) & 1u;
                        if (dy) blocked |= (occ[uz * XB + uy + dy] >> ux) & 1u;
                        if (dz) blocked |= (occ[(uz + dz) * XB + uy] >> ux) & 1u;
                    }
                    if (!blocked && gu != W::inf()) cand = W::add(gu, M, lane);
                }
                const GT best = warp_min(cand);
                if (best == W::inf()) { result = 0; c = INF_D; finished = 1; break; }
                const u32 who = __ballot_sync(0xFFFFFFFFu, cand == best);
                const int m = __ffs(who) - 1;
                c += __shfl_sync(0xFFFFFFFFu, mylen, m);
                vx -= __shfl_sync(0xFFFFFFFFu, dx, m); vy -= __shfl_sync(0xFFFFFFFFu, dy, m); vz -= __shfl_sync(0xFFFFFFFFu, dz, m);
                if (n > (int64_t)G.pcells) { result = 0; c = INF_D; finished = 1; break; }
            }
            if (lane == 0) { st[0] = vx; st[1] = vy; st[2] = vz; st[3] = finished; }
        }
        __syncthreads();
        vx = st[0]; vy = st[1]; vz = st[2];
        const int fin = st[3];
        __syncthreads();
        if (fin) break;
    }
    if (tid == 0) { plen[q] = result; cost[q] = c; }
}

extern "C" int pmp_extract_paths(const pmp_grid *g, int weight, int64_t nq,
                                 const int32_t *starts_dev, const int32_t *goals_dev, const void *g_dev,
                                 uint32_t *path_dev, int64_t path_cap, int32_t *path_len_dev,
                                 double *cost_dev, void *stream) {
    if (g && pmp_grid_upload_fence((pmp_grid *)g, (cudaStream_t)stream)) return -2; // a streamed occupancy upload must have landed
    if (!g || nq < 1 || !starts_dev || !goals_dev || !g_dev || !path_dev || path_cap < 1 || !path_len_dev || !cost_dev) {
        pmp_set_error("pmp_extract_paths: bad arguments");
        return -1;
    }
    const unsigned blocks = (unsigned)((nq * 32 + 127) / 128);
    cudaStream_t s = (cudaStream_t)stream;
    if (nq <= 64 && (weight == PMP_W_F64 || weight == PMP_W_U32) && !getenv("PMP_B200_WARP_EXTRACT")) {
        // few queries: one CTA per query with the field staged in shared memory
        if (weight == PMP_W_F64)
            k_extract_paths_box<WF64><<<(unsigned)nq, 256, 0, s>>>(g->d, g->mot, (u32)nq, starts_dev, goals_dev, (const u64 *)g_dev, path_dev, path_cap, path_len_dev, cost_dev);
        else
            k_extract_paths_box<WU32><<<(unsigned)nq, 256, 0, s>>>(g->d, g->mot, (u32)nq, starts_dev, goals_dev, (const u32 *)g_dev, path_dev, path_cap, path_len_dev, cost_dev);
        PMP_CHECK_CUDA(cudaGetLastError());
        return 0;
    }
    if (weight == PMP_W_F64)
        k_extract_paths<WF64><<<blocks, 128, 0, s>>>(g->d, g->mot, (u32)nq, starts_dev, goals_dev, (const u64 *)g_dev, path_dev, path_cap, path_len_dev, cost_dev);
    else if (weight == PMP_W_U32)
        k_extract_paths<WU32><<<blocks, 128, 0, s>>>(g->d, g->mot, (u32)nq, starts_dev, goals_dev, (const u32 *)g_dev, path_dev, path_cap, path_len_dev, cost_dev);
    else { pmp_set_error("pmp_extract_paths: bad weight type"); return -1; }
    PMP_CHECK_CUDA(cudaGetLastError());
    return 0;
}

// ---------------------------------------------------------------------------------------------
// shortest-path certificate (no reference counterpart; used by tests and the slab driver).
// A field g is THE single-source shortest-path field iff g(source) = 0 and every other free
// cell v satisfies g(v) = min over legal in-edges u->v of fl(g(u) + w) (+inf if no in-neighbour
// is reached).  Counts the cells that violate this, the reached cells and the cells that still
// carry a pending flag, for the planes z in [z_lo, z_hi).
// ---------------------------------------------------------------------------------------------
template <class W>
__global__ void k_verify_field(const GridDev G, const __grid_constant__ Motions M,
                               const typename W::G *g, const int32_t *source, int z_lo, int z_hi,
                               u64 *out) {
    typedef typename W::G GT;
    const size_t plane = (size_t)G.PX * G.PY;
    const size_t lo = (size_t)z_lo * plane, hi = (size_t)z_hi * plane;
    int sx = -1, sy = -1, sz = -1;
    if (source) { sx = source[0]; sy = source[1]; sz = G.ndim == 3 ? source[2] : 0; }
    u64 bad = 0, reached = 0, pend = 0;
    for (size_t i = lo + blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < hi; i += (size_t)gridDim.x * blockDim.x) {
        const int x = (int)(i % G.PX), y = (int)((i / G.PX) % G.PY), z = (int)(i / plane);
        const GT sw = g[i];
        const GT v = W::dec(sw);
        if (cell_blocked(G, x, y, z)) { if (sw != W::ginf()) ++bad; continue; }
        if (v != W::inf()) { ++reached; if (!(sw & 1)) ++pend; }
        GT best = W::inf();
        for (int m = 0; m < M.k; ++m) {
            const int ux = x - M.dx[m], uy = y - M.dy[m], uz = z - M.dz[m];
            if (!edge_legal(G, M, m, ux, uy, uz)) continue;
            const GT gu = W::dec(g[((size_t)uz * G.PY + uy) * G.PX + ux]);
            if (gu == W::inf()) continue;
            const GT c = W::add(gu, M, m);
            best = c < best ? c : best;
        }
        if (x == sx && y == sy && z == sz) { if (v != 0) ++bad; }
        else if (v != best) ++bad;
    }
    for (int o = 16; o; o >>= 1) {
        bad += __shfl_down_sync(0xFFFFFFFFu, bad, o);
        reached += __shfl_down_sync(0xFFFFFFFFu, reached, o);
        pend += __shfl_down_sync(0xFFFFFFFFu, pend, o);
    }
    if ((threadIdx.x & 31) == 0) {
        if (bad) atomicAdd(&out[0], bad);
        if (reached) atomicAdd(&out[1], reached);
        if (pend) atomicAdd(&out[2], pend);
    }
}

extern "C" int pmp_verify_field(const pmp_grid *g, int weight, const void *g_dev,
                                const int32_t *source_dev, int z_lo, int z_hi, uint64_t *out_dev,
                                void *stream) {
    if (g && pmp_grid_upload_fence((pmp_grid *)g, (cudaStream_t)stream)) return -2; // a streamed occupancy upload must have landed
    if (!g || !g_dev || !out_dev || z_lo < 0 || z_hi > g->d.PZ || z_lo > z_hi) {
        pmp_set_error("pmp_verify_field: bad arguments");
        return -1;
    }
    cudaStream_t s = (cudaStream_t)stream;
    PMP_CHECK_CUDA(cudaMemsetAsync(out_dev, 0, 3 * 8, s));
    if (z_lo == z_hi) return 0;
    if (weight == PMP_W_F64)
        k_verify_field<WF64><<<148 * 8, 256, 0, s>>>(g->d, g->mot, (const u64 *)g_dev, source_dev, z_lo, z_hi, (u64 *)out_dev);
    else if (weight == PMP_W_U32)
        k_verify_field<WU32><<<148 * 8, 256, 0, s>>>(g->d, g->mot, (const u32 *)g_dev, source_dev, z_lo, z_hi, (u64 *)out_dev);
    else { pmp_set_error("pmp_verify_field: bad weight type"); return -1; }
    PMP_CHECK_CUDA(cudaGetLastError());
    return 0;
}

// ---------------------------------------------------------------------------------------------
// greedy best-first search
// ---------------------------------------------------------------------------------------------
struct __align__(16) HeapRec {
    double h;
    u32 cell;
    u32 aux; // (push counter << 5) | motion index that produced it (31 = start)
};

// 2D: Node.__lt__ with g = 0 (gbfs.py:75): f = h, so a < b iff a.h < b.h.
// 3D: tuple (h, counter) (gbfs3d.py:46,79).
template <int FAMILY>
__device__ static inline bool rec_lt(const HeapRec &a, const HeapRec &b) {
    if (FAMILY == 2) return a.h < b.h;
    if (a.h != b.h) return a.h < b.h;
    return (a.aux >> 5) < (b.aux >> 5);
}

template <int FAMILY>
__device__ static void heap_sift_down(HeapRec *a, int64_t startpos, int64_t pos) { // heapq._siftdown
    const HeapRec item = a[pos];
    while (pos > startpos) {
        const int64_t parent = (pos - 1) >> 1;
        const HeapRec pr = a[parent];
        if (rec_lt<FAMILY>(item, pr)) { a[pos] = pr; pos = parent; continue; }
        break;
    }
    a[pos] = item;
}

template <int FAMILY>
__device__ static void heap_sift_up(HeapRec *a, int64_t n, int64_t pos) { // heapq._siftup
    const int64_t startpos = pos;
    const HeapRec item = a[pos];
    int64_t child = 2 * pos + 1;
    while (child < n) {
        const int64_t right = child + 1;
        HeapRec cr = a[child];
        if (right < n) {
            const HeapRec rr = a[right];
            if (!rec_lt<FAMILY>(cr, rr)) { child = right; cr = rr; }
        }
        a[pos] = cr;
        pos = child;
        child = 2 * pos + 1;
    }
    a[pos] = item;
    heap_sift_down<FAMILY>(a, startpos, pos);
}

// GraphSearcher.h (graph_search.py:41-44: hypot / manhattan) and GraphSearcher3D.h
// (graph_search_3d.py:46-49); exact: integer d2 -> double, IEEE sqrt.
__device__ static inline double heuristic_exact(int heuristic, int dx, int dy, int dz) {
    dx = abs(dx); dy = abs(dy); dz = abs(dz);
    if (heuristic == PMP_H_MANHATTAN) return (double)(dx + dy + dz);
    return sqrt((double)((long long)dx * dx + (long long)dy * dy + (long long)dz * dz));
}

template <int FAMILY>
__global__ void k_gbfs(const GridDev G, const __grid_constant__ Motions M, int heuristic, u32 nq,
                       const int32_t *starts, const int32_t *goals, u32 *parent, HeapRec *heap_all,
                       int64_t heap_cap, u32 *path, int64_t cap, int32_t *plen, double *cost,
                       int32_t *n_expand) {
    const u32 q = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    const int lane = threadIdx.x & 31;
    if (q >= nq) return;
    const double INF_D = __longlong_as_double(0x7ff0000000000000LL);
    u32 *par = parent + (size_t)q * G.pcells; // 0xFFFFFFFF = not in CLOSED
    HeapRec *heap = heap_all + (size_t)q * heap_cap;
    const int sx = starts[3 * q], sy = starts[3 * q + 1], sz = G.ndim == 3 ? starts[3 * q + 2] : 0;
    const int gx = goals[3 * q], gy = goals[3 * q + 1], gz = G.ndim == 3 ? goals[3 * q + 2] : 0;
    const bool ok = gx >= 0 && gy >= 0 && gz >= 0 && gx < G.X && gy < G.Y && gz < G.Z &&
                    sx >= 0 && sy >= 0 && sz >= 0 && sx < G.X && sy < G.Y && sz < G.Z &&
                    !cell_blocked(G, sx, sy, sz);
    if (!ok) {
        if (lane == 0) { plen[q] = 0; cost[q] = INF_D; n_expand[q] = 0; }
        return;
    }
    const u32 start_cell = (u32)(((size_t)sz * G.PY + sy) * G.PX + sx);
    const u32 goal_cell = (u32)(((size_t)gz * G.PY + gy) * G.PX + gx);
    int64_t hn = 0;   // heap size (lane 0's copy is authoritative, broadcast each iteration)
    u32 counter = 0;
    int closed = 0;
    int result = 0;   // 1 reached, -2 heap overflow
    if (lane == 0) {
        HeapRec r;
        // 2D: Node(start, start, 0, 0) (planner.py:15); 3D: h(start) (gbfs3d.py:29,46)
        r.h = FAMILY == 2 ? 0.0 : heuristic_exact(heuristic, gx - sx, gy - sy, gz - sz);
        r.cell = start_cell;
        r.aux = (counter++ << 5) | 31u;
        heap[0] = r;
        hn = 1;
    }
    for (;;) {
        hn = __shfl_sync(0xFFFFFFFFu, hn, 0);
        if (hn == 0 || result) break;
        HeapRec node;
        node.h = 0.0; node.cell = 0; node.aux = 0;
        int skip = 0;
        if (lane == 0) {
            // heapq.heappop
            const HeapRec last = heap[--hn];
            if (hn > 0) { node = heap[0]; heap[0] = last; heap_sift_up<FAMILY>(heap, hn, 0); }
            else node = last;
            if (par[node.cell] != 0xFFFFFFFFu) skip = 1; // gbfs.py:54-55, gbfs3d.py:55-56
        }
        skip = __shfl_sync(0xFFFFFFFFu, skip, 0);
        if (skip) continue;
        const u32 cell = __shfl_sync(0xFFFFFFFFu, node.cell, 0);
        const u32 aux = __shfl_sync(0xFFFFFFFFu, node.aux, 0);
        const int cx = (int)(cell % G.PX), cy = (int)((cell / G.PX) % G.PY), cz = (int)(cell / ((size_t)G.PX * G.PY));
        if (lane == 0) {
            const int pm = aux & 31;
            const u32 pcell = pm == 31 ? cell
                : (u32)(((size_t)(cz - M.dz[pm]) * G.PY + (cy - M.dy[pm])) * G.PX + (cx - M.dx[pm]));
            par[cell] = pcell; // CLOSED[node.current] = node
            ++closed;
        }
        if (cell == goal_cell) { result = 1; continue; }
        // lanes: legality, closed test and h of each neighbour (getNeighbor, a_star.py:85-96)
        bool legal = false;
        double hv = 0.0;
        u32 ncell = 0;
        if (lane < M.k) {
            const int nx = cx + M.dx[lane], ny = cy + M.dy[lane], nz = cz + M.dz[lane];
            if (edge_legal(G, M, lane, cx, cy, cz)) {
                ncell = (u32)(((size_t)nz * G.PY + ny) * G.PX + nx);
                legal = true;
                hv = heuristic_exact(heuristic, gx - nx, gy - ny, gz - nz);
            }
        }
        __syncwarp();
        if (legal && par[ncell] != 0xFFFFFFFFu) legal = false; // gbfs.py:70-71, gbfs3d.py:74-75
        const u32 mask = __ballot_sync(0xFFFFFFFFu, legal);
        // lane 0 pushes in motion order
        for (int m = 0; m < M.k; ++m) {
            if (!((mask >> m) & 1u)) continue;
            const double h_m = __shfl_sync(0xFFFFFFFFu, hv, m);
            const u32 c_m = __shfl_sync(0xFFFFFFFFu, ncell, m);
            if (lane == 0) {
                if (hn >= heap_cap) { result = -2; }
                else {
                    HeapRec r;
                    r.h = h_m; r.cell = c_m; r.aux = (counter++ << 5) | (u32)m;
                    heap[hn] = r;
                    ++hn;
                    heap_sift_down<FAMILY>(heap, 0, hn - 1); // heapq.heappush
                }
            }
            if (FAMILY == 2 && c_m == goal_cell) break; // gbfs.py:77-79
        }
        result = __shfl_sync(0xFFFFFFFFu, result, 0);
    }
    result = __shfl_sync(0xFFFFFFFFu, result, 0);
    if (lane != 0) return;
    n_expand[q] = closed;
    if (result != 1) {
        plen[q] = result == -2 ? -2 : 0;
        cost[q] = INF_D;
        return;
    }
    // extractPath (a_star.py:98-117): goal -> start, cost += dist
    int64_t n = 0;
    double c = 0.0;
    u32 cell = goal_cell;
    for (;;) {
        if (n < cap) path[(size_t)q * cap + n] = cell;
        ++n;
        if (cell == start_cell) break;
        const u32 p = par[cell];
        const int dx = (int)(cell % G.PX) - (int)(p % G.PX);
        const int dy = (int)((cell / G.PX) % G.PY) - (int)((p / G.PX) % G.PY);
        const int dz = (int)(cell / ((size_t)G.PX * G.PY)) - (int)(p / ((size_t)G.PX * G.PY));
        c += sqrt((double)(dx * dx + dy * dy + dz * dz));
        cell = p;
    }
    plen[q] = n <= cap ? (int32_t)n : -1;
    cost[q] = c;
}

extern "C" int pmp_gbfs(const pmp_grid *g, int heuristic, int64_t nq,
                        const int32_t *starts_dev, const int32_t *goals_dev,
                        uint32_t *parent_dev, void *heap_dev, int64_t heap_cap,
                        uint32_t *path_dev, int64_t path_cap, int32_t *path_len_dev,
                        double *cost_dev, int32_t *n_expand_dev, void *stream) {
    if (g && pmp_grid_upload_fence((pmp_grid *)g, (cudaStream_t)stream)) return -2; // a streamed occupancy upload must have landed
    if (!g || nq < 1 || !starts_dev || !goals_dev || !parent_dev || !heap_dev || heap_cap < 1 ||
        !path_dev || path_cap < 1 || !path_len_dev || !cost_dev || !n_expand_dev) {
        pmp_set_error("pmp_gbfs: bad arguments");
        return -1;
    }
    cudaStream_t s = (cudaStream_t)stream;
    PMP_CHECK_CUDA(cudaMemsetAsync(parent_dev, 0xFF, (size_t)nq * g->d.pcells * 4, s));
    const unsigned blocks = (unsigned)((nq * 32 + 63) / 64);
    if (g->d.ndim == 2)
        k_gbfs<2><<<blocks, 64, 0, s>>>(g->d, g->mot, heuristic, (u32)nq, starts_dev, goals_dev, parent_dev, (HeapRec *)heap_dev, heap_cap, path_dev, path_cap, path_len_dev, cost_dev, n_expand_dev);
    else
        k_gbfs<3><<<blocks, 64, 0, s>>>(g->d, g->mot, heuristic, (u32)nq, starts_dev, goals_dev, parent_dev, (HeapRec *)heap_dev, heap_cap, path_dev, path_cap, path_len_dev, cost_dev, n_expand_dev);
    PMP_CHECK_CUDA(cudaGetLastError());
    return 0;
}
