// pmp_common.cuh -- shared device/host definitions of libpmp_b200 (sm_100a).
#pragma once

#include <cuda_runtime.h>
#include <stdint.h>

#include "../../include/pmp_b200.h"
#include "pmp_griddev.h"

#define PMP_MAXM 26


#define KEY_INF (~0ull)

// Motion table (env.motions) in device form.  `need[m]` is the set of cells of the 3x3x3
// neighbourhood of the SOURCE cell that must be free for the edge u -> u + m to exist
// (bit (dz+1)*9 + (dy+1)*3 + (dx+1)): the target, plus for 2-axis moves the two in-plane face
// neighbours, plus for 3-axis moves the three face neighbours of u -- exactly the cells tested by
// GraphSearcher.isCollision (graph_search.py:61-87) and GraphSearcher3D.isCollision
// (graph_search_3d.py:66-107).
struct Motions {
    int k;
    int cut; // 1: diagonal moves may cut corners (only the target must be free) -- the graph of the reference's JPS (jps.py:85-121)
    int8_t dx[PMP_MAXM], dy[PMP_MAXM], dz[PMP_MAXM];
    u32 need[PMP_MAXM];
    u32 wi[PMP_MAXM];     // integer cost (u32 fields)
    double w[PMP_MAXM];   // cost accumulated along a path
    double len[PMP_MAXM]; // Euclidean length of the offset (Planner.dist, planner.py:22-23)
};

struct pmp_grid {
    GridDev d;
    u32 *bits; // owned
    Motions mot;
    int device;
    double wmin;     // smallest motion cost
    double he_scale; // largest a with a*euclid admissible for this motion table
    double hm_scale; // largest a with a*manhattan admissible
    bool integer_costs;
    int cut; // corner cutting allowed (pmp_grid_set_corner_cutting)
    // streamed occupancy upload (pmp_grid_set_packed_host_streamed): flags + the event that marks its end
    u32 *ready_dev;      // owned; n_ready words
    u32 *ones_host;      // owned, pinned: the source of the flag writes
    int n_ready;
    cudaEvent_t ev_reset, ev_uploaded; // owned (created on first use)
    bool upload_pending; // kernels that cannot wait for chunks must wait for ev_uploaded first (pmp_grid_upload_fence)
};

// make `stream` wait for a streamed occupancy upload still in flight; every entry point that reads the
// bitmask and cannot wait chunk by chunk calls this first
int pmp_grid_upload_fence(pmp_grid *g, cudaStream_t stream);

// ---- weight traits: how a cost field element is stored, compared and advanced ----
struct WF64 {
    typedef u64 G; // bit pattern of a non-negative double: orders like the double
    static constexpr int kind = PMP_W_F64;
    // Stored word = (value bits << 1) | clean; clean = 0 marks a cell that still has to
    // propagate ("pending").  One atomicMin therefore publishes a lower value AND its pending
    // flag; the sign bit of a non-negative double is free, so the shift is lossless.
    __host__ __device__ static inline G ginf() { return ~0ull; }      // stored "unreached"
    __host__ __device__ static inline G inf() { return ~0ull >> 1; }  // decoded "unreached"
    __host__ __device__ static inline G enc(G v, bool pending) { return (v << 1) | (pending ? 0ull : 1ull); }
    __host__ __device__ static inline G dec(G s) { return s >> 1; }
    __device__ static inline G add(G g, const Motions &M, int m) {
        return (G)__double_as_longlong(__longlong_as_double((long long)g) + M.w[m]);
    }
    __device__ static inline u64 wbits(const Motions &M, int m) { return (u64)__double_as_longlong(M.w[m]); }
    __device__ static inline G addw(G g, u64 w) {
        return (G)__double_as_longlong(__longlong_as_double((long long)g) + __longlong_as_double((long long)w));
    }
    // ordering key of a cell: bits of (g + h)
    __device__ static inline u64 key(G g, double h) {
        return (u64)__double_as_longlong(__longlong_as_double((long long)g) + h);
    }
    __device__ static inline u64 key_add(u64 k, double d) {
        return (u64)__double_as_longlong(__longlong_as_double((long long)k) + d);
    }
    __device__ static inline double to_double(G g) {
        return g == inf() ? __longlong_as_double(0x7ff0000000000000LL)
                          : __longlong_as_double((long long)g);
    }
};

struct WU32 {
    typedef u32 G;
    static constexpr int kind = PMP_W_U32;
    __host__ __device__ static inline G ginf() { return 0xFFFFFFFFu; }
    __host__ __device__ static inline G inf() { return 0x7FFFFFFFu; }
    __host__ __device__ static inline G enc(G v, bool pending) { return (v << 1) | (pending ? 0u : 1u); }
    __host__ __device__ static inline G dec(G s) { return s >> 1; }
    __device__ static inline G add(G g, const Motions &M, int m) { return g + M.wi[m]; }
    __device__ static inline u64 wbits(const Motions &M, int m) { return (u64)M.wi[m]; }
    __device__ static inline G addw(G g, u64 w) { return g + (u32)w; }
    __device__ static inline u64 key(G g, double h) { return (u64)g + (u64)h; }
    __device__ static inline u64 key_add(u64 k, double d) { return k + (u64)d; }
    __device__ static inline double to_double(G g) {
        return g == inf() ? __longlong_as_double(0x7ff0000000000000LL) : (double)g;
    }
};

// L2-coherent loads (the persistent kernels must not read stale L1 lines)
__device__ static inline u64 ld_cg(const u64 *p) { return __ldcg(p); }
__device__ static inline u32 ld_cg(const u32 *p) { return __ldcg(p); }
__device__ static inline int ld_cg(const int *p) { return __ldcg(p); }

__device__ static inline void red_min(u64 *p, u64 v) { atomicMin(p, v); }
__device__ static inline void red_min(u32 *p, u32 v) { atomicMin(p, v); }
__device__ static inline void cas_word(u64 *p, u64 cmp, u64 v) { atomicCAS(p, cmp, v); }
__device__ static inline void cas_word(u32 *p, u32 cmp, u32 v) { atomicCAS(p, cmp, v); }
// release/acquire fence at device scope (cheaper than the sequentially consistent __threadfence)
__device__ static inline void fence_gpu() { asm volatile("fence.acq_rel.gpu;" ::: "memory"); }

// n (<= 34) occupancy bits of bitmask row (y, z) for x in [x0, x0 + n); bit i <-> x0 + i.
// Anything outside the padded grid reads as blocked.
__device__ static inline u64 row_bits(const GridDev &G, int x0, int y, int z, int n) {
    const u64 nmask = (n >= 64) ? ~0ull : ((1ull << n) - 1ull);
    if (y < 0 || y >= G.PY || z < 0 || z >= G.PZ) return nmask;
    const u32 *row = G.bits + ((size_t)z * G.PY + y) * G.wpr;
    const int wi = x0 >> 5; // arithmetic shift: x0 = -1 -> -1
    const int sh = x0 & 31;
    const u32 a = (wi >= 0 && wi < G.wpr) ? __ldg(row + wi) : 0xFFFFFFFFu;
    const u32 b = (wi + 1 >= 0 && wi + 1 < G.wpr) ? __ldg(row + wi + 1) : 0xFFFFFFFFu;
    const u32 c = (wi + 2 >= 0 && wi + 2 < G.wpr) ? __ldg(row + wi + 2) : 0xFFFFFFFFu;
    u64 lo = (u64)a | ((u64)b << 32);
    u64 out = lo >> sh;
    if (sh) out |= (u64)c << (64 - sh);
    return out & nmask;
}

__device__ static inline int cell_blocked(const GridDev &G, int x, int y, int z) {
    if (x < 0 || y < 0 || z < 0 || x >= G.PX || y >= G.PY || z >= G.PZ) return 1;
    const u32 w = __ldg(G.bits + ((size_t)z * G.PY + y) * G.wpr + (x >> 5));
    return (w >> (x & 31)) & 1u;
}

// legality of the edge u -> u + m, evaluated at u (global bitmask reads)
__device__ static inline bool edge_legal(const GridDev &G, const Motions &M, int m, int ux, int uy, int uz) {
    const int dx = M.dx[m], dy = M.dy[m], dz = M.dz[m];
    if (cell_blocked(G, ux, uy, uz) || cell_blocked(G, ux + dx, uy + dy, uz + dz)) return false;
    const int nz = (dx != 0) + (dy != 0) + (dz != 0);
    if (nz >= 2 && !M.cut) {
        if (dx && cell_blocked(G, ux + dx, uy, uz)) return false;
        if (dy && cell_blocked(G, ux, uy + dy, uz)) return false;
        if (dz && cell_blocked(G, ux, uy, uz + dz)) return false;
    }
    return true;
}

// Admissible heuristic used for ordering / pruning only: max(he * euclid_lo, hm * manhattan),
// euclid_lo a round-down float sqrt (never above the reference's math.hypot / math.sqrt,
// graph_search.py:41-44, graph_search_3d.py:46-49).
__device__ static inline double heuristic_lo(int dx, int dy, int dz, double he, double hm) {
    dx = abs(dx); dy = abs(dy); dz = abs(dz);
    const int d2 = dx * dx + dy * dy + dz * dz;
    const float e = __fsqrt_rd(__int2float_rd(d2));
    const double a = he * (double)e;
    const double b = hm * (double)(dx + dy + dz);
    return a > b ? a : b;
}

void pmp_set_error(const char *fmt, ...);
#define PMP_CHECK_CUDA(expr)                                                                     \
    do {                                                                                         \
        cudaError_t _e = (expr);                                                                 \
        if (_e != cudaSuccess) {                                                                 \
            pmp_set_error("%s failed: %s (%s:%d)", #expr, cudaGetErrorString(_e), __FILE__,      \
                          __LINE__);                                                             \
            return -2;                                                                           \
        }                                                                                        \
    } while (0)
