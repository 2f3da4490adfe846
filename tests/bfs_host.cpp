// bfs_host.cpp -- CPU harness of the unit-weight field kernel (test infrastructure only).
//
// Compiles python_motion_planning_b200/csrc/pmp_bfs_core.cuh UNCHANGED against tests/simt_host.h
// (fibers as lanes, random warp interleaving) and checks the field it produces against a plain
// queue BFS -- the same six unit moves the reference's Dijkstra3D walks when env.motions is
// filtered to cost-1 moves (dijkstra3d.py:39-126, env3d.py:56-70).  Several z-slabs ("ranks")
// share one address space here, so the peer-memory protocol of the multi-GPU path runs too.
//
//   bfs_host X Y Z seed density_percent ranks warps_per_rank [mode]
//     mode: 0 random cells, 1 random boxes, 2 serpentine walls (long in-tile paths), 3 empty
#include <algorithm>
#include <queue>
#include <vector>

#include "simt_host.h"

#include "../python_motion_planning_b200/csrc/pmp_griddev.h"
#include "../python_motion_planning_b200/csrc/pmp_bfs_core.cuh"

struct Rank {
    BfsArgs A;
    std::vector<u32> bits, g, state, ctr;
    std::vector<u64> ring;
    std::vector<u64> stats;
    int z0, z1;
};

static std::vector<Rank> R;
static int g_warps = 4;

// streamed occupancy upload (GridDev.ready): an extra "copy engine" fiber sets the chunk flags one by one
// while the search runs; the occupancy words of a chunk are garbage until its flag is set
static std::vector<u32> g_ready, g_bits_final;
static int g_stream_shift = -1;

static void entry(int warp, int lane, void *) {
    if (warp == (int)R.size() * g_warps) { // the copy-engine fiber (lane 0 only; the other lanes leave at once)
        if (lane != 0) return;
        Rank &K = R[0];
        const int cp = 1 << g_stream_shift, PZ = K.A.G.PZ;
        const size_t plane = (size_t)K.A.G.wpr * K.A.G.PY;
        for (int c = 0; c * cp < PZ; ++c) {
            for (int k = 0; k < 40; ++k) simt::yield();                       // the copy takes a while
            const size_t a = (size_t)c * cp * plane, b = std::min((size_t)(c + 1) * cp, (size_t)PZ) * plane;
            for (size_t i = a; i < b; ++i) K.bits[i] = g_bits_final[i];          // the chunk lands ...
            g_ready[c] = 1;                                                      // ... then its flag
        }
        return;
    }
    const int r = warp / g_warps;
    if (R.size() > 1) bfs_warp_main<true>(R[r].A, lane, warp % g_warps == 0);
    else bfs_warp_main<false>(R[r].A, lane, warp % g_warps == 0);
}

static uint64_t rng_state;
static uint32_t rnd32() {
    rng_state ^= rng_state << 13; rng_state ^= rng_state >> 7; rng_state ^= rng_state << 17;
    return (uint32_t)(rng_state >> 11);
}

int main(int argc, char **argv) {
    if (argc < 8) { fprintf(stderr, "usage: bfs_host X Y Z seed density ranks warps [mode]\n"); return 2; }
    const int X = atoi(argv[1]), Y = atoi(argv[2]), Z = atoi(argv[3]);
    const uint64_t seed = strtoull(argv[4], nullptr, 10);
    const int dens = atoi(argv[5]), nr = atoi(argv[6]);
    g_warps = atoi(argv[7]);
    const int mode = argc > 8 ? atoi(argv[8]) : 0;
    g_stream_shift = argc > 9 ? atoi(argv[9]) : -1; // >= 3: planes per chunk = 2^shift, single rank only
    rng_state = seed * 0x9E3779B97F4A7C15ull + 99;
    for (int i = 0; i < 8; ++i) rnd32();

    // ---- occupancy ----
    std::vector<uint8_t> occ((size_t)X * Y * Z, 0);
    auto at = [&](int x, int y, int z) -> uint8_t & { return occ[((size_t)z * Y + y) * X + x]; };
    if (mode == 0) {
        for (auto &c : occ) c = (int)(rnd32() % 100) < dens;
    } else if (mode == 1) {
        const int nb = (int)((size_t)X * Y * Z * dens / 100 / 40) + 1;
        for (int b = 0; b < nb; ++b) {
            const int cx = rnd32() % X, cy = rnd32() % Y, cz = rnd32() % Z;
            const int ex = 1 + rnd32() % 6, ey = 1 + rnd32() % 6, ez = 1 + rnd32() % 6;
            for (int z = cz; z < std::min(Z, cz + ez); ++z)
                for (int y = cy; y < std::min(Y, cy + ey); ++y)
                    for (int x = cx; x < std::min(X, cx + ex); ++x) at(x, y, z) = 1;
        }
    } else if (mode == 2) {
        // walls at every other y with a gap alternating between the x ends, and at every other z
        // with a gap alternating between the y ends: one long corridor
        for (int z = 0; z < Z; ++z)
            for (int y = 0; y < Y; ++y)
                for (int x = 0; x < X; ++x) {
                    if (z % 2 == 1) { at(x, y, z) = !(((z / 2) % 2 == 0) ? (y == Y - 1 && x == X - 1) : (y == 0 && x == 0)); continue; }
                    if (y % 2 == 1) at(x, y, z) = !(((y / 2) % 2 == 0) ? x == X - 1 : x == 0);
                }
    }
    int sx = 0, sy = 0, sz = 0;
    if (mode != 2) { sx = rnd32() % X; sy = rnd32() % Y; sz = rnd32() % Z; }
    at(sx, sy, sz) = 0;

    // ---- reference: plain BFS ----
    std::vector<uint32_t> ref((size_t)X * Y * Z, 0xFFFFFFFFu);
    {
        std::queue<size_t> q;
        const size_t s = ((size_t)sz * Y + sy) * X + sx;
        ref[s] = 0;
        q.push(s);
        const int dx[6] = {1, -1, 0, 0, 0, 0}, dy[6] = {0, 0, 1, -1, 0, 0}, dz[6] = {0, 0, 0, 0, 1, -1};
        while (!q.empty()) {
            const size_t c = q.front(); q.pop();
            const int x = (int)(c % X), y = (int)((c / X) % Y), z = (int)(c / ((size_t)X * Y));
            for (int d = 0; d < 6; ++d) {
                const int nx = x + dx[d], ny = y + dy[d], nz = z + dz[d];
                if (nx < 0 || ny < 0 || nz < 0 || nx >= X || ny >= Y || nz >= Z) continue;
                const size_t n = ((size_t)nz * Y + ny) * X + nx;
                if (occ[n] || ref[n] != 0xFFFFFFFFu) continue;
                ref[n] = ref[c] + 1;
                q.push(n);
            }
        }
    }

    // ---- slabs ----
    const int PX = (X + 15) / 16 * 16, PY = (Y + 7) / 8 * 8;
    const int ztiles = (Z + 7) / 8;
    R.resize(nr);
    for (int r = 0; r < nr; ++r) {
        const int t0 = (int)((long long)ztiles * r / nr), t1 = (int)((long long)ztiles * (r + 1) / nr);
        R[r].z0 = t0 * 8;
        R[r].z1 = std::min(Z, t1 * 8);
        if (t1 <= t0) { fprintf(stderr, "too many ranks for %d z tiles\n", ztiles); return 2; }
    }
    for (int r = 0; r < nr; ++r) {
        Rank &K = R[r];
        const int Zl = K.z1 - K.z0, PZ = (Zl + 7) / 8 * 8;
        GridDev G;
        memset(&G, 0, sizeof G);
        G.ndim = 3; G.X = X; G.Y = Y; G.Z = Zl; G.PX = PX; G.PY = PY; G.PZ = PZ;
        G.wpr = (PX + 31) / 32;
        G.pcells = (u64)PX * PY * PZ;
        K.bits.assign((size_t)G.wpr * PY * PZ, 0xFFFFFFFFu);
        for (int z = 0; z < Zl; ++z)
            for (int y = 0; y < Y; ++y)
                for (int x = 0; x < X; ++x)
                    if (!at(x, y, K.z0 + z)) K.bits[((size_t)z * PY + y) * G.wpr + (x >> 5)] &= ~(1u << (x & 31));
        G.bits = K.bits.data();
        BfsArgs &A = K.A;
        memset(&A, 0, sizeof A);
        A.G = G;
        A.nbx = G.wpr; A.nby = PY / 8; A.nbz = PZ / 8;
        A.ntiles = (u32)(A.nbx * A.nby * A.nbz);
        K.g.assign(G.pcells, 0xDEADBEEFu); // garbage: the kernel must not rely on a pre-filled field
        K.state.assign(A.ntiles, 0);
        A.ring_cap = A.ntiles + 64 + (u32)(g_warps);
        K.ring.assign(A.ring_cap, 0);
        K.ctr.assign(BFS_C_WORDS, 0);
        K.stats.assign(32, 0);
        A.g = K.g.data(); A.state = K.state.data(); A.ring = K.ring.data(); A.ctr = K.ctr.data();
        A.stats = K.stats.data();
        A.rank = r; A.world = nr;
        A.timeout_ns = 0;
        A.max_pops = 64u * A.ntiles; // the product's activation budget (tiles put back during a streamed upload do not count)
        A.pace_window = nr > 1 ? 8 : 0; // tight: entries are held often
        A.pace_hold_polls = 40;
    }
    for (int r = 0; r < nr; ++r) {
        BfsArgs &A = R[r].A;
        for (int p = 0; p < nr; ++p) A.peer_ctr[p] = R[p].ctr.data();
        A.active = &R[0].ctr[BFS_C_ACTIVE];
        if (r > 0) {
            Rank &P = R[r - 1];
            A.lo.g = P.g.data(); A.lo.state = P.state.data(); A.lo.ring = P.ring.data(); A.lo.ctr = P.ctr.data();
            A.lo.ring_cap = P.A.ring_cap; A.lo.nbz = P.A.nbz; A.lo.zadj = (P.z1 - P.z0) - 1; A.lo.bzadj = P.A.nbz - 1;
        }
        if (r < nr - 1) {
            Rank &P = R[r + 1];
            A.hi.g = P.g.data(); A.hi.state = P.state.data(); A.hi.ring = P.ring.data(); A.hi.ctr = P.ctr.data();
            A.hi.ring_cap = P.A.ring_cap; A.hi.nbz = P.A.nbz; A.hi.zadj = 0; A.hi.bzadj = 0;
        }
    }
    // ---- init (what k_bfs_init does on the device) ----
    R[0].ctr[BFS_C_ACTIVE] = 1;
    for (int r = 0; r < nr; ++r) {
        Rank &K = R[r];
        if (sz < K.z0 || sz >= K.z1) continue;
        const int lz = sz - K.z0;
        const u32 t = (u32)(((lz / 8) * K.A.nby + sy / 8) * K.A.nbx + sx / 32);
        K.state[t] = BFS_SRC | BFS_Q;
        K.ring[0] = (u64)(t + 1);
        K.ctr[BFS_C_TAIL] = 1;
        K.ctr[BFS_C_INFLIGHT] = 1;
        K.ctr[BFS_C_SRC] = (u32)sx; K.ctr[BFS_C_SRC + 1] = (u32)sy; K.ctr[BFS_C_SRC + 2] = (u32)lz;
    }

    if (g_stream_shift >= 3 && nr == 1) {
        Rank &K = R[0];
        g_bits_final = K.bits;
        for (auto &w : K.bits) w = 0xA5A5A5A5u ^ (u32)(&w - K.bits.data()); // garbage until the chunk "arrives"
        g_ready.assign((K.A.G.PZ >> g_stream_shift) + 2, 0);
        K.A.G.ready = g_ready.data();
        K.A.G.ready_shift = g_stream_shift;
        // the source's chunk must be there before the init step reads it (k_bfs_init waits for it on the device)
    } else g_stream_shift = -1;
    simt::run(nr * g_warps + (g_stream_shift >= 3 ? 1 : 0), entry, nullptr, seed, 3);

    // ---- fill untouched tiles, compare ----
    size_t bad = 0, reached = 0;
    u64 act = 0, idle = 0, lowered = 0, levels = 0, passes = 0, redo = 0, remote = 0, tiles = 0, touched_tiles = 0;
    for (int r = 0; r < nr; ++r) {
        Rank &K = R[r];
        if (K.ctr[BFS_C_ERR]) {
            printf("FAIL rank %d error flag %u (tickets drawn %u, entries pushed %u, put back %u, activations %llu, budget %u)\n", r, K.ctr[BFS_C_ERR],
                   K.ctr[BFS_C_HEAD], K.ctr[BFS_C_TAIL], K.ctr[BFS_C_PUTBACK], (unsigned long long)K.stats[1], K.A.max_pops);
            return 1;
        }
        if (K.ctr[BFS_C_INFLIGHT]) { printf("FAIL rank %d inflight %u at exit\n", r, K.ctr[BFS_C_INFLIGHT]); return 1; }
        for (u32 t = 0; t < K.A.ntiles; ++t) {
            ++tiles;
            if (K.state[t] & ~BFS_T) { printf("FAIL rank %d tile %u state %x at exit\n", r, t, K.state[t]); return 1; }
            if (K.state[t] & BFS_T) { ++touched_tiles; continue; }
            for (int lane = 0; lane < 32; ++lane) bfs_fill_tile(K.A, t, lane);
        }
        const int Zl = K.z1 - K.z0;
        for (int z = 0; z < Zl; ++z)
            for (int y = 0; y < Y; ++y)
                for (int x = 0; x < X; ++x) {
                    const u32 raw = K.g[((size_t)z * PY + y) * PX + x];
                    const u32 want = ref[((size_t)(K.z0 + z) * Y + y) * X + x];
                    const u32 want_raw = want == 0xFFFFFFFFu ? 0xFFFFFFFFu : (want << 1) | 1u;
                    if (want != 0xFFFFFFFFu) ++reached;
                    if (raw != want_raw) {
                        if (bad < 10) printf("MISMATCH rank %d cell (%d,%d,%d): got raw %08x want %08x (value %u)\n", r, x, y, K.z0 + z, raw, want_raw, want);
                        ++bad;
                    }
                }
        // padding cells must read as unreached
        for (size_t i = 0; i < K.g.size(); ++i) {
            const int x = (int)(i % PX), y = (int)((i / PX) % PY), z = (int)(i / ((size_t)PX * PY));
            if ((x >= X || y >= Y || z >= Zl) && K.g[i] != 0xFFFFFFFFu) {
                if (bad < 10) printf("PADDING rank %d (%d,%d,%d) = %08x\n", r, x, y, z, K.g[i]);
                ++bad;
            }
        }
        act += K.stats[1]; lowered += K.stats[2]; levels += K.stats[12]; idle += K.stats[13]; passes += K.stats[14];
        redo += K.stats[5]; remote += K.stats[7];
    }
    printf("%s cells=%zu reached=%zu bad=%zu tiles=%llu touched=%llu activations=%llu idle=%llu lowered=%llu levels=%llu passes=%llu requeues=%llu remote=%llu\n",
           bad ? "FAIL" : "OK", occ.size(), reached, bad, tiles, touched_tiles, act, idle, lowered, levels, passes, redo, remote);
    return bad ? 1 : 0;
}
