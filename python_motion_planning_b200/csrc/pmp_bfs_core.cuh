// pmp_bfs_core.cuh -- unit-weight cost-to-go fields on 6-connected voxel grids: one WARP per tile,
// the whole tile solve in registers (no shared memory, no CTA barrier).
//
// Replaces, for the integer-weight 6-connected case (BASELINE.json config 5), the expansion loop of
// Dijkstra3D.plan / getNeighbor (global_planner/graph_search/dijkstra3d.py:39-126) run to
// exhaustion with `env.motions` filtered to the six unit moves: with unit weights the cost of a
// cell is its BFS level, so the same field is produced bit for bit.
//
// Scheme
//   * tile = 32(x) x 8(y) x 8(z) cells.  Lane l of the warp owns the column x = x0 + l: 64 cells
//     as two 32-bit words, bit (z & 3) * 8 + y of word z >> 2.  A BFS level step is then a handful
//     of shifts inside the lane (y, z neighbours) plus four warp shuffles (x neighbours); the
//     level of every cell is recorded in bit-sliced form (7 planes x 2 words) and turned back into
//     integers by 8x8 bit-matrix transposes -- row stores of the field are naturally coalesced
//     (one 128-byte line per (y, z) row per instruction).
//   * PULL model: field words are final values in the library's (value << 1 | 1) format, written
//     only by the warp that owns the tile.  A tile that lowered boundary cells sets a dirty-face
//     bit in the neighbour tile's state word and queues the tile; the neighbour reads the face
//     straight out of the owner's field rows.  Because min-plus relaxation is linear, an
//     activation only sweeps from the NEW seeds (halo + 1 that beat the own boundary value) and
//     min-merges the result -- a tile that was never touched is not even read (lazy
//     initialisation: no field fill before the search; untouched tiles are filled afterwards).
//   * scheduling: persistent warps pop tile ids from one FIFO ring (near-BFS order, so a tile is
//     activated about once); per tile a state word {dirty faces, queued, touched} gives mutual
//     exclusion; an in-flight counter ends the search.
//   * multi-GPU (z-slabs): the tiles of the lowest / highest z-row read their z halo from the PEER
//     GPU's field over NVLink and arm / queue the peer's tile with remote atomics -- the same
//     protocol as between two tiles of one GPU; a counter of active ranks detects termination.
//
// The file is compiled by nvcc for sm_100a (pmp_bfs.cu) and, unchanged, by g++ against
// tests/simt_host.h (a fiber-based SIMT emulator) for the CPU tests.
#pragma once

#define BFS_FULL 0xFFFFFFFFu
#define BFS_SRC 0x40u    // state: the source cell lies in this tile and has not been injected yet
#define BFS_Q 0x80u      // state: the tile is in the ring or being processed
#define BFS_T 0x100u     // state: the tile's field words are initialised
#define BFS_DIRTY 0x3Fu  // state bits 0..5: face f has new values on the neighbour's side
#define BFS_INF 0x7FFFFFFFu
#define BFS_RAWINF 0xFFFFFFFFu
#define BFS_CINV 0xFFFFFFFFu
#define BFS_NREC 7
#define BFS_LIM 128      // levels one pass can record (2^BFS_NREC)
#ifndef BFS_LB
#define BFS_LB 3      // the level loop is unrolled 2^BFS_LB times (code size vs. static low level bits; 2 -> 3 after the per-block overheads were trimmed: 17.2 -> 16.8 ms)
#endif
#define BFS_UB (1 << BFS_LB)
#define BFS_SEEDMAX 63u  // largest seed level relative to the pass base (6 seed planes)

// control block (u32 words; hot counters on separate 128-byte lines)
#define BFS_C_HEAD 0
#define BFS_C_TAIL 32
#define BFS_C_INFLIGHT 64
#define BFS_C_DONE 96
#define BFS_C_ERR 97
#define BFS_C_PUTBACK 98 // streamed upload: tiles put back into the ring because their occupancy had not arrived (not activations)
#define BFS_C_ACTIVE 128 // ranks with work in flight (only rank 0's word is used)
#define BFS_C_SRC 160    // local source cell x, y, z and a valid flag
// multi-GPU pacing (bfs_seq_*): entries pushed / started on this rank, and the other ranks' counts as they
// last reported them
#define BFS_C_PUSHED 192
#define BFS_C_STARTED 224
#define BFS_C_GP 256     // + r: entries pushed by rank r
#define BFS_C_GS 288     // + r: entries started by rank r
#define BFS_C_WORDS 320

// hook: the kernel that keeps the warps of a CTA in step defines this as __syncthreads()
#ifndef BFS_PHASE_SYNC
#define BFS_PHASE_SYNC()
#endif
#ifndef BFS_CLAIM_FENCE
#define BFS_CLAIM_FENCE 1 // experiment: 0 = no fence between the claim and the halo loads (single GPU)
#endif
#ifndef BFS_PROF
#define BFS_PROF 0 // experiment builds: 1 / 2 = the phase counters [6], [11], [3], [0] hold four sub-phases of publish / setup instead
#endif
#ifndef BFS_CLOCKS
#define BFS_CLOCKS (BFS_PROF != 0) // 1: phase times per activation in stats [4], [6], [11], [3], [0] (experiment builds: ten registers and ~50 instructions per activation)
#endif
#define BFS_NOW() (BFS_CLOCKS ? clock64() : 0ll)
#define BFS_MARK(which, field) do { if (BFS_PROF == (which)) { const long long now_ = clock64(); cnt.field += (u64)(now_ - tmark); tmark = now_; } } while (0)
#ifndef BFS_PUTBACK_NS
#define BFS_PUTBACK_NS 400 // pause before a tile whose occupancy has not arrived goes back into the ring
#endif
#ifndef BFS_SLEEP_CAP
#define BFS_SLEEP_CAP 128 // longest back-off (ns) of a warp polling an empty ring slot (1024 -> 128: -3..6 % at 1536^3)
#endif
#ifndef BFS_SYNC_MASK
#define BFS_SYNC_MASK 7 // which phase boundaries carry a barrier: 1 setup|sweep, 2 sweep|extract, 4 extract|publish
#endif

#define BFS_ERR_TIMEOUT 1u
#define BFS_ERR_BUDGET 2u

struct BfsPeer {
    u32 *g;      // peer field (nullptr: no neighbour slab on this side)
    u32 *state;
    u64 *ring;
    u32 *ctr;
    u32 ring_cap;
    int nbz;     // z tile rows of the peer
    int zadj;    // the peer's plane adjacent to this slab (its local z)
    int bzadj;   // the peer's tile row adjacent to this slab
};

struct BfsArgs {
    GridDev G;   // the LOCAL box (a whole grid, or one z-slab of it)
    int nbx, nby, nbz;
    u32 ntiles;
    u32 *g;
    u32 *state;
    u64 *ring;   // entries: (global sequence number << 32) | (tile id + 1); 0 = empty slot
    u32 ring_cap;
    u32 *ctr;
    u64 *stats;
    int rank, world;
    BfsPeer lo, hi;      // slabs below (z - 1) / above (z + 1)
    u32 *peer_ctr[8];    // control blocks of all ranks (termination broadcast); [rank] = own
    u32 *active;         // rank 0's BFS_C_ACTIVE word
    u64 timeout_ns;      // spin-loop safety net (0: none)
    u32 max_pops;        // activation budget of this rank (0: none): a search that needs more is aborted
    u32 pace_window;     // multi-GPU: an entry waits while its sequence number is more than this ahead of the number of
                         // entries started on all ranks (bfs_seq_*; 0: no pacing -- single-GPU runs never pace)
    u32 pace_hold_polls; // polls (~0.25 us each) an entry is held at most before it is served regardless
};

struct BfsCount {
    u32 act, idle, lowered, levels, passes, redo, remote;
    u32 all_up;    // streamed occupancy upload: the last chunk has been seen, no more flag checks
    u32 peer_rows; // 128-byte rows of a neighbour GPU's field read over NVLink (halo pulls + arming checks)
    u64 c_pop, c_setup, c_sweep, c_extract, c_publish; // cycles (lane 0's clock) per phase, summed over activations
};

#ifdef __CUDACC__
__device__ static inline u64 bfs_now_ns() {
    u64 t;
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
    return t;
}
__device__ static inline void bfs_store(u32 *p, u32 v) { *p = v; }
__device__ static inline void bfs_store_volatile(u32 *p, u32 v) { *(volatile u32 *)p = v; }
// release / acquire fence at device scope (the sequentially consistent __threadfence costs more)
__device__ __forceinline__ void bfs_fence() { asm volatile("fence.acq_rel.gpu;" ::: "memory"); }
// a field word of the neighbour GPU (NVLink): system-scope load, never served from this SM's L1
__device__ __forceinline__ u32 bfs_ld_peer(const u32 *p) {
    u32 v;
    asm volatile("ld.relaxed.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
    return v;
}
__device__ __forceinline__ u32 bfs_ld_acquire_sys(const u32 *p) {
    u32 v;
    asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
    return v;
}
// all-ones if bit 7 of byte k of v is set, else 0: PRMT in its sign-replicating form (selector
// nibble 8 | k) -- __byte_perm() masks the selector to 3 bits, so it takes inline PTX
__device__ __forceinline__ u32 bfs_byte_sign(u32 v, int k) {
    u32 d;
    asm("prmt.b32 %0, %1, %2, %3;" : "=r"(d) : "r"(v), "r"(0u), "r"(0x1111u * (u32)(8 | k)));
    return d;
}
#else
static inline u64 bfs_now_ns() { return 0; }
static inline void bfs_store(u32 *p, u32 v) { *p = v; }
static inline void bfs_store_volatile(u32 *p, u32 v) { *(volatile u32 *)p = v; }
static inline void bfs_fence() { __threadfence(); }
static inline u32 bfs_ld_peer(const u32 *p) { return *(const volatile u32 *)p; }
static inline u32 bfs_ld_acquire_sys(const u32 *p) { return *(const volatile u32 *)p; }
static inline u32 bfs_byte_sign(u32 v, int k) { return ((v >> (8 * k)) & 0x80u) ? 0xFFFFFFFFu : 0u; }
#endif

// the 64-bit value hi:lo shifted right / left by s (0 < s < 32): its low / high word
__device__ __forceinline__ u32 bfs_funnel_r(u32 lo, u32 hi, int s) { return (lo >> s) | (hi << (32 - s)); }
__device__ __forceinline__ u32 bfs_funnel_l(u32 lo, u32 hi, int s) { return (hi << s) | (lo >> (32 - s)); }
__device__ __forceinline__ u32 bfs_spread4(u32 n) { return ((n & 0xFu) * 0x00204081u) & 0x01010101u; }
__device__ __forceinline__ u32 bfs_compress4(u32 x) { return (((x & 0x01010101u) * 0x01020408u) >> 24) & 0xFu; }

// 32 x 32 bit-matrix transpose across the warp: afterwards lane x holds the word whose bit l is
// bit x of the word lane l passed in.  Three matrices at once (the tile's two occupancy words and
// the rows behind its faces).  Butterfly step j: a lane with bit j set keeps its high j-blocks and
// takes the partner's high blocks moved down, the others the mirror image; both are "rotate the
// partner's word, then select by mask" -- the rotation count and the mask depend on the lane and the
// step only, so the three words share them: shuffle + funnel shift + one LOP3 per word and step.
__device__ __forceinline__ void bfs_transpose32x3(u32 &a, u32 &b, u32 &c, const int lane) {
#pragma unroll
    for (int j = 16; j >= 1; j >>= 1) {
        const u32 m = j == 16 ? 0x0000FFFFu : (j == 8 ? 0x00FF00FFu : (j == 4 ? 0x0F0F0F0Fu : (j == 2 ? 0x33333333u : 0x55555555u)));
        const bool up = (lane & j) != 0;
        const u32 keep = up ? ~m : m;          // bits of the own word that stay
        const int rs = up ? j : 32 - j;        // rotate the partner's word right by this much
        const u32 oa = __shfl_xor_sync(BFS_FULL, a, j), ob = __shfl_xor_sync(BFS_FULL, b, j), oc = __shfl_xor_sync(BFS_FULL, c, j);
        a = (a & keep) | (((oa >> rs) | (oa << (32 - rs))) & ~keep);
        b = (b & keep) | (((ob >> rs) | (ob << (32 - rs))) & ~keep);
        c = (c & keep) | (((oc >> rs) | (oc << (32 - rs))) & ~keep);
    }
}

// 8 x 8 bit-matrix transpose of the 64-bit value hi:lo (row k = byte k).  The first two butterfly steps
// never move a bit across the two words, the third swaps the high nibbles of lo's bytes with the low
// nibbles of hi's -- written on the 32-bit halves (the 64-bit form costs funnel shifts: 36 instead of 20
// instructions, 16 times per activation).
__device__ __forceinline__ void bfs_transpose8(u32 &lo, u32 &hi) {
    u32 t;
    t = (lo ^ (lo >> 7)) & 0x00AA00AAu; lo = lo ^ t ^ (t << 7);
    t = (hi ^ (hi >> 7)) & 0x00AA00AAu; hi = hi ^ t ^ (t << 7);
    t = (lo ^ (lo >> 14)) & 0x0000CCCCu; lo = lo ^ t ^ (t << 14);
    t = (hi ^ (hi >> 14)) & 0x0000CCCCu; hi = hi ^ t ^ (t << 14);
    const u32 nl = (lo & 0x0F0F0F0Fu) | ((hi << 4) & 0xF0F0F0F0u);
    hi = (hi & 0xF0F0F0F0u) | ((lo >> 4) & 0x0F0F0F0Fu);
    lo = nl;
}

// ---- ring ------------------------------------------------------------------------------------
// slot of a ticket: the capacity is a power of two wherever the host sized the ring itself (a mask instead of a
// 32-bit division on every poll); any capacity works
__device__ __forceinline__ u32 bfs_ring_slot(u32 ticket, u32 cap) { return (cap & (cap - 1u)) == 0u ? (ticket & (cap - 1u)) : ticket % cap; }
__device__ static inline void bfs_push(u64 *ring, u32 cap, u32 *ctr, u32 tile, u32 level) {
    const u32 ticket = atomicAdd(&ctr[BFS_C_TAIL], 1u);
    *(volatile u64 *)&ring[bfs_ring_slot(ticket, cap)] = ((u64)level << 32) | (u64)(tile + 1u);
}

// ---- multi-GPU: one FIFO served by all GPUs --------------------------------------------------------
// One GPU keeps the wave in order for free: a single FIFO, so a tile is served about one queue latency
// (~1 ms at 1536^3) after it was armed, by which time its other neighbours have delivered too.  With
// one FIFO per slab that is lost: a slab with a short queue serves a tile the moment it is armed, runs
// ahead along whatever route reaches it first, and lowers the same cells again and again as the bulk of
// the wave arrives (8 GPUs, 1536^3: 1.6x the cells lowered, the far slabs 2.8x -- profiles/r02_notes.md).
// So the slabs emulate ONE queue: every entry carries a global sequence number (the number of entries
// pushed on all ranks so far, as far as the pusher knows), every rank counts the entries it has started,
// and a rank starts its head entry only when the entries started on all ranks together come within
// `pace_window` of that number -- i.e. when a single FIFO over all slabs would have reached it.  The
// counts are exchanged every BFS_SEQ_PUB events per rank (plain remote stores, no atomics); an entry is
// held for at most pace_hold_polls polls, so pacing can delay but never deadlock.
#define BFS_SEQ_PUB 128u
__device__ static inline u32 bfs_seq_sum(const BfsArgs &A, const int base_word, const int own_word) {
    u32 s = __ldcg(&A.ctr[own_word]); // own count: exact
    for (int r = 0; r < A.world; ++r)
        if (r != A.rank) s += __ldcg(&A.ctr[base_word + r]);
    return s;
}
// one more entry pushed / started on this rank; tell the others every BFS_SEQ_PUB events
__device__ static inline void bfs_seq_bump(const BfsArgs &A, const int own_word, const int base_word) {
    const u32 n = atomicAdd(&A.ctr[own_word], 1u) + 1u;
    if ((n & (BFS_SEQ_PUB - 1u)) == 0u)
        for (int r = 0; r < A.world; ++r)
            if (r != A.rank) bfs_store_volatile(&A.peer_ctr[r][base_word + A.rank], n);
}

__device__ __noinline__ static void bfs_fail(const BfsArgs &A, u32 code) {
    atomicOr(&A.ctr[BFS_C_ERR], code);
    for (int r = 0; r < A.world; ++r) bfs_store_volatile(&A.peer_ctr[r][BFS_C_DONE], 1u);
}

// Streamed occupancy upload (GridDev.ready): have the planes [zlo, zhi] of the bitmask arrived?  Lanes 0 and 1 look
// at the flags of the first and the last chunk involved (chunks arrive in z order, and a tile's rows span at most
// two) with acquire loads -- no fence, this sits on every activation's path; lane 2 looks at the very last chunk.
// Returns 0: not yet (the caller puts the tile back into the ring: the tip of the wave runs along the z axis far
// ahead of the upload, and a warp that waited here would hold up its whole CTA at the next phase barrier), 1: there,
// 2: there and the upload is complete (the caller stops asking).
__device__ __noinline__ static u32 bfs_planes_there(const BfsArgs &A, int zlo, int zhi, const int lane) {
    const GridDev &Gd = A.G;
    u32 v = 1u;
    if (lane < 3) {
        zlo = zlo < 0 ? 0 : zlo;
        zhi = zhi > Gd.PZ - 1 ? Gd.PZ - 1 : zhi;
        v = bfs_ld_acquire_sys(Gd.ready + ((lane == 0 ? zlo : (lane == 1 ? zhi : Gd.PZ - 1)) >> Gd.ready_shift));
    }
    const u32 m = __ballot_sync(BFS_FULL, v != 0u);
    return (m & 3u) != 3u ? 0u : ((m & 4u) ? 2u : 1u);
}
// the tile's occupancy has not arrived: back into the ring with it (state untouched: it is still queued and counted
// in flight), after a short pause -- early in an upload the ring holds little else.  Bounded by the kernel's time-out.
__device__ __noinline__ static void bfs_put_back(const BfsArgs &A, const u64 entry, const int lane) {
    if (lane == 0) {
        __nanosleep(BFS_PUTBACK_NS);
        if (A.timeout_ns && A.stats && bfs_now_ns() - __ldcg(&A.stats[9]) > A.timeout_ns) bfs_fail(A, BFS_ERR_TIMEOUT);
        atomicAdd(&A.ctr[BFS_C_PUTBACK], 1u);
        bfs_push(A.ring, A.ring_cap, A.ctr, (u32)entry - 1u, (u32)(entry >> 32));
    }
    __syncwarp();
}

// A warp's position in the ring: the ticket stays valid across calls of bfs_pop_warp until it is
// served (identical in all lanes).
struct BfsPop {
    u32 ticket, have, held_polls;
    u64 held;       // an entry taken from the ring that waits for its turn in the global order (0: none)
    u64 t0;         // lane 0: when the current wait began (time-out safety net)
};
#define BFS_TOMB (~0ull) // ring slot: "the item of this ticket was taken by another warp"

// The whole warp: try to take a tile.  Returns the ring entry (sequence number << 32 | tile id + 1); or 0
// with `done` set when the search is over; or 0 with `done` clear when nothing turned up within
// max_polls polls (the ticket, or an entry held back by the pacing, is kept for the next call).
//
// Tickets spread the consumers over the slots, but an item is not bound to its ticket holder: a warp
// whose own slot is still empty also looks at the 31 slots before it and takes the oldest item it
// finds there.  Those belong to warps that cannot serve their ticket right now (in the phase-locked
// kernel a warp without a tile sits out its CTA's round at the barriers); without this an item
// could wait a whole round while other warps idle, and the wave would run out of order.  The taker
// leaves a tombstone, so the holder knows to draw a new ticket.
// MULTI: this rank is one of several slabs (pacing, peer queues); false strips all of it from the single-GPU kernel
template <bool MULTI>
__device__ static inline u64 bfs_pop_warp(const BfsArgs &A, BfsPop &P, const int lane, const unsigned max_polls,
                                          bool &done) {
    done = false;
    unsigned ns = 32;
    for (unsigned polls = 0;; ++polls) {
        u64 got = P.held;
        if (!got) {
            if (!P.have) {
                u32 t = 0;
                if (lane == 0) t = atomicAdd(&A.ctr[BFS_C_HEAD], 1u);
                P.ticket = __shfl_sync(BFS_FULL, t, 0);
                P.have = 1u;
                if (A.max_pops && P.ticket > A.max_pops) {
                    // tiles put back during a streamed upload are not activations (the count may have passed this
                    // ticket meanwhile: a "negative" difference is within budget)
                    const u32 used = P.ticket - (A.G.ready ? __ldcg(&A.ctr[BFS_C_PUTBACK]) : 0u);
                    if ((int)used > 0 && used > A.max_pops) {
                        if (lane == 0) bfs_fail(A, BFS_ERR_BUDGET);
                        done = true;
                        return 0ull;
                    }
                }
            }
            const u32 idx = P.ticket - 31u + (u32)lane; // lane 31: my own slot
            u64 *slot = &A.ring[bfs_ring_slot(idx, A.ring_cap)];
            const u64 v = (int)idx >= 0 ? __ldcg(slot) : 0ull;
            const u32 own_lo = __shfl_sync(BFS_FULL, (u32)v, 31), own_hi = __shfl_sync(BFS_FULL, (u32)(v >> 32), 31);
            if ((own_lo & own_hi) == 0xFFFFFFFFu) { // my item went to another warp: clear the slot, draw a new ticket
                if (lane == 31) *(volatile u64 *)slot = 0ull;
                P.have = 0u;
                continue;
            }
            const u32 m = __ballot_sync(BFS_FULL, v != 0ull && v != BFS_TOMB);
            if (m) {
                const int l = (m >> 31) ? 31 : __ffs(m) - 1; // my own, else the oldest one waiting for a busy warp
                u64 g = 0;
                if (lane == l) g = atomicCAS((unsigned long long *)slot, (unsigned long long)v, l == 31 ? 0ull : BFS_TOMB) == v ? v : 0ull;
                got = ((u64)__shfl_sync(BFS_FULL, (u32)(g >> 32), l) << 32) | (u64)__shfl_sync(BFS_FULL, (u32)g, l);
                if (!got) continue; // somebody else was faster: look again at once
                // no fence here: the entry only names a tile; the claim that follows (an atomic on the tile's
                // state word, then a fence) is what orders the field reads behind the neighbours' stores
                if (l == 31) P.have = 0u;
                P.t0 = 0;
                P.held_polls = 0;
            }
        }
        if (got) {
            // pacing (multi-GPU only): would a single FIFO over all slabs have reached this entry yet?
            u32 hold = 0;
            if (MULTI && A.pace_window) {
                if (lane == 0 && P.held_polls < A.pace_hold_polls && !__ldcg(&A.ctr[BFS_C_DONE]))
                    hold = (int)((u32)(got >> 32) - (bfs_seq_sum(A, BFS_C_GS, BFS_C_STARTED) + A.pace_window)) > 0 ? 1u : 0u;
                hold = __shfl_sync(BFS_FULL, hold, 0);
            }
            if (!hold) {
                P.held = 0ull;
                if (MULTI && A.pace_window && lane == 0) bfs_seq_bump(A, BFS_C_STARTED, BFS_C_GS);
                return got;
            }
            P.held = got;
            ++P.held_polls;
            if (polls >= max_polls) return 0ull;
            __nanosleep(256);
            continue;
        }
        u32 flag = 0;
        if (lane == 0) {
            if (__ldcg(&A.ctr[BFS_C_DONE])) flag = 1u;
            else if (A.timeout_ns && (polls & 1023u) == 1023u) {
                const u64 now = bfs_now_ns();
                if (!P.t0) P.t0 = now;
                else if (now - P.t0 > A.timeout_ns) { bfs_fail(A, BFS_ERR_TIMEOUT); flag = 1u; }
            }
        }
        flag = __shfl_sync(BFS_FULL, flag, 0);
        if (flag) { done = true; return 0ull; }
        if (polls >= max_polls) return 0ull;
        __nanosleep(ns);
        if (ns < BFS_SLEEP_CAP) ns *= 2;
    }
}

// this rank has nothing in flight any more: leave the set of active ranks; the last one ends
// the search everywhere
__device__ __noinline__ static void bfs_rank_idle(const BfsArgs &A) {
    __threadfence_system();
    if (atomicSub(A.active, 1u) == 1u) {
        for (int r = 0; r < A.world; ++r) bfs_store_volatile(&A.peer_ctr[r][BFS_C_DONE], 1u);
    }
}

// ---- the sweep ---------------------------------------------------------------------------------
// Multi-source BFS on the tile's bit lanes.  Seeds are given bit-sliced (6 planes: level relative
// to the pass base, valid mask sv; the candidates of all six faces merged, the smaller level where an
// edge cell has two); `init` cells start at level 0; `R` enters as the cells that must not be entered
// (blocked by an earlier pass) and returns with the cells reached added.  Levels are recorded
// bit-sliced in rec[7][2].  Returns in `carry` the cells that would get level BFS_LIM (0 if the
// wave died out), and the number of levels swept.
struct BfsSeeds {
    u32 sp[6][2];
    u32 sv[2];
    int last; // largest seed level (uniform), -1: none
};

// one block of BFS_UB levels starting at level rb.  (A second copy without the seed injection for the blocks beyond
// the last seed level was slower: +0.6 ms at 1536^3 -- code size.)
__device__ __forceinline__ void bfs_sweep_block(const int rb, const u32 free0, const u32 free1, const BfsSeeds &SA,
                                                u32 &c0, u32 &c1, u32 &R0, u32 &R1, u32 (&rec)[BFS_NREC][2]) {
    // seeds of this block: the high bits of their level equal those of rb
    u32 ha0 = 0, ha1 = 0;
    const bool useA = rb <= SA.last;
    if (useA) {
        ha0 = SA.sv[0]; ha1 = SA.sv[1];
#pragma unroll
        for (int b = BFS_LB; b < 6; ++b) {
            const u32 flip = ((rb >> b) & 1) ? 0u : BFS_FULL;
            ha0 &= SA.sp[b][0] ^ flip;
            ha1 &= SA.sp[b][1] ^ flip;
        }
    }
    const u32 rb0 = R0, rb1 = R1; // reached before this block
#pragma unroll
    for (int k = 0; k < BFS_UB; ++k) {
        u32 f0 = c0, f1 = c1;
        if (useA) {
            u32 m0 = ha0, m1 = ha1;
#pragma unroll
            for (int b = 0; b < BFS_LB; ++b) {
                m0 &= ((k >> b) & 1) ? SA.sp[b][0] : ~SA.sp[b][0];
                m1 &= ((k >> b) & 1) ? SA.sp[b][1] : ~SA.sp[b][1];
            }
            f0 |= m0 & ~R0;
            f1 |= m1 & ~R1;
        }
        R0 |= f0; R1 |= f1;
#pragma unroll
        for (int b = 0; b < BFS_LB; ++b)
            if ((k >> b) & 1) { rec[b][0] |= f0; rec[b][1] |= f1; }
        // neighbours of the frontier: y +- 1 inside a byte, z +- 1 = byte shifts across the two
        // words, x +- 1 = the adjacent lanes
        u32 n0 = ((f0 << 1) & 0xFEFEFEFEu) | ((f0 >> 1) & 0x7F7F7F7Fu) | (f0 << 8) | bfs_funnel_r(f0, f1, 8);
        u32 n1 = ((f1 << 1) & 0xFEFEFEFEu) | ((f1 >> 1) & 0x7F7F7F7Fu) | (f1 >> 8) | bfs_funnel_l(f0, f1, 8);
        const u32 u0 = __shfl_up_sync(BFS_FULL, f0, 1), u1 = __shfl_up_sync(BFS_FULL, f1, 1);
        const u32 d0 = __shfl_down_sync(BFS_FULL, f0, 1), d1 = __shfl_down_sync(BFS_FULL, f1, 1);
        // (lane 0 / 31 get their own word back from the shuffle: those cells are in R already)
        n0 |= u0 | d0;
        n1 |= u1 | d1;
        c0 = n0 & free0 & ~R0;
        c1 = n1 & free1 & ~R1;
    }
    const u32 bf0 = R0 & ~rb0, bf1 = R1 & ~rb1; // cells that got a level inside this block
#pragma unroll
    for (int b = BFS_LB; b < BFS_NREC; ++b)
        if ((rb >> b) & 1) { rec[b][0] |= bf0; rec[b][1] |= bf1; }
}

__device__ __forceinline__ u32 bfs_sweep(const u32 free0, const u32 free1, const BfsSeeds &SA,
                                         const u32 init0, const u32 init1, u32 &R0, u32 &R1, u32 (&rec)[BFS_NREC][2],
                                         u32 &carry0, u32 &carry1) {
#pragma unroll
    for (int b = 0; b < BFS_NREC; ++b) rec[b][0] = rec[b][1] = 0u;
    u32 c0 = init0, c1 = init1;
    const int last = SA.last;
    u32 nlev = 0;
    bool fin = false;
    for (int rb = 0; rb < BFS_LIM && !fin; rb += BFS_UB) {
        bfs_sweep_block(rb, free0, free1, SA, c0, c1, R0, R1, rec);
        nlev += BFS_UB;
        // the wave has died out (and no seed is left): asked once per block of levels -- a vote per level costs more
        // than the up to BFS_UB - 1 empty levels swept past the end
        if (rb + BFS_UB - 1 >= last && !__any_sync(BFS_FULL, (c0 | c1) != 0u)) fin = true;
    }
    carry0 = fin ? 0u : c0;
    carry1 = fin ? 0u : c1;
    return nlev;
}

// One (z, word) group of 8 cells: byte zb of the 7 planes + the "not reached" plane -> 8 x 8 bit
// matrix -> one level byte per cell -> field words.  Returns the lowered cells as 8 bits.
// TOUCHED: the tile already holds values -> min-merge, store only lowered cells; otherwise every word of the
// group is written (unreached / blocked cells = all-ones) and the lowered cells are the reached ones.
template <bool TOUCHED>
__device__ __forceinline__ u32 bfs_extract8(const u32 r0, const u32 r1, const u32 r2, const u32 r3, const u32 r4,
                                            const u32 r5, const u32 r6, const u32 Nw, const u32 zb, const u32 base2,
                                            u32 *pz, const size_t strideY) {
    const u32 sel = zb | ((4u + zb) << 4);
    const u32 t01 = __byte_perm(r0, r1, sel), t23 = __byte_perm(r2, r3, sel);
    const u32 t45 = __byte_perm(r4, r5, sel), t67 = __byte_perm(r6, ~Nw, sel);
    u32 lo = __byte_perm(t01, t23, 0x5410u), hi = __byte_perm(t45, t67, 0x5410u);
    bfs_transpose8(lo, hi);
    const u32 nb = (Nw >> (zb * 8u)) & 0xFFu; // cells of this group reached in this pass
    u32 low = 0;
    u32 *p = pz;
    const u32 lo2 = (lo << 1) & 0xFEFEFEFEu, hi2 = (hi << 1) & 0xFEFEFEFEu; // the level bytes doubled (one shift per four cells)
#pragma unroll
    for (int y = 0; y < 8; ++y) {
        const u32 src = y < 4 ? lo : hi;
        const u32 lv2 = __byte_perm(y < 4 ? lo2 : hi2, 0u, 0x4440u | (u32)(y & 3));
        // a cell that was not reached has bit 7 of its byte set -> an all-ones word ("unreached")
        const u32 word = bfs_byte_sign(src, y & 3) | (base2 + lv2);
        if (TOUCHED) {
            // min-merge: only reached cells can be lower (they are free cells, hence inside the row)
            if ((nb >> y) & 1u) {
                const u32 own = __ldcg(p);
                if (word < own) { bfs_store(p, word); low |= 1u << y; } // stored words order like their values
            }
        } else {
            bfs_store(p, word);
        }
        p += strideY;
    }
    return TOUCHED ? low : nb;
}

// ---- levels -> field words -----------------------------------------------------------------------
// For the cells in N (reached in this pass) the value is base + recorded level.  Adds the lowered cells
// to low[].  A runtime loop over the four byte positions keeps the code small; the two forms (first
// visit of a tile: plain stores of all 64 words of the lane's column, no branches; later visits:
// min-merge of the reached cells) have a loop each.
template <bool TOUCHED>
__device__ __forceinline__ void bfs_extract_loop(const u32 (&rec)[BFS_NREC][2], const u32 N0, const u32 N1, const u32 base2,
                                                 u32 *pz, const size_t strideY, const size_t strideZ, u32 &low0, u32 &low1) {
#pragma unroll 1
    for (u32 zb = 0; zb < 4; ++zb) {
        low0 |= bfs_extract8<TOUCHED>(rec[0][0], rec[1][0], rec[2][0], rec[3][0], rec[4][0], rec[5][0], rec[6][0], N0, zb, base2,
                                      pz, strideY) << (zb * 8u);
        low1 |= bfs_extract8<TOUCHED>(rec[0][1], rec[1][1], rec[2][1], rec[3][1], rec[4][1], rec[5][1], rec[6][1], N1, zb, base2,
                                      pz + 4 * strideZ, strideY) << (zb * 8u);
        pz += strideZ;
    }
}
__device__ __forceinline__ void bfs_extract(const u32 (&rec)[BFS_NREC][2], const u32 N0, const u32 N1, const u32 base,
                                            u32 *cell0, const bool inx, const size_t strideY, const size_t strideZ,
                                            const bool touched, u32 &low0, u32 &low1) {
    const u32 base2 = (base << 1) | 1u; // stored word of level 0
    if (touched) bfs_extract_loop<true>(rec, N0, N1, base2, cell0, strideY, strideZ, low0, low1);
    else if (inx) bfs_extract_loop<false>(rec, N0, N1, base2, cell0, strideY, strideZ, low0, low1);
    // (a lane whose column lies outside the box has no free cell: N = 0, nothing lowered)
}


// level of a candidate relative to the pass base as a byte (0x80: no seed in this pass); seeds
// beyond the six seed planes are left for a later activation (`far`)
__device__ __forceinline__ u32 bfs_rel(const u32 c, const u32 base, u32 &far, u32 &any, u32 &last) {
    if (c == BFS_CINV) return 0x80u;
    const u32 r = c - base;
    if (r > BFS_SEEDMAX) { far = 1u; return 0x80u; }
    any = 1u;
    last = r > last ? r : last;
    return r;
}

// Candidates read across a y face FA and a z face FB on the same side (0 = y-, 1 = y+, 2 = z-, 3 = z+): for each of the
// eight boundary cells of this lane's column the halo value + 1 if the cell is free and that beats its own value, else
// BFS_CINV.  hrow / orow: the first of the eight halo / own rows, `step` apart; f*_lo / f*_hi: the free-cell words holding
// the boundary cells 0..3 / 4..7.  The rows of both faces are requested before either is used.  (An unreached halo word
// gives 0x80000000, above every own value: no explicit test.)
template <int FA, int FB>
__device__ __forceinline__ void bfs_face_cand2(u32 (&ca)[8], u32 (&cb)[8], const bool da, const bool db,
                                               const u32 *hrowA, const u32 *orowA, const size_t stepA,
                                               const u32 *hrowB, const u32 *orowB, const size_t stepB, const bool peerB,
                                               const bool inx, const bool touched,
                                               const u32 fa_lo, const u32 fa_hi, const u32 fb_lo, const u32 fb_hi) {
#pragma unroll
    for (int i = 0; i < 8; ++i) ca[i] = cb[i] = BFS_CINV;
    if (!(da || db)) return;
    u32 ha[8], hb[8], oa[8], ob[8];
#pragma unroll
    for (int i = 0; i < 8; ++i) {
        ha[i] = (da && inx) ? __ldcg(hrowA + (size_t)i * stepA) : BFS_RAWINF;
        hb[i] = (db && inx) ? (peerB ? bfs_ld_peer(hrowB + (size_t)i * stepB) : __ldcg(hrowB + (size_t)i * stepB)) : BFS_RAWINF;
    }
    if (touched) {
#pragma unroll
        for (int i = 0; i < 8; ++i) {
            oa[i] = (da && inx) ? __ldcg(orowA + (size_t)i * stepA) : BFS_RAWINF;
            ob[i] = (db && inx) ? __ldcg(orowB + (size_t)i * stepB) : BFS_RAWINF;
        }
    } else {
#pragma unroll
        for (int i = 0; i < 8; ++i) oa[i] = ob[i] = BFS_RAWINF;
    }
    if (da) {
#pragma unroll
        for (int i = 0; i < 8; ++i) {
            const u32 cv = (ha[i] >> 1) + 1u;
            const u32 sh = (u32)((i & 3) * 8) + (FA ? 7u : 0u);
            if ((((i < 4 ? fa_lo : fa_hi) >> sh) & 1u) && cv < (oa[i] >> 1)) ca[i] = cv;
        }
    }
    if (db) {
#pragma unroll
        for (int i = 0; i < 8; ++i) {
            const u32 cv = (hb[i] >> 1) + 1u;
            const u32 sh = (u32)i + (FB == 3 ? 24u : 0u);
            if ((((i < 4 ? fb_lo : fb_hi) >> sh) & 1u) && cv < (ob[i] >> 1)) cb[i] = cv;
        }
    }
}

// The eight candidates of one y / z face (face F: 0 = y-, 1 = y+, 2 = z-, 3 = z+; index = z for y faces, y for z
// faces) -> seed planes.  Packed as bytes: level relative to the base, saturated at 0x80 (= no candidate); a byte
// with bit 6 set is a candidate beyond the seed planes (`farw` collects the bytes).
//   y faces: cell (y, z) of word z >> 2 sits at bit (z & 3) * 8 + y -- the packed bytes line up (y = 0 / 7);
//   z faces: cell (y, z = 0) is bit y of word 0, (y, z = 7) bit 24 + y of word 1 -- an 8 x 8 bit transpose of the
//   eight bytes yields the planes as bytes.
template <int F>
__device__ __forceinline__ void bfs_seed_face(const u32 (&c)[8], const u32 base, BfsSeeds &S, u32 &farw) {
    u32 p0 = 0, p1 = 0;
#pragma unroll
    for (int i = 0; i < 8; ++i) {
        const u32 d = c[i] - base; // >= 0; an absent candidate gives a huge value
        const u32 r = (d < 0x80u ? d : 0x80u) << ((i & 3) * 8);
        if (i < 4) p0 |= r; else p1 |= r;
    }
    farw |= p0 | p1;
    if (F < 2) {
        constexpr u32 sh = F ? 7u : 0u;
        const u32 v0 = (~(p0 | (p0 << 1)) >> 7) & 0x01010101u, v1 = (~(p1 | (p1 << 1)) >> 7) & 0x01010101u; // bits 6, 7 clear
        S.sv[0] |= v0 << sh; S.sv[1] |= v1 << sh;
#pragma unroll
        for (int b = 0; b < 6; ++b) {
            S.sp[b][0] |= ((p0 >> b) & v0) << sh;
            S.sp[b][1] |= ((p1 >> b) & v1) << sh;
        }
    } else {
        u32 lo = p0, hi = p1;
        bfs_transpose8(lo, hi); // byte b of hi:lo = bit b of the eight bytes: planes 0..3 in lo, 4, 5 in hi, then "far", "none"
        const u32 v = ~((hi >> 16) | (hi >> 24)) & 0xFFu;
        if (F == 2) {
            S.sv[0] |= v;
#pragma unroll
            for (int b = 0; b < 6; ++b) S.sp[b][0] |= ((b < 4 ? lo : hi) >> ((b & 3) * 8)) & v;
        } else {
            S.sv[1] |= v << 24;
#pragma unroll
            for (int b = 0; b < 6; ++b) S.sp[b][1] |= (((b < 4 ? lo : hi) >> ((b & 3) * 8)) & v) << 24;
        }
    }
}

// ---- one tile activation -----------------------------------------------------------------------
// SYNC: the phase-locked kernel (CTA barriers at the phase boundaries); false: free-running warps
template <bool SYNC, bool MULTI>
__device__ __forceinline__ void bfs_process_tile(const BfsArgs &A, const u64 entry, const int lane, BfsCount &cnt) {
    const u32 t = (u32)entry - 1u; // tile id (the upper half, the entry's sequence number, has served its purpose)
    const GridDev &Gd = A.G;
    const int bx = (int)(t % (u32)A.nbx), by = (int)((t / (u32)A.nbx) % (u32)A.nby), bz = (int)(t / ((u32)A.nbx * (u32)A.nby));
    const int x0 = bx * 32, y0 = by * 8, z0 = bz * 8;
    const size_t sY = (size_t)Gd.PX, sZ = (size_t)Gd.PX * Gd.PY;
    const bool inx = x0 + lane < Gd.PX;
    u32 *cell0 = A.g + (size_t)z0 * sZ + (size_t)y0 * sY + (size_t)(x0 + lane);
    const bool peer_lo = MULTI && bz == 0 && A.lo.g != nullptr, peer_hi = MULTI && bz == A.nbz - 1 && A.hi.g != nullptr;

    const long long tc0 = BFS_NOW();
    long long tmark = tc0;
    if (Gd.ready && !cnt.all_up) { // occupancy still uploading (e2e overlap)
        const u32 there = bfs_planes_there(A, z0 - 1, z0 + 8, lane);
        if (there == 0u) {
            bfs_put_back(A, entry, lane);
            if (SYNC && (BFS_SYNC_MASK & 1)) { BFS_PHASE_SYNC(); } // the phase boundaries the CTA's other warps take
            if (SYNC && (BFS_SYNC_MASK & 2)) { BFS_PHASE_SYNC(); }
            if (SYNC && (BFS_SYNC_MASK & 4)) { BFS_PHASE_SYNC(); }
            return;
        }
        cnt.all_up = there == 2u ? 1u : 0u;
    }
    // ---- occupancy words (read-only during a search): issued before the claim so that their
    // latency overlaps the claim's round trip ----
    const int ly = lane & 7, lzq = lane >> 3; // the (y, z & 3) this lane handles in row-distributed steps
    const u32 *brow = Gd.bits + ((size_t)(z0 + lzq) * Gd.PY + (size_t)(y0 + ly)) * Gd.wpr + bx;
    const size_t w1 = (size_t)4 * Gd.PY * Gd.wpr;
    const u32 oc0 = __ldg(brow), oc1 = __ldg(brow + w1);
    u32 onb = BFS_FULL; // rows behind the y / z faces; outside the box: blocked
    {
        const int i = lane & 7;
        if (lane < 8) { if (by > 0) onb = __ldg(Gd.bits + ((size_t)(z0 + i) * Gd.PY + (size_t)(y0 - 1)) * Gd.wpr + bx); }
        else if (lane < 16) { if (by < A.nby - 1) onb = __ldg(Gd.bits + ((size_t)(z0 + i) * Gd.PY + (size_t)(y0 + 8)) * Gd.wpr + bx); }
        else if (lane < 24) {
            if (bz > 0) onb = __ldg(Gd.bits + ((size_t)(z0 - 1) * Gd.PY + (size_t)(y0 + i)) * Gd.wpr + bx);
            else if (peer_lo) onb = 0u; // the peer's occupancy is not mapped here: assume free (conservative)
        } else {
            if (bz < A.nbz - 1) onb = __ldg(Gd.bits + ((size_t)(z0 + 8) * Gd.PY + (size_t)(y0 + i)) * Gd.wpr + bx);
            else if (peer_hi) onb = 0u;
        }
    }
    u32 xa0 = BFS_FULL, xa1 = BFS_FULL, xb0 = BFS_FULL, xb1 = BFS_FULL; // words left / right of the tile's
    if (bx > 0) { xa0 = __ldg(brow - 1); xa1 = __ldg(brow - 1 + w1); }
    if (bx < A.nbx - 1) { xb0 = __ldg(brow + 1); xb1 = __ldg(brow + 1 + w1); }

    // ---- claim: take the dirty faces, keep Q (nobody else may process or queue the tile) ----
    u32 old = 0;
    if (lane == 0) old = atomicAnd(&A.state[t], BFS_Q | BFS_T);
    old = __shfl_sync(BFS_FULL, old, 0);
    // acquire: the neighbours' rows written before they set our dirty bits (system scope where a
    // neighbour is another GPU)
    if (peer_lo || peer_hi) __threadfence_system(); else if (BFS_CLAIM_FENCE) bfs_fence();
    BFS_MARK(2, c_setup); // occupancy loads issued, claim, fence
    const u32 dirty = old & BFS_DIRTY;
    bool touched = (old & BFS_T) != 0u;
    if (MULTI && cnt.act == 0 && lane == 0 && A.stats && !__ldcg(&A.stats[18])) // [18]: when this rank started its first activation
        atomicCAS((unsigned long long *)&A.stats[18], 0ull, (unsigned long long)bfs_now_ns());
    ++cnt.act;

    // ---- occupancy rows -> this lane's column bits; free cells on the far side of each face (a
    // neighbour tile is only armed through them) ----
    u32 free0 = oc0, free1 = oc1, nf = onb;
    bfs_transpose32x3(free0, free1, nf, lane);
    free0 = ~free0; free1 = ~free1; nf = ~nf;
    u32 ymf0, ymf1, ypf0, ypf1, zmf, zpf, xmf0, xmf1, xpf0, xpf1;
    {
        ymf0 = bfs_spread4(nf); ymf1 = bfs_spread4(nf >> 4);
        ypf0 = bfs_spread4(nf >> 8) << 7; ypf1 = bfs_spread4(nf >> 12) << 7;
        zmf = (nf >> 16) & 0xFFu;
        zpf = nf & 0xFF000000u;
        xmf0 = __ballot_sync(BFS_FULL, !(xa0 >> 31)); xmf1 = __ballot_sync(BFS_FULL, !(xa1 >> 31));
        xpf0 = __ballot_sync(BFS_FULL, !(xb0 & 1u)); xpf1 = __ballot_sync(BFS_FULL, !(xb1 & 1u));
    }

    BFS_MARK(2, c_sweep); // transposes, face masks
    // ---- candidates across the dirty faces: halo + 1 where it beats the own boundary value ----
    // Face 0 = y-, 1 = y+, 2 = z-, 3 = z+ (dirty bit 4 << f); a clean face costs one uniform branch.
    u32 cYm[8], cYp[8], cZm[8], cZp[8], cXm[2], cXp[2];
    cXm[0] = cXm[1] = cXp[0] = cXp[1] = BFS_CINV;
    u32 *colrow = A.g + (size_t)(z0 + lzq) * sZ + (size_t)(y0 + ly) * sY + (size_t)x0; // x faces, row-distributed: cells (y = ly, z = lzq + 4w)
    {
        // halo rows: the neighbour's boundary rows (on the lowest / highest tile row of a slab the peer GPU's)
        const u32 *zlo = peer_lo ? A.lo.g + (size_t)A.lo.zadj * sZ + (size_t)y0 * sY + (size_t)(x0 + lane) : cell0 - sZ;
        const u32 *zhi = peer_hi ? A.hi.g + (size_t)A.hi.zadj * sZ + (size_t)y0 * sY + (size_t)(x0 + lane) : cell0 + 8 * sZ;
        // A wave usually enters through the faces of one corner (x-, y-, z- far from the source's corner, ...):
        // the halo words of the three "minus" faces are requested before the first is used, then those of the
        // three "plus" faces -- one L2 round trip per group instead of one per dirty face.
        u32 xh[2][2], xo[2][2];
#pragma unroll
        for (int w = 0; w < 2; ++w) {
            xh[0][w] = (dirty & 1u) ? __ldcg(colrow + (size_t)(4 * w) * sZ - 1) : BFS_RAWINF;
            xo[0][w] = ((dirty & 1u) && touched) ? __ldcg(colrow + (size_t)(4 * w) * sZ) : BFS_RAWINF;
        }
        bfs_face_cand2<0, 2>(cYm, cZm, (dirty & 4u) != 0u, (dirty & 16u) != 0u, cell0 - sY, cell0, sZ, zlo, cell0, sY, peer_lo, inx, touched,
                             free0, free1, free0, free0);
#pragma unroll
        for (int w = 0; w < 2; ++w) {
            xh[1][w] = (dirty & 2u) ? __ldcg(colrow + (size_t)(4 * w) * sZ + 32) : BFS_RAWINF;
            xo[1][w] = ((dirty & 2u) && touched) ? __ldcg(colrow + (size_t)(4 * w) * sZ + 31) : BFS_RAWINF;
        }
        bfs_face_cand2<1, 3>(cYp, cZp, (dirty & 8u) != 0u, (dirty & 32u) != 0u, cell0 + 8 * sY, cell0 + 7 * sY, sZ, zhi, cell0 + 7 * sZ, sY, peer_hi,
                             inx, touched, free0, free1, free1, free1);
        if (dirty & 3u) { // x-: halo column x0 - 1, own column x0 (lane 0's cells); x+: x0 + 32 / x0 + 31 (lane 31's)
            const u32 fm0 = __shfl_sync(BFS_FULL, free0, 0), fm1 = __shfl_sync(BFS_FULL, free1, 0);
            const u32 fp0 = __shfl_sync(BFS_FULL, free0, 31), fp1 = __shfl_sync(BFS_FULL, free1, 31);
#pragma unroll
            for (int w = 0; w < 2; ++w) {
                const u32 cm = (xh[0][w] >> 1) + 1u, cp = (xh[1][w] >> 1) + 1u;
                if ((((w ? fm1 : fm0) >> lane) & 1u) && cm < (xo[0][w] >> 1)) cXm[w] = cm;
                if ((((w ? fp1 : fp0) >> lane) & 1u) && cp < (xo[1][w] >> 1)) cXp[w] = cp;
            }
        }
        if (MULTI) cnt.peer_rows += (((dirty & 16u) && peer_lo) ? 8u : 0u) + (((dirty & 32u) && peer_hi) ? 8u : 0u);
    }
    // corner cells of the y / z faces have two candidates: keep the smaller one in the y array
    {
        u32 m;
        m = cYm[0] < cZm[0] ? cYm[0] : cZm[0]; cYm[0] = m; cZm[0] = BFS_CINV;
        m = cYp[0] < cZm[7] ? cYp[0] : cZm[7]; cYp[0] = m; cZm[7] = BFS_CINV;
        m = cYm[7] < cZp[0] ? cYm[7] : cZp[0]; cYm[7] = m; cZp[0] = BFS_CINV;
        m = cYp[7] < cZp[7] ? cYp[7] : cZp[7]; cYp[7] = m; cZp[7] = BFS_CINV;
    }
    // the source cell
    u32 init0 = 0, init1 = 0;
    bool src = false;
    if (old & BFS_SRC) {
        const int sx = (int)__ldcg(&A.ctr[BFS_C_SRC]) - x0, sy = (int)__ldcg(&A.ctr[BFS_C_SRC + 1]) - y0,
                  sz = (int)__ldcg(&A.ctr[BFS_C_SRC + 2]) - z0;
        src = true;
        if (lane == sx) { if (sz < 4) init0 = 1u << (sz * 8 + sy); else init1 = 1u << ((sz - 4) * 8 + sy); }
    }

    BFS_MARK(2, c_extract); // candidates of all faces
    // ---- pass base = smallest candidate; the largest one bounds the last seed level ----
    // (an absent candidate is all-ones: the largest unsigned and, read as signed, the smallest value)
    u32 base;
    int maxA = -1;
    {
        u32 m = BFS_CINV;
#pragma unroll
        for (int i = 0; i < 8; ++i) {
            m = cYm[i] < m ? cYm[i] : m; m = cYp[i] < m ? cYp[i] : m;
            m = cZm[i] < m ? cZm[i] : m; m = cZp[i] < m ? cZp[i] : m;
            maxA = (int)cYm[i] > maxA ? (int)cYm[i] : maxA; maxA = (int)cYp[i] > maxA ? (int)cYp[i] : maxA;
            maxA = (int)cZm[i] > maxA ? (int)cZm[i] : maxA; maxA = (int)cZp[i] > maxA ? (int)cZp[i] : maxA;
        }
        m = cXm[0] < m ? cXm[0] : m; m = cXm[1] < m ? cXm[1] : m;
        m = cXp[0] < m ? cXp[0] : m; m = cXp[1] < m ? cXp[1] : m;
        base = __reduce_min_sync(BFS_FULL, m);
        if (src) base = 0u;
    }

    u32 low0 = 0, low1 = 0; // cells lowered by this activation
    u32 redo = 0;           // dirty faces to look at again (seeds beyond the range of one pass)
    const bool active = base != BFS_CINV; // uniform: something to sweep from
    BfsSeeds SA;
#pragma unroll
    for (int b = 0; b < 6; ++b) SA.sp[b][0] = SA.sp[b][1] = 0u;
    SA.sv[0] = SA.sv[1] = 0u;
    SA.last = -1;
    if (active) {
        // ---- bit-sliced seed levels ----
        // y / z faces: the eight candidates of a face become eight bytes (level relative to the base; 0x40..0x7F: beyond
        // the six seed planes, left for a later activation; 0x80: no candidate) and the planes are cut from the packed
        // words -- a few instructions per face instead of a dozen per candidate.
        u32 farw = 0;
        bfs_seed_face<0>(cYm, base, SA, farw);
        bfs_seed_face<1>(cYp, base, SA, farw);
        bfs_seed_face<2>(cZm, base, SA, farw);
        bfs_seed_face<3>(cZp, base, SA, farw);
        u32 far = farw & 0x40404040u, lastB = 0, anyB = 0;
        const bool anyA = __any_sync(BFS_FULL, (SA.sv[0] | SA.sv[1]) != 0u); // any y / z seed within the planes
        // x faces: the ballots assemble lane 0's (x-) and lane 31's (x+) words, which are then merged into the same
        // planes.  An edge cell of those two columns can have a candidate from either side: the smaller level stays
        // (a bit-sliced compare, run only when there are x candidates at all) -- one seed set for the sweep.
#pragma unroll
        for (int w = 0; w < 2; ++w) {
            const u32 rm = bfs_rel(cXm[w], base, far, anyB, lastB), rp = bfs_rel(cXp[w], base, far, anyB, lastB);
            const bool any_x = __any_sync(BFS_FULL, (rm & rp & 0x80u) == 0u);
            if (any_x) {
                const u32 vm = __ballot_sync(BFS_FULL, !(rm & 0x80u)), vp = __ballot_sync(BFS_FULL, !(rp & 0x80u));
                const u32 vb = lane == 0 ? vm : (lane == 31 ? vp : 0u);
                u32 xb[6];
#pragma unroll
                for (int b = 0; b < 6; ++b) {
                    const u32 bm = __ballot_sync(BFS_FULL, (rm >> b) & 1u & ~(rm >> 7)), bp = __ballot_sync(BFS_FULL, (rp >> b) & 1u & ~(rp >> 7));
                    xb[b] = lane == 0 ? bm : (lane == 31 ? bp : 0u);
                }
                const u32 both = SA.sv[w] & vb;
                u32 lt = 0u, eq = both; // cells where the y / z candidate is below / equal to the x candidate so far (from the top bit down)
#pragma unroll
                for (int b = 5; b >= 0; --b) {
                    lt |= eq & ~SA.sp[b][w] & xb[b];
                    eq &= ~(SA.sp[b][w] ^ xb[b]);
                }
                const u32 take = vb & ~(lt | eq); // the x candidate wins: alone, or strictly smaller
#pragma unroll
                for (int b = 0; b < 6; ++b) SA.sp[b][w] = (SA.sp[b][w] & ~take) | (xb[b] & take);
                SA.sv[w] |= vb;
            }
        }
        {
            // last seed level: exact unless some y / z candidate lies beyond the planes (then 63, a bound)
            const int ma = __reduce_max_sync(BFS_FULL, maxA);
            const u32 la = (u32)ma - base;
            const u32 lb = __reduce_max_sync(BFS_FULL, lastB);
            const int lastA = anyA ? (int)(la < BFS_SEEDMAX ? la : BFS_SEEDMAX) : -1;
            const int lastX = __any_sync(BFS_FULL, anyB != 0u) ? (int)lb : -1;
            SA.last = lastA > lastX ? lastA : lastX;
            if (__any_sync(BFS_FULL, far != 0u)) redo = dirty;
        }

    }
    BFS_MARK(2, c_publish); // base + seed planes
    if (!BFS_PROF) cnt.c_setup += (u64)(BFS_NOW() - tc0);

    // ---- sweep, extract, store; continuation passes if a wave outlives the recordable range ----
    // (BFS_PHASE_SYNC: in the phase-locked kernel all warps of the CTA enter each phase together, so
    // that they share its instructions in the instruction cache)
    // One instance of the sweep and of the extraction in the code (the kernel is instruction-cache bound):
    // a wave that outlives the recordable range of a pass (rare: corridors winding inside a tile) goes round
    // the same loop again without the barriers -- every warp of the CTA takes each barrier exactly once.
    u32 R0 = 0, R1 = 0, pbase = base;
    u32 rec[BFS_NREC][2];
    u32 carry0 = 0, carry1 = 0;
    if (SYNC && (BFS_SYNC_MASK & 1)) { BFS_PHASE_SYNC(); }
    const long long tc1 = BFS_NOW();
    long long tsw = 0;
    for (int pass = 0;; ++pass) {
        const u32 P0 = R0, P1 = R1;
        const long long ta = BFS_NOW();
        if (active) {
            cnt.levels += bfs_sweep(free0, free1, SA, init0, init1, R0, R1, rec, carry0, carry1);
            ++cnt.passes;
        }
        tsw += BFS_NOW() - ta;
        if (SYNC && (BFS_SYNC_MASK & 2) && pass == 0) { BFS_PHASE_SYNC(); }
        if (!active) break;
        bfs_extract(rec, R0 & ~P0, R1 & ~P1, pbase, cell0, inx, sY, sZ, touched, low0, low1);
        touched = true;
        if (!__any_sync(BFS_FULL, (carry0 | carry1) != 0u)) break;
        init0 = carry0; init1 = carry1;
        pbase += BFS_LIM;
        SA.last = -1;
    }
    if (active) cnt.lowered += (u32)(__popc(low0) + __popc(low1));
    else ++cnt.idle;
    if (!BFS_PROF) { cnt.c_sweep += (u64)tsw; cnt.c_extract += (u64)(BFS_NOW() - tc1 - tsw); }
    if (SYNC && (BFS_SYNC_MASK & 4)) { BFS_PHASE_SYNC(); }

    // ---- publish: every store above is visible before a neighbour is armed ----
    const long long tc3 = BFS_NOW();
    tmark = tc3;
    if (peer_lo || peer_hi) __threadfence_system(); else bfs_fence();
    __syncwarp();
    // Which neighbours can use what this activation lowered?  Across a face that was NOT read in
    // this activation: any lowered boundary cell with a free cell behind it.  Across a face that
    // WAS read (the wave came from there -- the common case): only if a lowered cell actually beats
    // the neighbour's current value by more than the step, checked against the neighbour's
    // boundary cells (re-read: they can only have become smaller, so "no" stays "no").  Without
    // this every tile would re-activate its upstream neighbours for nothing.
    BFS_MARK(1, c_setup); // fence
    u32 arms = 0;
    {
        // x faces, row-distributed: this lane looks at the cells (y = ly, z = lzq + 4w) of lane 0 / 31
        const u32 a0 = __shfl_sync(BFS_FULL, low0 & xmf0, 0), a1 = __shfl_sync(BFS_FULL, low1 & xmf1, 0);
        const u32 b0 = __shfl_sync(BFS_FULL, low0 & xpf0, 31), b1 = __shfl_sync(BFS_FULL, low1 & xpf1, 31);
        if ((a0 | a1) != 0u) {
            bool imp = !(dirty & 1u);
            if (!imp) {
#pragma unroll
                for (int w = 0; w < 2; ++w)
                    if (((w ? a1 : a0) >> lane) & 1u) {
                        const u32 own = __ldcg(colrow + (size_t)(4 * w) * sZ), h = __ldcg(colrow + (size_t)(4 * w) * sZ - 1);
                        imp |= own + 2u < h;
                    }
            }
            arms |= (__any_sync(BFS_FULL, imp) ? 1u : 0u) << 0;
        }
        if ((b0 | b1) != 0u) {
            bool imp = !(dirty & 2u);
            if (!imp) {
#pragma unroll
                for (int w = 0; w < 2; ++w)
                    if (((w ? b1 : b0) >> lane) & 1u) {
                        const u32 own = __ldcg(colrow + (size_t)(4 * w) * sZ + 31), h = __ldcg(colrow + (size_t)(4 * w) * sZ + 32);
                        imp |= own + 2u < h;
                    }
            }
            arms |= (__any_sync(BFS_FULL, imp) ? 1u : 0u) << 1;
        }
        // y / z faces (this lane's own column) through one loop body, like the candidates: face f = 0 y-, 1 y+,
        // 2 z-, 3 z+ (arm bit 2 + f).  The lowered cells that have a free cell behind the face are gathered
        // into 8 bits (one per row of the face).
        const u32 cm[4] = {bfs_compress4(low0 & ymf0) | (bfs_compress4(low1 & ymf1) << 4),
                           bfs_compress4((low0 & ypf0) >> 7) | (bfs_compress4((low1 & ypf1) >> 7) << 4),
                           low0 & zmf, (low1 & zpf) >> 24};
    #pragma unroll 1
        for (int f = 0; f < 4; ++f) {
            const u32 c = f == 0 ? cm[0] : (f == 1 ? cm[1] : (f == 2 ? cm[2] : cm[3]));
            if (!__any_sync(BFS_FULL, c != 0u)) continue;
            bool imp = c != 0u;
            if ((dirty >> (f + 2)) & 1u) { // the wave came through this face: only real improvements count
                imp = false;
                const bool yface = f < 2, hiside = (f & 1) != 0;
                const bool peer = f == 2 ? peer_lo : (f == 3 ? peer_hi : false);
                const size_t step = yface ? sZ : sY;
                const u32 *hrow = yface ? (hiside ? cell0 + 8 * sY : cell0 - sY)
                                        : (f == 2 ? (peer_lo ? A.lo.g + (size_t)A.lo.zadj * sZ + (size_t)y0 * sY + (size_t)(x0 + lane) : cell0 - sZ)
                                                  : (peer_hi ? A.hi.g + (size_t)A.hi.zadj * sZ + (size_t)y0 * sY + (size_t)(x0 + lane) : cell0 + 8 * sZ));
                const u32 *orow = yface ? (hiside ? cell0 + 7 * sY : cell0) : (hiside ? cell0 + 7 * sZ : cell0);
                if (MULTI && peer) cnt.peer_rows += (u32)__popc(__reduce_or_sync(BFS_FULL, c));
#pragma unroll
                for (int i = 0; i < 8; ++i)
                    if ((c >> i) & 1u) {
                        const u32 own = __ldcg(orow + (size_t)i * step);
                        const u32 h = (MULTI && peer) ? bfs_ld_peer(hrow + (size_t)i * step) : __ldcg(hrow + (size_t)i * step);
                        imp |= own + 2u < h;
                    }
            }
            arms |= (__any_sync(BFS_FULL, imp) ? 1u : 0u) << (f + 2);
        }
    }

    BFS_MARK(1, c_sweep); // arming checks
    // sequence number of everything this activation queues: the entries pushed so far on all ranks
    u32 seq_out = 0;
    if (MULTI && A.pace_window) {
        if (lane == 0) seq_out = bfs_seq_sum(A, BFS_C_GP, BFS_C_PUSHED);
        seq_out = __shfl_sync(BFS_FULL, seq_out, 0);
    }
    // lanes 1..6 arm the six neighbours, lane 0 releases the tile -- all in parallel
    bool need_push = false, requeue = false;
    u32 ntile = 0;
    if (lane >= 1 && lane <= 6 && ((arms >> (lane - 1)) & 1u)) {
        const int f = lane - 1;
        const int nx = bx + (f == 0 ? -1 : (f == 1 ? 1 : 0)), ny = by + (f == 2 ? -1 : (f == 3 ? 1 : 0));
        const int nz = bz + (f == 4 ? -1 : (f == 5 ? 1 : 0));
        const u32 bit = (1u << (f ^ 1)) | BFS_Q;
        if (MULTI && (nz < 0 || nz >= A.nbz)) { // the neighbour tile lives on the peer GPU
            const BfsPeer &P = nz < 0 ? A.lo : A.hi;
            ntile = (u32)((P.bzadj * A.nby + ny) * A.nbx + nx);
            const u32 was = atomicOr(&P.state[ntile], bit);
            if (!(was & BFS_Q)) {
                // queue it over there: first make the peer (and, if it was idle, the set of active
                // ranks) count the tile, then hand it over
                if (atomicAdd(&P.ctr[BFS_C_INFLIGHT], 1u) == 0u)
                    cnt.remote += atomicAdd(A.active, 1u) == 0xFFFFFFFFu; // the returned value: performed before we go on
                __threadfence_system();
                if (A.pace_window) bfs_seq_bump(A, BFS_C_PUSHED, BFS_C_GP);
                bfs_push(P.ring, P.ring_cap, P.ctr, ntile, seq_out);
                ++cnt.remote;
            }
        } else {
            ntile = (u32)((nz * A.nby + ny) * A.nbx + nx);
            const u32 was = atomicOr(&A.state[ntile], bit);
            need_push = !(was & BFS_Q);
        }
    }
    if (lane == 0) {
        const u32 tbit = touched ? BFS_T : 0u; // the tile's words are initialised from now on
        if (redo) {
            atomicOr(&A.state[t], redo | tbit);
            requeue = true;
        } else {
            const u32 expect = BFS_Q | (old & BFS_T);
            const u32 was = atomicCAS(&A.state[t], expect, tbit);
            if (was != expect) { // new dirty faces arrived meanwhile: stay queued
                if (tbit) atomicOr(&A.state[t], tbit);
                requeue = true;
            }
        }
    }
    const u32 pushes = __ballot_sync(BFS_FULL, need_push || requeue);
    BFS_MARK(1, c_extract); // arming atomics + release
    const int np = __popc(pushes);
    if (np >= 2) {
        if (lane == 0) { atomicAdd(&A.ctr[BFS_C_INFLIGHT], (u32)(np - 1)); bfs_fence(); }
        __syncwarp();
    }
    if (need_push || requeue) {
        if (MULTI && A.pace_window) bfs_seq_bump(A, BFS_C_PUSHED, BFS_C_GP);
        bfs_push(A.ring, A.ring_cap, A.ctr, requeue ? t : ntile, seq_out);
        if (requeue) ++cnt.redo;
    }
    if (np == 0 && lane == 0) {
        if (atomicSub(&A.ctr[BFS_C_INFLIGHT], 1u) == 1u) bfs_rank_idle(A);
    }
    __syncwarp();
    BFS_MARK(1, c_publish); // pushes
    if (!BFS_PROF) cnt.c_publish += (u64)(BFS_NOW() - tc3);
    // [16]: when this rank finished its last activation (its busy span = [16] - [9]; the kernel itself runs
    // until the LAST rank is done)
    if (MULTI && lane == 0 && A.stats) atomicMax((unsigned long long *)&A.stats[16], (unsigned long long)bfs_now_ns());
}

// ---- persistent warp loop --------------------------------------------------------------------------
__device__ static inline void bfs_count_init(BfsCount &cnt) {
    cnt.act = cnt.idle = cnt.lowered = cnt.levels = cnt.passes = cnt.redo = cnt.remote = 0;
    cnt.peer_rows = 0;
    cnt.all_up = 0;
    cnt.c_pop = cnt.c_setup = cnt.c_sweep = cnt.c_extract = cnt.c_publish = 0;
}

__device__ static inline void bfs_count_flush(const BfsArgs &A, BfsCount &cnt, const int lane) {
    cnt.lowered = __reduce_add_sync(BFS_FULL, cnt.lowered); // per-lane counters
    cnt.remote = __reduce_add_sync(BFS_FULL, cnt.remote);
    if (lane == 0 && A.stats && cnt.act) {
        atomicAdd((unsigned long long *)&A.stats[1], (unsigned long long)cnt.act);
        atomicAdd((unsigned long long *)&A.stats[2], (unsigned long long)cnt.lowered);
        // phase cycles: [4] waiting for work, [6] claim + occupancy + halo candidates + seeds, [11] level
        // sweeps, [3] levels -> words + stores, [0] fence + arming + release + pushes
        atomicAdd((unsigned long long *)&A.stats[4], (unsigned long long)cnt.c_pop);
        atomicAdd((unsigned long long *)&A.stats[6], (unsigned long long)cnt.c_setup);
        atomicAdd((unsigned long long *)&A.stats[11], (unsigned long long)cnt.c_sweep);
        atomicAdd((unsigned long long *)&A.stats[3], (unsigned long long)cnt.c_extract);
        atomicAdd((unsigned long long *)&A.stats[0], (unsigned long long)cnt.c_publish);
        atomicAdd((unsigned long long *)&A.stats[12], (unsigned long long)cnt.levels);
        atomicAdd((unsigned long long *)&A.stats[13], (unsigned long long)cnt.idle);
        atomicAdd((unsigned long long *)&A.stats[14], (unsigned long long)cnt.passes);
        atomicAdd((unsigned long long *)&A.stats[5], (unsigned long long)cnt.redo);
        atomicAdd((unsigned long long *)&A.stats[7], (unsigned long long)cnt.remote);
        if (cnt.peer_rows) atomicAdd((unsigned long long *)&A.stats[14 + 3], (unsigned long long)cnt.peer_rows); // [17]
    }
}

// the first warp of a rank ends the search at once when the source was rejected (blocked /
// outside the box: k_bfs_init left a flag and queued nothing)
__device__ __noinline__ static void bfs_reject_check(const BfsArgs &A, const int lane, const bool first) {
    if (first && lane == 0 && __ldcg(&A.ctr[BFS_C_SRC + 3]))
        for (int r = 0; r < A.world; ++r) bfs_store_volatile(&A.peer_ctr[r][BFS_C_DONE], 1u);
}

// free-running form: every warp loops pop -> process on its own (the CPU emulator and the slab
// group kernel; the single-rank kernel k_bfs keeps the warps of a CTA in the same phase instead)
template <bool MULTI>
__device__ static inline void bfs_warp_main(const BfsArgs &A, const int lane, const bool first) {
    bfs_reject_check(A, lane, first);
    BfsCount cnt;
    bfs_count_init(cnt);
    BfsPop P;
    P.ticket = 0; P.have = 0; P.t0 = 0; P.held = 0; P.held_polls = 0;
    for (;;) {
        const long long tp = BFS_NOW();
        bool done = false;
        const u64 item = bfs_pop_warp<MULTI>(A, P, lane, 0xFFFFFFFFu, done);
        if (item == 0ull) break;
        cnt.c_pop += (u64)(BFS_NOW() - tp);
        bfs_process_tile<false, MULTI>(A, item, lane, cnt);
    }
    bfs_count_flush(A, cnt, lane);
}

// after the search: tiles that were never touched hold no values yet -> all-ones (unreached)
__device__ static inline void bfs_fill_tile(const BfsArgs &A, const u32 t, const int lane) {
    const GridDev &Gd = A.G;
    const int bx = (int)(t % (u32)A.nbx), by = (int)((t / (u32)A.nbx) % (u32)A.nby), bz = (int)(t / ((u32)A.nbx * (u32)A.nby));
    const size_t sY = (size_t)Gd.PX, sZ = (size_t)Gd.PX * Gd.PY;
    if (bx * 32 + lane >= Gd.PX) return;
    u32 *cell0 = A.g + (size_t)(bz * 8) * sZ + (size_t)(by * 8) * sY + (size_t)(bx * 32 + lane);
#pragma unroll
    for (int z = 0; z < 8; ++z)
#pragma unroll
        for (int y = 0; y < 8; ++y) bfs_store(cell0 + (size_t)z * sZ + (size_t)y * sY, BFS_RAWINF);
}
