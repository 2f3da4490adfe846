// simt_host.h -- a small SIMT emulator for the CPU tests (test infrastructure only).
//
// The unit-weight field kernel (python_motion_planning_b200/csrc/pmp_bfs_core.cuh) is written
// against a handful of warp intrinsics and atomics.  This header provides the same names for a
// plain g++ build: every lane of every warp is a ucontext fiber, warp collectives (__shfl_sync,
// __ballot_sync, ...) are rendezvous points of the 32 fibers of a warp, atomics are plain
// read-modify-writes (the fibers are cooperative, so they are atomic by construction), and the
// scheduler switches to another warp at atomics / __nanosleep with a seeded RNG so that the
// queue protocol is exercised under many interleavings.  The kernel source itself is compiled
// unchanged, so its bit arithmetic and its protocol logic are checked here before any GPU runs.
#pragma once
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include <ucontext.h>

#include <vector>

#define __device__
#define __host__
#define __forceinline__ inline
#define __noinline__

namespace simt {

struct Fiber {
    ucontext_t ctx;
    char *stack = nullptr;
    int warp = 0, lane = 0;
    bool done = false;
    bool waiting = false; // parked at a warp collective
};

struct Warp {
    uint64_t slot[32];
    int arrived = 0;
    int gen = 0;
    int alive = 32;
    uint64_t result = 0;
};

struct Sched {
    std::vector<Fiber> fibers;
    std::vector<Warp> warps;
    ucontext_t main_ctx;
    int cur = -1;
    uint64_t rng = 0x9E3779B97F4A7C15ull;
    uint64_t clock = 0;
    int preempt_every = 3; // switch warps at every n-th atomic on average
    int hint = -1;         // warp whose lanes should run next (a lane of it parked at a collective)
    void (*entry)(int warp, int lane, void *arg) = nullptr;
    void *arg = nullptr;
};

static Sched *S = nullptr;

static inline uint64_t rnd() {
    S->rng ^= S->rng << 13; S->rng ^= S->rng >> 7; S->rng ^= S->rng << 17;
    return S->rng;
}

static void trampoline() {
    Fiber &f = S->fibers[S->cur];
    S->entry(f.warp, f.lane, S->arg);
    f.done = true;
    Warp &w = S->warps[f.warp];
    --w.alive;
    swapcontext(&f.ctx, &S->main_ctx);
}

static inline void switch_to_main() {
    Fiber &f = S->fibers[S->cur];
    swapcontext(&f.ctx, &S->main_ctx);
}

// run all fibers until every one has finished
static inline void run(int n_warps, void (*entry)(int, int, void *), void *arg, uint64_t seed, int preempt_every = 3) {
    Sched sched;
    S = &sched;
    sched.entry = entry;
    sched.arg = arg;
    sched.rng = seed * 0x9E3779B97F4A7C15ull + 0x1234567ull;
    sched.preempt_every = preempt_every;
    sched.fibers.resize((size_t)n_warps * 32);
    sched.warps.resize(n_warps);
    const size_t STK = 256 << 10;
    for (int i = 0; i < n_warps * 32; ++i) {
        Fiber &f = sched.fibers[i];
        f.warp = i / 32;
        f.lane = i % 32;
        f.stack = (char *)malloc(STK);
        getcontext(&f.ctx);
        f.ctx.uc_stack.ss_sp = f.stack;
        f.ctx.uc_stack.ss_size = STK;
        f.ctx.uc_link = nullptr;
        makecontext(&f.ctx, (void (*)())trampoline, 0);
    }
    // scheduler loop: pick a runnable fiber (not done, not parked at an incomplete collective)
    size_t live = sched.fibers.size();
    int rr = 0;
    uint64_t idle_spins = 0;
    while (live) {
        const int n = (int)sched.fibers.size();
        int pick = -1;
        (void)rr;
        if (sched.hint >= 0) { // finish the collective of that warp first (keeps the emulation fast)
            for (int l = 0; l < 32; ++l) {
                Fiber &f = sched.fibers[(size_t)sched.hint * 32 + l];
                if (!f.done && !f.waiting) { pick = sched.hint * 32 + l; break; }
            }
            sched.hint = -1;
        }
        if (pick < 0) {
            const int start = (int)(rnd() % (uint64_t)n);
            for (int k = 0; k < n; ++k) {
                const int i = (start + k) % n;
                Fiber &f = sched.fibers[i];
                if (!f.done && !f.waiting) { pick = i; break; }
            }
        }
        if (pick < 0) {
            fprintf(stderr, "simt: deadlock -- every live fiber waits at a warp collective\n");
            abort();
        }
        sched.cur = pick;
        swapcontext(&sched.main_ctx, &sched.fibers[pick].ctx);
        if (sched.fibers[pick].done) { --live; idle_spins = 0; }
        if (++idle_spins > (1ull << 34)) { fprintf(stderr, "simt: livelock\n"); abort(); }
    }
    for (auto &f : sched.fibers) free(f.stack);
    S = nullptr;
}

static inline int lane_id() { return S->fibers[S->cur].lane; }
static inline int warp_id() { return S->fibers[S->cur].warp; }

// yield the processor to some other fiber (possibly of another warp)
static inline void yield() { switch_to_main(); }

static inline void maybe_preempt() {
    if (S->preempt_every > 0 && (rnd() % (uint64_t)S->preempt_every) == 0) yield();
}

} // namespace simt

// Warp rendezvous: every lane of the warp deposits a value and continues once all 32 have
// arrived.  Double-buffered on the generation parity: a released lane may deposit its next value
// before a slower lane has read the previous generation.
namespace simt {
struct WarpBuf {
    uint64_t slot[2][32];
};
static inline const uint64_t *rendezvous(uint64_t v) {
    Fiber &f = S->fibers[S->cur];
    Warp &w = S->warps[f.warp];
    static thread_local std::vector<WarpBuf> bufs;
    if (bufs.size() < S->warps.size()) bufs.resize(S->warps.size());
    if (w.alive != 32) { fprintf(stderr, "simt: collective after a lane of the warp exited\n"); abort(); }
    const int gen = w.gen;
    WarpBuf &b = bufs[f.warp];
    b.slot[gen & 1][f.lane] = v;
    ++w.arrived;
    if (w.arrived == 32) {
        w.arrived = 0;
        ++w.gen;
        for (int l = 0; l < 32; ++l) S->fibers[(size_t)f.warp * 32 + l].waiting = false;
    } else {
        f.waiting = true;
        while (w.gen == gen) { S->hint = f.warp; switch_to_main(); }
    }
    // A lane can be at most one generation ahead of the slowest lane (it blocks in the next
    // rendezvous until everybody has arrived there, i.e. has left this one), so the buffer of
    // parity gen & 1 is intact until every lane has read it.
    return b.slot[gen & 1];
}
} // namespace simt

#define SIMT_FULL 0xFFFFFFFFu

static inline uint32_t __shfl_sync(unsigned, uint32_t v, int src) { return (uint32_t)simt::rendezvous(v)[src & 31]; }
static inline uint32_t __shfl_up_sync(unsigned, uint32_t v, int d) {
    const int l = simt::lane_id();
    const uint64_t *s = simt::rendezvous(v);
    return l - d >= 0 ? (uint32_t)s[l - d] : v;
}
static inline uint32_t __shfl_down_sync(unsigned, uint32_t v, int d) {
    const int l = simt::lane_id();
    const uint64_t *s = simt::rendezvous(v);
    return l + d < 32 ? (uint32_t)s[l + d] : v;
}
static inline uint32_t __shfl_xor_sync(unsigned, uint32_t v, int m) {
    const int l = simt::lane_id();
    return (uint32_t)simt::rendezvous(v)[(l ^ m) & 31];
}
static inline uint32_t __ballot_sync(unsigned, int p) {
    const uint64_t *s = simt::rendezvous(p ? 1 : 0);
    uint32_t r = 0;
    for (int l = 0; l < 32; ++l) r |= (uint32_t)(s[l] & 1) << l;
    return r;
}
static inline int __any_sync(unsigned m, int p) { return __ballot_sync(m, p) != 0; }
static inline int __all_sync(unsigned m, int p) { return __ballot_sync(m, p) == 0xFFFFFFFFu; }
static inline uint32_t __reduce_min_sync(unsigned, uint32_t v) {
    const uint64_t *s = simt::rendezvous(v);
    uint32_t r = 0xFFFFFFFFu;
    for (int l = 0; l < 32; ++l) r = (uint32_t)s[l] < r ? (uint32_t)s[l] : r;
    return r;
}
static inline uint32_t __reduce_max_sync(unsigned, uint32_t v) {
    const uint64_t *s = simt::rendezvous(v);
    uint32_t r = 0;
    for (int l = 0; l < 32; ++l) r = (uint32_t)s[l] > r ? (uint32_t)s[l] : r;
    return r;
}
static inline uint32_t __reduce_or_sync(unsigned, uint32_t v) {
    const uint64_t *s = simt::rendezvous(v);
    uint32_t r = 0;
    for (int l = 0; l < 32; ++l) r |= (uint32_t)s[l];
    return r;
}
static inline uint32_t __reduce_add_sync(unsigned, uint32_t v) {
    const uint64_t *s = simt::rendezvous(v);
    uint32_t r = 0;
    for (int l = 0; l < 32; ++l) r += (uint32_t)s[l];
    return r;
}
static inline void __syncwarp(unsigned = 0xFFFFFFFFu) { simt::rendezvous(0); }

static inline int __popc(uint32_t v) { return __builtin_popcount(v); }
static inline int __ffs(uint32_t v) { return __builtin_ffs((int)v); }
static inline int __clz(uint32_t v) { return v ? __builtin_clz(v) : 32; }
static inline uint32_t __funnelshift_l(uint32_t lo, uint32_t hi, uint32_t s) {
    s &= 31;
    return s ? (hi << s) | (lo >> (32 - s)) : hi;
}
static inline uint32_t __funnelshift_r(uint32_t lo, uint32_t hi, uint32_t s) {
    s &= 31;
    return s ? (lo >> s) | (hi << (32 - s)) : lo;
}
static inline uint32_t __byte_perm(uint32_t a, uint32_t b, uint32_t sel) {
    const uint64_t v = (uint64_t)a | ((uint64_t)b << 32);
    uint32_t r = 0;
    for (int i = 0; i < 4; ++i) {
        const uint32_t n = (sel >> (4 * i)) & 0xF;
        const uint32_t byte = (uint32_t)(v >> (8 * (n & 7))) & 0xFFu; // __byte_perm uses 3 bits per selector: no sign mode
        r |= byte << (8 * i);
    }
    return r;
}

// atomics: cooperative fibers -> plain read-modify-write, then maybe hand the CPU to another warp
template <class T> static inline T atomicAdd(T *p, T v) { T o = *p; *p = o + v; simt::maybe_preempt(); return o; }
template <class T> static inline T atomicSub(T *p, T v) { T o = *p; *p = o - v; simt::maybe_preempt(); return o; }
template <class T> static inline T atomicOr(T *p, T v) { T o = *p; *p = o | v; simt::maybe_preempt(); return o; }
template <class T> static inline T atomicAnd(T *p, T v) { T o = *p; *p = o & v; simt::maybe_preempt(); return o; }
template <class T> static inline T atomicExch(T *p, T v) { T o = *p; *p = v; simt::maybe_preempt(); return o; }
template <class T> static inline T atomicMin(T *p, T v) { T o = *p; *p = o < v ? o : v; simt::maybe_preempt(); return o; }
template <class T> static inline T atomicMax(T *p, T v) { T o = *p; *p = o > v ? o : v; simt::maybe_preempt(); return o; }
template <class T> static inline T atomicCAS(T *p, T c, T v) { T o = *p; if (o == c) *p = v; simt::maybe_preempt(); return o; }
template <class T> static inline T __ldcg(const T *p) { return *(const volatile T *)p; }
template <class T> static inline T __ldg(const T *p) { return *p; }
static inline void __threadfence() { simt::maybe_preempt(); }
static inline void __threadfence_system() { simt::maybe_preempt(); }
static inline void __nanosleep(unsigned) { simt::S->clock += 100; simt::yield(); }
static inline long long clock64() { return (long long)(simt::S->clock += 1); }
