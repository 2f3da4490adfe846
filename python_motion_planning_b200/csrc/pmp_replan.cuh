// pmp_replan.cuh -- incremental replanning on a resident cost field (included by pmp_sssp.cu).
//
// Replaces the repair loop of the reference's incremental planners: LPAStar.OnPress /
// LPAStar3D.apply_change + computeShortestPath + updateVertex
// (global_planner/graph_search/lpa_star.py:102-177, lpa_star3d.py:98-177) and the modify loop
// of DStar3D.apply_dynamic_obstacles (d_star3d.py:115-151).  The reference keeps (g, rhs) per
// cell and a priority queue of inconsistent cells; here the field g of the previous search
// stays in HBM in the search's own word format and a change is applied in three steps:
//
//   1. k_replan_bits   flips the occupancy bits;
//   2. k_replan_seed / k_replan_invalidate
//                      RAISE: a cell whose value is no longer explained by any legal in-edge
//                      (no u with fl(g(u) + w) == g(v)) is reset to "unreached" and its
//                      out-neighbours are re-examined, wave by wave, until every remaining value
//                      has a support chain back to the source (values only shrink along a chain,
//                      so the fixed point does not depend on the order).  Every reached
//                      in-neighbour of a reset cell, and every reached neighbour of a freed
//                      cell, gets its pending flag back and arms its tile;
//   3. the ordinary k_sssp kernel resumes from that state (LOWER).  Label correcting converges
//      to the shortest-path field from any state in which every value is the length of a real
//      path and every cell that could still lower a neighbour is pending -- which is exactly
//      what step 2 leaves behind -- so the repaired field equals the field of a search from
//      scratch, bit for bit.
//
// Searches that are going to be repaired run without the g[goal] pruning of relax_record: a
// relaxation dropped because it could not beat the old cost may be needed once the cost rises.

#define PMP_REPLAN_CTL 16 // u32 control words in front of the two candidate buffers

// mark a reached cell pending again and make sure its tile is scheduled
template <class C>
__device__ static inline void replan_arm_cell(const SsspArgs &A, int x, int y, int z, int gx, int gy, int gz) {
    typedef typename C::W W;
    typedef typename C::G G;
    const GridDev &Gd = A.G;
    G *p = (G *)A.g + ((size_t)z * Gd.PY + y) * Gd.PX + x;
    const G w = ld_cg(p);
    if (w == W::ginf()) return;
    if (w & 1) cas_word(p, w, w & ~(G)1); // a failed CAS means: already pending, or just reset
    const double h = A.use_h ? heuristic_lo(x - gx, y - gy, z - gz, A.he, A.hm) : 0.0;
    const u32 t = (u32)(((z / C::TZ) * Gd.nty + (y / C::TY)) * Gd.ntx + (x / C::TX));
    atomicMin(&A.minkey[t], W::key(W::dec(w), h));
}

__global__ void k_replan_bits(GridDev Gd, u32 *bits, const int32_t *cells, const uint8_t *blocked, u32 n) {
    for (u32 i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) {
        const int x = cells[3 * i], y = cells[3 * i + 1], z = Gd.ndim == 3 ? cells[3 * i + 2] : 0;
        if (x < 0 || y < 0 || z < 0 || x >= Gd.X || y >= Gd.Y || z >= Gd.Z) continue;
        u32 *w = bits + ((size_t)z * Gd.PY + y) * Gd.wpr + (x >> 5);
        if (blocked[i]) atomicOr(w, 1u << (x & 31));
        else atomicAnd(w, ~(1u << (x & 31)));
    }
}

// one warp per changed cell, one lane per cell of its 3x3x3 neighbourhood
template <class C>
__global__ void k_replan_seed(SsspArgs A, const int32_t *cells, const uint8_t *blocked, u32 n,
                              u32 *ctl, u32 *queue, u32 qcap) {
    typedef typename C::W W;
    typedef typename C::G G;
    const GridDev &Gd = A.G;
    const int lane = threadIdx.x & 31;
    const u32 wid = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, nw = (gridDim.x * blockDim.x) >> 5;
    const int sx = A.starts[0], sy = A.starts[1], sz = Gd.ndim == 3 ? A.starts[2] : 0;
    int gx = 0, gy = 0, gz = 0;
    if (A.goals) { gx = A.goals[0]; gy = A.goals[1]; gz = Gd.ndim == 3 ? A.goals[2] : 0; }
    for (u32 i = wid; i < n; i += nw) {
        const int cx = cells[3 * i], cy = cells[3 * i + 1], cz = Gd.ndim == 3 ? cells[3 * i + 2] : 0;
        if (cx < 0 || cy < 0 || cz < 0 || cx >= Gd.X || cy >= Gd.Y || cz >= Gd.Z) continue;
        const bool now_blocked = blocked[i] != 0;
        const int x = cx + lane % 3 - 1, y = cy + (lane / 3) % 3 - 1, z = cz + lane / 9 - 1;
        const bool in = lane < 27 && x >= 0 && y >= 0 && z >= 0 && x < Gd.X && y < Gd.Y && z < Gd.Z &&
                        (Gd.ndim == 3 || z == 0);
        G *p = (G *)A.g + ((size_t)(in ? z : 0) * Gd.PY + (in ? y : 0)) * Gd.PX + (in ? x : 0);
        bool cand = false;
        if (in && lane == 13) {
            if (now_blocked) {
                *p = W::ginf(); // a blocked cell never carries a value
            } else if (x == sx && y == sy && z == sz && *p == W::ginf()) {
                *p = W::enc((G)0, true); // the source itself was freed
                const double h = A.use_h ? heuristic_lo(x - gx, y - gy, z - gz, A.he, A.hm) : 0.0;
                const u32 t = (u32)(((z / C::TZ) * Gd.nty + (y / C::TY)) * Gd.ntx + (x / C::TX));
                atomicMin(&A.minkey[t], W::key((G)0, h));
            }
        } else if (in && !cell_blocked(Gd, x, y, z)) {
            if (now_blocked) cand = ld_cg(p) != W::ginf(); // may have lost its support
            else replan_arm_cell<C>(A, x, y, z, gx, gy, gz); // may reach the freed cell / use new edges
        }
        const u32 m = __ballot_sync(0xFFFFFFFFu, cand);
        if (m) {
            u32 pos = 0;
            if (lane == 0) pos = atomicAdd(&ctl[0], (u32)__popc(m));
            pos = __shfl_sync(0xFFFFFFFFu, pos, 0) + __popc(m & ((1u << lane) - 1u));
            if (cand && pos < qcap) queue[pos] = (u32)(((size_t)z * Gd.PY + y) * Gd.PX + x);
        }
    }
}

// RAISE phase.  Cooperative launch; a wave = the candidates queued by the previous wave, one warp
// per candidate, one lane per motion.  ctl[0..2]: entry counts of wave w, w+1, w+2 (rotating, so
// that a counter is reset a full wave before it is filled again); ctl[3]: overflow; ctl[4]: cells
// reset; ctl[5]: waves.  Candidate buffers: queue[(w & 1) * qcap ...].
template <class C>
__global__ void __launch_bounds__(256) k_replan_invalidate(const __grid_constant__ SsspArgs A,
                                                           const __grid_constant__ Motions M,
                                                           u32 *ctl, u32 *queue, u32 qcap) {
    typedef typename C::W W;
    typedef typename C::G G;
    cg::grid_group grid = cg::this_grid();
    const GridDev &Gd = A.G;
    const int lane = threadIdx.x & 31;
    const u32 wid = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, nw = (gridDim.x * blockDim.x) >> 5;
    const int sx = A.starts[0], sy = A.starts[1], sz = Gd.ndim == 3 ? A.starts[2] : 0;
    int gx = 0, gy = 0, gz = 0;
    if (A.goals) { gx = A.goals[0]; gy = A.goals[1]; gz = Gd.ndim == 3 ? A.goals[2] : 0; }
    G *gq = (G *)A.g;
    const bool has = lane < M.k;
    const int dx = has ? M.dx[lane] : 0, dy = has ? M.dy[lane] : 0, dz = has ? M.dz[lane] : 0;
    u32 n_reset = 0;
    for (u32 w = 0;; ++w) {
        const u32 n = ld_cg(&ctl[w % 3]);
        if (n == 0) break;
        if (n > qcap) { // candidates were dropped: the caller recomputes from scratch
            if (blockIdx.x == 0 && threadIdx.x == 0) ctl[3] = 1;
            break;
        }
        if (blockIdx.x == 0 && threadIdx.x == 0) { ctl[(w + 2) % 3] = 0; ctl[5] = w + 1; }
        const u32 *in = queue + (size_t)(w & 1) * qcap;
        u32 *out = queue + (size_t)((w + 1) & 1) * qcap;
        u32 *out_n = &ctl[(w + 1) % 3];
        for (u32 i = wid; i < n; i += nw) {
            const u32 v = ld_cg(in + i);
            const int vx = (int)(v % Gd.PX), vy = (int)((v / Gd.PX) % Gd.PY), vz = (int)(v / ((size_t)Gd.PX * Gd.PY));
            const G word = ld_cg(gq + v);
            if (word == W::ginf()) continue;                    // reset by an earlier candidate entry
            if (vx == sx && vy == sy && vz == sz) continue;     // the source needs no support
            const G gv = W::dec(word);
            bool supported = false;
            if (has) {
                const int ux = vx - dx, uy = vy - dy, uz = vz - dz;
                if (edge_legal(Gd, M, lane, ux, uy, uz)) { // false outside the padded box
                    const G gu = W::dec(ld_cg(gq + ((size_t)uz * Gd.PY + uy) * Gd.PX + ux));
                    supported = gu != W::inf() && W::add(gu, M, lane) == gv;
                }
            }
            if (__any_sync(0xFFFFFFFFu, supported)) continue;
            G old = 0;
            if (lane == 0) old = atomicExch(gq + v, W::ginf());
            old = __shfl_sync(0xFFFFFFFFu, old, 0);
            if (old == W::ginf()) continue;                     // another warp got there first
            ++n_reset;
            // out-neighbours: re-examine; in-neighbours: they have to propagate into v again
            bool push = false;
            u32 wcell = 0;
            if (has) {
                const int wx = vx + dx, wy = vy + dy, wz = vz + dz;
                if ((unsigned)wx < (unsigned)Gd.PX && (unsigned)wy < (unsigned)Gd.PY && (unsigned)wz < (unsigned)Gd.PZ) {
                    wcell = (u32)(((size_t)wz * Gd.PY + wy) * Gd.PX + wx);
                    push = ld_cg(gq + wcell) != W::ginf();
                }
                const int ux = vx - dx, uy = vy - dy, uz = vz - dz;
                if ((unsigned)ux < (unsigned)Gd.PX && (unsigned)uy < (unsigned)Gd.PY && (unsigned)uz < (unsigned)Gd.PZ)
                    replan_arm_cell<C>(A, ux, uy, uz, gx, gy, gz);
            }
            const u32 m = __ballot_sync(0xFFFFFFFFu, push);
            if (m) {
                u32 pos = 0;
                if (lane == 0) pos = atomicAdd(out_n, (u32)__popc(m));
                pos = __shfl_sync(0xFFFFFFFFu, pos, 0) + __popc(m & ((1u << lane) - 1u));
                if (push && pos < qcap) out[pos] = wcell;
            }
        }
        grid.sync();
    }
    if (lane == 0 && n_reset) atomicAdd(&ctl[4], n_reset);
}

// scheduling state for the resumed search: nothing is in any window yet (T = 0), so the first
// round only measures the smallest pending key and the second one starts working
__global__ void k_replan_arm(SsspArgs A, const u32 *ctl) {
    A.qmin[0][0] = KEY_INF; A.qmin[1][0] = KEY_INF;
    A.T[0][0] = 0; A.T[1][0] = 0;
    A.done[0] = ctl[3] ? PMP_ST_REPLAN_OVERFLOW : 0;
    for (int i = 0; i < 16; ++i) A.ctr[i] = 0;
}

// ---------------------------------------------------------------------------------------------
// host side
// ---------------------------------------------------------------------------------------------
static int sssp_from_scratch(pmp_grid *g, const pmp_sssp_params *P, int64_t nq, const int32_t *starts_dev,
                             const int32_t *goals_dev, void *g_dev, void *ws_dev, int64_t ws_bytes,
                             double *gcost_dev, int32_t *status_dev, uint64_t *stats_dev, void *stream,
                             int no_prune);

static int replan_check(pmp_grid *g, const pmp_sssp_params *P, const int32_t *start_dev, const int32_t *goal_dev,
                        void *g_dev, void *ws_dev, int64_t ws_bytes, WsLayout &L) {
    if (!g || !P || !start_dev || !g_dev || !ws_dev) { pmp_set_error("pmp_replan: bad arguments"); return -1; }
    if (!goal_dev && !P->full_field) { pmp_set_error("pmp_replan: goal_dev is NULL but full_field is 0"); return -1; }
    if (P->weight != PMP_W_F64 && P->weight != PMP_W_U32) { pmp_set_error("pmp_replan: bad weight type"); return -1; }
    if (P->weight == PMP_W_U32 && !g->integer_costs) { pmp_set_error("pmp_replan: PMP_W_U32 needs integer motion costs"); return -1; }
    if (P->algo != PMP_ASTAR && P->algo != PMP_DIJKSTRA) { pmp_set_error("pmp_replan: algo must be PMP_ASTAR or PMP_DIJKSTRA"); return -1; }
    L = ws_layout(g, 1);
    if ((size_t)ws_bytes < L.total) { pmp_set_error("pmp_replan: workspace too small"); return -1; }
    return 0;
}

extern "C" int pmp_replan_begin(pmp_grid *g, const pmp_sssp_params *P, const int32_t *start_dev,
                                const int32_t *goal_dev, void *g_dev, void *ws_dev, int64_t ws_bytes,
                                double *gcost_dev, int32_t *status_dev, uint64_t *stats_dev, void *stream) {
    if (g && pmp_grid_upload_fence((pmp_grid *)g, (cudaStream_t)stream)) return -2; // a streamed occupancy upload must have landed
    WsLayout L;
    if (replan_check(g, P, start_dev, goal_dev, g_dev, ws_dev, ws_bytes, L)) return -1;
    return sssp_from_scratch(g, P, 1, start_dev, goal_dev, g_dev, ws_dev, ws_bytes, gcost_dev, status_dev,
                             stats_dev, stream, 1);
}

template <class C>
static int launch_replan(pmp_grid *g, SsspArgs &A, const int32_t *cells_dev, const uint8_t *blocked_dev, u32 n,
                         u32 *ctl, u32 *queue, u32 qcap, cudaStream_t s) {
    const unsigned blocks = (n * 32 + 255) / 256;
    k_replan_seed<C><<<blocks > 1184 ? 1184 : blocks, 256, 0, s>>>(A, cells_dev, blocked_dev, n, ctl, queue, qcap);
    PMP_CHECK_CUDA(cudaGetLastError());
    static int n_sm = 0;
    if (!n_sm) {
        int dev = 0;
        PMP_CHECK_CUDA(cudaGetDevice(&dev));
        PMP_CHECK_CUDA(cudaDeviceGetAttribute(&n_sm, cudaDevAttrMultiProcessorCount, dev));
    }
    void *args[] = {(void *)&A, (void *)&g->mot, (void *)&ctl, (void *)&queue, (void *)&qcap};
    PMP_CHECK_CUDA(cudaLaunchCooperativeKernel((void *)k_replan_invalidate<C>, dim3((unsigned)(2 * n_sm)), dim3(256), args, 0, s));
    return 0;
}

extern "C" int pmp_replan_update(pmp_grid *g, const pmp_sssp_params *P, const int32_t *start_dev,
                                 const int32_t *goal_dev, void *g_dev, void *ws_dev, int64_t ws_bytes,
                                 const int32_t *cells_dev, const uint8_t *blocked_dev, int64_t n_changes,
                                 uint32_t *scratch_dev, int64_t scratch_words,
                                 double *gcost_dev, int32_t *status_dev, uint64_t *stats_dev, void *stream) {
    if (g && pmp_grid_upload_fence((pmp_grid *)g, (cudaStream_t)stream)) return -2; // a streamed occupancy upload must have landed
    WsLayout L;
    if (replan_check(g, P, start_dev, goal_dev, g_dev, ws_dev, ws_bytes, L)) return -1;
    if (n_changes < 0 || n_changes > 0x7FFFFFF || (n_changes && (!cells_dev || !blocked_dev))) {
        pmp_set_error("pmp_replan_update: bad change list");
        return -1;
    }
    if (!scratch_dev || scratch_words < PMP_REPLAN_CTL + 2 * 64) {
        pmp_set_error("pmp_replan_update: scratch_dev needs at least %d words", PMP_REPLAN_CTL + 2 * 64);
        return -1;
    }
    cudaStream_t s = (cudaStream_t)stream;
    SsspArgs A;
    fill_args(g, P, 1, start_dev, goal_dev, g_dev, ws_dev, L, A, 1);
    u32 *ctl = scratch_dev, *queue = scratch_dev + PMP_REPLAN_CTL;
    int64_t cap64 = (scratch_words - PMP_REPLAN_CTL) / 2;
    const u32 qcap = cap64 > 0x7FFFFFFF ? 0x7FFFFFFFu : (u32)cap64;
    PMP_CHECK_CUDA(cudaMemsetAsync(ctl, 0, PMP_REPLAN_CTL * 4, s));
    PMP_CHECK_CUDA(cudaMemsetAsync(A.stats, 0, 16 * 8, s));
    if (n_changes) {
        const u32 n = (u32)n_changes;
        const unsigned blocks = (n + 255) / 256;
        k_replan_bits<<<blocks > 1184 ? 1184 : blocks, 256, 0, s>>>(g->d, g->bits, cells_dev, blocked_dev, n);
        PMP_CHECK_CUDA(cudaGetLastError());
        int rc;
        if (g->d.ndim == 2)
            rc = P->weight == PMP_W_F64 ? launch_replan<C2F>(g, A, cells_dev, blocked_dev, n, ctl, queue, qcap, s)
                                        : launch_replan<C2U>(g, A, cells_dev, blocked_dev, n, ctl, queue, qcap, s);
        else
            rc = P->weight == PMP_W_F64 ? launch_replan<C3F>(g, A, cells_dev, blocked_dev, n, ctl, queue, qcap, s)
                                        : launch_replan<C3U>(g, A, cells_dev, blocked_dev, n, ctl, queue, qcap, s);
        if (rc) return rc;
    }
    k_replan_arm<<<1, 1, 0, s>>>(A, ctl);
    PMP_CHECK_CUDA(cudaGetLastError());
    const int rc = dispatch_sssp(g, P->weight, A, gcost_dev, status_dev, s, 2 | 4);
    if (rc) return rc;
    if (stats_dev) {
        PMP_CHECK_CUDA(cudaMemcpyAsync(stats_dev, A.stats, 16 * 8, cudaMemcpyDeviceToDevice, s));
    }
    return 0;
}
