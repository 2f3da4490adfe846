// bestfirst_host.cpp -- TEST INFRASTRUCTURE.  Compiles csrc/pmp_bestfirst_core.h (the sequential
// part of the pmp_bestfirst kernel) with g++ and drives it the way k_bestfirst does -- bf_pop on
// "lane 0", bf_eval once per motion against the CLOSED table as it stood before any push, pushes
// in motion order -- so the heap / CLOSED / Theta* logic of the kernel can be checked on a machine
// without a GPU against outputs of the unmodified reference (tests/test_bestfirst_core.py).
// Nothing under python_motion_planning_b200/ loads this; the product path is the CUDA kernel.
#include <stdlib.h>
#include <string.h>

#include <vector>

#include "../python_motion_planning_b200/csrc/pmp_jps_core.h"

typedef int (*los_fn_t)(int ndim, const int32_t *dims, const uint8_t *occ, int64_t n, const int32_t *a,
                        const int32_t *b, uint8_t *out);

struct HostEnv {
    int ndim;
    int32_t dims[3];
    const uint8_t *occ; // (z*Y + y)*X + x
    int nk;
    const int32_t *mot;
    los_fn_t los_fn;

    bool blocked(int x, int y, int z) const {
        if (x < 0 || y < 0 || z < 0 || x >= dims[0] || y >= dims[1] || z >= dims[2]) return true;
        return occ[((int64_t)z * dims[1] + y) * dims[0] + x] != 0;
    }
    // the edge rule of csrc/pmp_common.cuh:edge_legal (graph_search.py:61-87, graph_search_3d.py:66-107)
    bool legal(int m, int x, int y, int z) const {
        const int dx = mot[3 * m], dy = mot[3 * m + 1], dz = mot[3 * m + 2];
        if (blocked(x, y, z) || blocked(x + dx, y + dy, z + dz)) return false;
        if ((dx != 0) + (dy != 0) + (dz != 0) >= 2) {
            if (dx && blocked(x + dx, y, z)) return false;
            if (dy && blocked(x, y + dy, z)) return false;
            if (dz && blocked(x, y, z + dz)) return false;
        }
        return true;
    }
    bool los(int ax, int ay, int az, int bx, int by, int bz) const {
        const int32_t a[3] = {ax, ay, az}, b[3] = {bx, by, bz};
        uint8_t out = 0;
        los_fn(ndim, dims, occ, 1, a, b, &out);
        return out != 0;
    }
    bool inbox(int x, int y, int z) const {
        return x >= 0 && y >= 0 && z >= 0 && x < dims[0] && y < dims[1] && z < dims[2];
    }
    bool toggle(int x, int y, int z) const { // DStarLite.OnPress edits the obstacle set mid-walk; true = freed
        uint8_t *p = const_cast<uint8_t *>(occ) + ((int64_t)z * dims[1] + y) * dims[0] + x;
        const bool was = *p != 0;
        *p = was ? 0 : 1;
        return was;
    }
    bool obst(int x, int y, int z) const { return inbox(x, y, z) && blocked(x, y, z); }
    bool legal_dir(int dx, int dy, int dz, int x, int y, int z) const {
        if (blocked(x, y, z) || blocked(x + dx, y + dy, z + dz)) return false;
        if ((dx != 0) + (dy != 0) + (dz != 0) >= 2) {
            if (dx && blocked(x + dx, y, z)) return false;
            if (dy && blocked(x, y + dy, z)) return false;
            if (dz && blocked(x, y, z + dz)) return false;
        }
        return true;
    }
    int k() const { return nk; }
    void offset(int m, int &dx, int &dy, int &dz) const { dx = mot[3 * m]; dy = mot[3 * m + 1]; dz = mot[3 * m + 2]; }
    double len(int m) const {
        const int dx = mot[3 * m], dy = mot[3 * m + 1], dz = mot[3 * m + 2];
        return sqrt((double)(dx * dx + dy * dy + dz * dz));
    }
};

template <int FAMILY>
static void run(BfState &S, const HostEnv &env, const double *w, const int32_t *start) {
    bf_start<FAMILY>(S, start[0], start[1], start[2]);
    for (;;) {
        BfRec node;
        const int action = bf_pop<FAMILY>(S, env, node);
        if (action == BF_DONE) break;
        if (action == BF_SKIP) continue;
        BfRec cand[32];
        bool push[32];
        for (int lane = 0; lane < env.nk; ++lane)
            push[lane] = S.algo == BF_JPS
                ? (FAMILY == 2 ? jps2d_eval(S, env, node, lane, w[lane], cand[lane]) : jps3d_eval(S, env, node, lane, cand[lane]))
                : bf_eval<FAMILY>(S, env, node, lane, w[lane], cand[lane]);
        for (int m = 0; m < env.nk; ++m) {
            if (!push[m]) continue;
            if (bf_push<FAMILY>(S, cand[m])) break;
        }
        if (FAMILY == 2) bf_close(S, node);
        if (S.status != BF_RUNNING) break;
    }
}

// Same outputs as pmp_oracle_plan (oracle/pmp_oracle.h); motion_w = what one step adds to g
// (motion.g for ndim 2, the offset's length for ndim 3).  Returns 0, -2 on capacity overflow.
extern "C" int bf_host_plan(int ndim, const int32_t *dims, const uint8_t *occ, int n_motions,
                            const int32_t *motions, const double *motion_w, int algo, int heuristic,
                            const int32_t *start, const int32_t *goal, void *los_fn, double *cost_out,
                            int32_t *path_out, int64_t path_cap, int64_t *path_len,
                            int64_t *expand_cell, double *expand_g, double *expand_h,
                            int64_t *expand_parent, int64_t expand_cap, int64_t *n_expand,
                            int32_t *reached) {
    if (n_motions > 32) return -1;
    HostEnv env;
    env.ndim = ndim;
    env.dims[0] = dims[0]; env.dims[1] = dims[1]; env.dims[2] = ndim == 3 ? dims[2] : 1;
    env.occ = occ;
    env.nk = n_motions;
    env.mot = motions;
    env.los_fn = (los_fn_t)los_fn;
    const int64_t cells = (int64_t)env.dims[0] * env.dims[1] * env.dims[2];
    std::vector<BfRec> heap((size_t)cells * n_motions + 16);
    std::vector<double> cg(cells), ch(cells);
    std::vector<uint32_t> cpar(cells, BF_NONE), order(cells);
    BfState S;
    S.heap = heap.data(); S.heap_cap = (int64_t)heap.size();
    S.cg = cg.data(); S.ch = ch.data(); S.cpar = cpar.data();
    S.order = order.data(); S.order_cap = cells;
    S.algo = algo; S.heuristic = heuristic;
    S.PX = env.dims[0]; S.PY = env.dims[1];
    S.gx = goal[0]; S.gy = goal[1]; S.gz = ndim == 3 ? goal[2] : 0;
    const int32_t s3[3] = {start[0], start[1], ndim == 3 ? start[2] : 0};
    S.start_cell = bf_cell(S, s3[0], s3[1], s3[2]);
    S.goal_cell = bf_cell(S, S.gx, S.gy, S.gz);
    if (ndim == 2) run<2>(S, env, motion_w, s3);
    else run<3>(S, env, motion_w, s3);
    if (S.status == BF_OVERFLOW) return -2;
    *reached = S.status == BF_REACHED;
    *n_expand = S.n_order;
    if (S.n_order > expand_cap) return -2;
    for (int64_t i = 0; i < S.n_order; ++i) {
        const uint32_t c = S.order[i];
        expand_cell[i] = c; expand_g[i] = cg[c]; expand_h[i] = ch[c]; expand_parent[i] = cpar[c];
    }
    *path_len = 0;
    *cost_out = INFINITY;
    if (S.status == BF_REACHED) {
        std::vector<uint32_t> cellsv((size_t)path_cap);
        int32_t n = 0;
        bf_extract(S, cellsv.data(), path_cap, &n, cost_out);
        if (n < 0) return -2;
        for (int32_t i = 0; i < n; ++i) { // goal first
            int x, y, z;
            bf_coords(S, cellsv[i], x, y, z);
            path_out[3 * i] = x; path_out[3 * i + 1] = y; path_out[3 * i + 2] = z;
        }
        *path_len = n;
    }
    return 0;
}

// ---- LPA* / D* sessions (csrc/pmp_incr_core.h).  The occupancy array is referenced, not copied:
// the test edits it between calls like the planners edit env.obstacles. ----
#include "../python_motion_planning_b200/csrc/pmp_incr_core.h"

struct IncrHost {
    HostEnv env;
    int kind; // 0 LPA*, 1 D*, 2 D* Lite
    std::vector<double> a, b, key;
    std::vector<uint32_t> parent, list, expand;
    std::vector<uint8_t> tag;
    LpaState L;
    DsState D;
    int64_t cells;
};

static void host_geometry(BfState &S, const HostEnv &env, int heuristic, const int32_t *start, const int32_t *goal) {
    memset(&S, 0, sizeof S);
    S.heuristic = heuristic;
    S.PX = env.dims[0]; S.PY = env.dims[1];
    S.gx = goal[0]; S.gy = goal[1]; S.gz = env.ndim == 3 ? goal[2] : 0;
    S.start_cell = bf_cell(S, start[0], start[1], env.ndim == 3 ? start[2] : 0);
    S.goal_cell = bf_cell(S, S.gx, S.gy, S.gz);
}

extern "C" void *incr_host_create(int kind, int ndim, const int32_t *dims, const uint8_t *occ, int n_motions,
                                  const int32_t *motions, int heuristic, const int32_t *start,
                                  const int32_t *goal) {
    IncrHost *H = new IncrHost();
    H->kind = kind;
    H->env.ndim = ndim;
    H->env.dims[0] = dims[0]; H->env.dims[1] = dims[1]; H->env.dims[2] = ndim == 3 ? dims[2] : 1;
    H->env.occ = occ; H->env.nk = n_motions; H->env.mot = motions; H->env.los_fn = nullptr;
    const int64_t cells = H->cells = (int64_t)H->env.dims[0] * H->env.dims[1] * H->env.dims[2];
    H->a.assign(cells, INFINITY); H->b.assign(cells, INFINITY); H->key.assign(2 * cells, 0.0);
    H->parent.assign(cells, BF_NONE); H->tag.assign(cells, DS_NEW);
    H->list.assign(8 * cells + 64, 0); H->expand.assign(64 * cells + 64, 0);
    if (kind == 0 || kind == 2) {
        LpaState &L = H->L;
        L.dlite = kind == 2; L.km = 0.0;
        L.g = H->a.data(); L.rhs = H->b.data(); L.k1 = H->key.data(); L.k2 = H->key.data() + cells;
        L.U = H->list.data(); L.ucap = (int64_t)H->list.size(); L.un = 0;
        L.expand = H->expand.data(); L.expand_cap = (int64_t)H->expand.size(); L.n_expand = 0;
        L.family = ndim; L.status = BF_RUNNING;
        host_geometry(L.S, H->env, heuristic, start, goal);
    } else {
        DsState &D = H->D;
        D.h = H->a.data(); D.k = H->b.data(); D.parent = H->parent.data(); D.t = H->tag.data();
        D.OPEN = H->list.data(); D.ocap = (int64_t)H->list.size(); D.on = 0;
        D.expand = H->expand.data(); D.expand_cap = (int64_t)H->expand.size(); D.n_expand = 0;
        D.family = ndim; D.status = BF_RUNNING;
        host_geometry(D.S, H->env, 0, start, goal);
    }
    return H;
}

extern "C" void incr_host_destroy(void *h) { delete (IncrHost *)h; }

// op 0: plan from scratch; op 1: after an occupancy edit (LPA*: change = {x, y, z, freed}).
// path_out: coordinates in the order the core wrote them (LPA* goal first, D* start first).
extern "C" int incr_host_step(void *h, int op, const int32_t *change, double *cost_out, int32_t *path_out,
                              int64_t path_cap, int64_t *path_len, int64_t *expand_cell, int64_t expand_cap,
                              int64_t *n_expand) {
    IncrHost *H = (IncrHost *)h;
    std::vector<uint32_t> cellsv((size_t)path_cap);
    int32_t n = 0;
    const BfState *S;
    int status;
    if (H->kind == 2) {
        LpaState &L = H->L;
        if (op == 0) {
            lpa_init(L);
            lpa_compute(L, H->env);
            if (L.status != BF_OVERFLOW) dlite_extract(L, H->env, cellsv.data(), path_cap, &n, cost_out);
        } else {
            dlite_press(L, H->env, change[0], change[1], 0, cellsv.data(), path_cap, &n, cost_out);
        }
        status = L.status;
        *n_expand = L.n_expand;
        for (int64_t i = 0; i < L.n_expand && i < expand_cap; ++i) expand_cell[i] = L.expand[i];
        S = &L.S;
    } else if (H->kind == 0) {
        LpaState &L = H->L;
        if (op == 0) lpa_init(L);
        else lpa_apply_change(L, H->env, change[0], change[1], H->env.ndim == 3 ? change[2] : 0, change[3] != 0);
        lpa_compute(L, H->env);
        status = L.status;
        if (status != BF_OVERFLOW) lpa_extract(L, H->env, cellsv.data(), path_cap, &n, cost_out);
        *n_expand = L.n_expand;
        for (int64_t i = 0; i < L.n_expand && i < expand_cap; ++i) expand_cell[i] = L.expand[i];
        S = &L.S;
    } else {
        DsState &D = H->D;
        if (op == 0) {
            ds_init(D);
            ds_plan(D, H->env);
            if (D.status != BF_OVERFLOW) ds_extract(D, H->env, cellsv.data(), path_cap, &n, cost_out);
        } else {
            ds_replan(D, H->env, cellsv.data(), path_cap, &n, cost_out);
        }
        status = D.status;
        *n_expand = D.n_expand;
        for (int64_t i = 0; i < D.n_expand && i < expand_cap; ++i) expand_cell[i] = D.expand[i];
        S = &D.S;
    }
    if (status == BF_OVERFLOW || n < 0) return -2;
    for (int32_t i = 0; i < n; ++i) {
        int x, y, z;
        bf_coords(*S, cellsv[i], x, y, z);
        path_out[3 * i] = x; path_out[3 * i + 1] = y; path_out[3 * i + 2] = z;
    }
    *path_len = n;
    return 0;
}
