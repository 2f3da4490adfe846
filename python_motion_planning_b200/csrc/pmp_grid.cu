// pmp_grid.cu -- occupancy bitmask + motion table (the device form of Grid / Grid3D).
// Replaces utils/environment/env.py:41-80 and env3d.py:43-103 of the reference (a Python set of
// coordinate tuples and a cKDTree rebuilt on every update()).
#include <math.h>
#include <stdarg.h>
#include <stdio.h>
#include <string.h>

#include "pmp_common.cuh"

static thread_local char g_err[512] = "";

void pmp_set_error(const char *fmt, ...) {
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(g_err, sizeof g_err, fmt, ap);
    va_end(ap);
}

extern "C" const char *pmp_last_error(void) { return g_err; }
extern "C" int pmp_abi_version(void) { return PMP_B200_ABI_VERSION; }

extern "C" int pmp_device_count(void) {
    int n = 0;
    cudaError_t e = cudaGetDeviceCount(&n);
    if (e != cudaSuccess) {
        pmp_set_error("cudaGetDeviceCount: %s", cudaGetErrorString(e));
        cudaGetLastError();
        return -2;
    }
    return n;
}

// ---------------------------------------------------------------------------------------------
// kernels
// ---------------------------------------------------------------------------------------------

// All cells inside the box free, everything in the tile padding blocked.
__global__ void k_bits_reset(GridDev G, u32 *bits) {
    const size_t n_words = (size_t)G.wpr * G.PY * G.PZ;
    for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < n_words;
         i += (size_t)gridDim.x * blockDim.x) {
        const int wx = (int)(i % G.wpr);
        const size_t row = i / G.wpr;
        const int y = (int)(row % G.PY), z = (int)(row / G.PY);
        u32 w;
        if (y >= G.Y || z >= G.Z) {
            w = 0xFFFFFFFFu;
        } else {
            const int x0 = wx * 32;
            if (x0 + 32 <= G.X) w = 0u;
            else if (x0 >= G.X) w = 0xFFFFFFFFu;
            else w = 0xFFFFFFFFu << (G.X - x0);
        }
        bits[i] = w;
    }
}

// obstacle coordinate list -> bits (value 1) or clear bits (value 0); out-of-box ignored
__global__ void k_bits_scatter(GridDev G, u32 *bits, const int32_t *coords, int64_t n, int value) {
    for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < n;
         i += (int64_t)gridDim.x * blockDim.x) {
        const int x = coords[3 * i], y = coords[3 * i + 1], z = G.ndim == 3 ? coords[3 * i + 2] : 0;
        if (x < 0 || y < 0 || z < 0 || x >= G.X || y >= G.Y || z >= G.Z) continue;
        u32 *w = bits + ((size_t)z * G.PY + y) * G.wpr + (x >> 5);
        if (value) atomicOr(w, 1u << (x & 31));
        else atomicAnd(w, ~(1u << (x & 31)));
    }
}

// dense u8 box [(z*Y + y)*X + x] -> bits; one warp packs one 32-cell word with a ballot
__global__ void k_bits_pack_dense(GridDev G, u32 *bits, const uint8_t *occ) {
    const size_t n_words = (size_t)G.wpr * G.PY * G.PZ;
    const int lane = threadIdx.x & 31;
    const size_t warp0 = (blockIdx.x * (size_t)blockDim.x + threadIdx.x) >> 5;
    const size_t n_warps = ((size_t)gridDim.x * blockDim.x) >> 5;
    for (size_t i = warp0; i < n_words; i += n_warps) {
        const int wx = (int)(i % G.wpr);
        const size_t row = i / G.wpr;
        const int y = (int)(row % G.PY), z = (int)(row / G.PY);
        const int x = wx * 32 + lane;
        int blocked = 1;
        if (x < G.X && y < G.Y && z < G.Z) blocked = occ[((size_t)z * G.Y + y) * G.X + x] != 0;
        const u32 w = __ballot_sync(0xFFFFFFFFu, blocked);
        if (lane == 0) bits[i] = w;
    }
}

// dense u8 box in the caller's own index order [(x*Y + y)*Z + z] (arrays indexed [x, y(, z)] as
// the reference's coordinate tuples are) -> bits.  A block of 32 x 8 threads transposes a
// 32(x) x 128(z) patch through shared memory: 4-byte reads coalesced along z, ballots along x.
__global__ void k_bits_pack_xyz(GridDev G, u32 *bits, const uint8_t *occ, int vec) {
    __shared__ u32 tile[32][33]; // [x][z / 4]
    const int tx = threadIdx.x, ty = threadIdx.y;
    const int z0 = blockIdx.x * 128, y = blockIdx.y, wx = blockIdx.z;
#pragma unroll
    for (int k = 0; k < 4; ++k) {
        const int xi = ty + 8 * k, x = wx * 32 + xi, z = z0 + 4 * tx;
        u32 v = 0x01010101u; // outside the box: blocked
        if (x < G.X && z < G.Z) {
            const uint8_t *p = occ + ((size_t)x * G.Y + y) * G.Z + z;
            if (vec && z + 3 < G.Z) {
                v = *(const u32 *)p;
            } else {
                v = 0;
                for (int b = 0; b < 4; ++b) v |= (u32)((z + b < G.Z) ? (p[b] != 0) : 1) << (8 * b);
            }
        }
        tile[xi][tx] = v;
    }
    __syncthreads();
    // warp ty handles z = z0 + ty, ty + 8, ...: lane = x
#pragma unroll
    for (int k = 0; k < 16; ++k) {
        const int zi = ty + 8 * k; // 0..127
        const u32 word = tile[tx][zi >> 2];
        const u32 w = __ballot_sync(0xFFFFFFFFu, ((word >> (8 * (zi & 3))) & 0xFFu) != 0);
        const int z = z0 + zi;
        if (tx == 0 && z < G.Z) bits[((size_t)z * G.PY + y) * G.wpr + wx] = w;
    }
}

// caller-packed occupancy rows [z][y][ceil(X / 32)] (bit i of word w <-> x = 32 w + i, 1 = blocked)
// -> the grid's own rows; everything outside the box (tail bits, padding rows / planes) blocked
__global__ void k_bits_from_packed(GridDev G, u32 *bits, const u32 *rows, int wpr_in) {
    const size_t n_words = (size_t)G.wpr * G.PY * G.PZ;
    for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < n_words;
         i += (size_t)gridDim.x * blockDim.x) {
        const int wx = (int)(i % G.wpr);
        const size_t row = i / G.wpr;
        const int y = (int)(row % G.PY), z = (int)(row / G.PY);
        u32 w = 0xFFFFFFFFu;
        if (y < G.Y && z < G.Z && wx < wpr_in) {
            w = rows[((size_t)z * G.Y + y) * wpr_in + wx];
            const int x0 = wx * 32;
            if (x0 + 32 > G.X) w |= x0 >= G.X ? 0xFFFFFFFFu : 0xFFFFFFFFu << (G.X - x0);
        }
        bits[i] = w;
    }
}

__global__ void k_count_free(GridDev G, const u32 *bits, u64 *out) {
    const size_t n_words = (size_t)G.wpr * G.PY * G.PZ;
    u64 local = 0;
    for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < n_words;
         i += (size_t)gridDim.x * blockDim.x)
        local += 32 - __popc(bits[i]);
    for (int o = 16; o; o >>= 1) local += __shfl_down_sync(0xFFFFFFFFu, local, o);
    if ((threadIdx.x & 31) == 0 && local) atomicAdd(out, local);
}

// ---------------------------------------------------------------------------------------------
// host API
// ---------------------------------------------------------------------------------------------

static int round_up(int v, int m) { return (v + m - 1) / m * m; }

static int install_motions(pmp_grid *g, int k, const int32_t *off, const double *cost) {
    if (k < 1 || k > PMP_MAXM) {
        pmp_set_error("motion count %d outside [1, %d]", k, PMP_MAXM);
        return -1;
    }
    Motions M;
    memset(&M, 0, sizeof M);
    M.k = k;
    M.cut = g->cut;
    double wmin = INFINITY, he = INFINITY, hm = INFINITY;
    bool integer = true;
    for (int m = 0; m < k; ++m) {
        const int dx = off[3 * m], dy = off[3 * m + 1], dz = off[3 * m + 2];
        if (abs(dx) > 1 || abs(dy) > 1 || abs(dz) > 1 || (dx == 0 && dy == 0 && dz == 0) ||
            (g->d.ndim == 2 && dz != 0)) {
            pmp_set_error("motion %d = (%d,%d,%d): offsets must be non-zero and in {-1,0,1}", m, dx, dy, dz);
            return -1;
        }
        if (!(cost[m] > 0.0) || !isfinite(cost[m])) {
            pmp_set_error("motion %d has non-positive cost %g", m, cost[m]);
            return -1;
        }
        M.dx[m] = (int8_t)dx; M.dy[m] = (int8_t)dy; M.dz[m] = (int8_t)dz;
        M.w[m] = cost[m];
        M.wi[m] = (u32)cost[m];
        if ((double)M.wi[m] != cost[m]) integer = false;
        const int nz = (dx != 0) + (dy != 0) + (dz != 0);
        M.len[m] = sqrt((double)(dx * dx + dy * dy + dz * dz));
        auto bit = [](int ax, int ay, int az) { return 1u << ((az + 1) * 9 + (ay + 1) * 3 + (ax + 1)); };
        u32 need = bit(dx, dy, dz);
        if (nz >= 2 && !g->cut) { // face neighbours of the source along every moving axis
            if (dx) need |= bit(dx, 0, 0);
            if (dy) need |= bit(0, dy, 0);
            if (dz) need |= bit(0, 0, dz);
        }
        M.need[m] = need;
        if (cost[m] < wmin) wmin = cost[m];
        const double e = cost[m] / M.len[m], a = cost[m] / (double)nz;
        if (e < he) he = e;
        if (a < hm) hm = a;
    }
    g->mot = M;
    g->wmin = wmin;
    g->he_scale = he;
    g->hm_scale = hm;
    g->integer_costs = integer;
    return 0;
}

extern "C" int pmp_grid_create(int ndim, const int32_t *dims, pmp_grid **out) {
    if (!dims || !out || (ndim != 2 && ndim != 3)) {
        pmp_set_error("pmp_grid_create: bad arguments");
        return -1;
    }
    const int X = dims[0], Y = dims[1], Z = ndim == 3 ? dims[2] : 1;
    if (X < 1 || Y < 1 || Z < 1) {
        pmp_set_error("pmp_grid_create: non-positive dims");
        return -1;
    }
    pmp_grid *g = new pmp_grid();
    memset(g, 0, sizeof *g);
    GridDev &d = g->d;
    d.ndim = ndim;
    d.X = X; d.Y = Y; d.Z = Z;
    if (ndim == 2) { d.TX = 32; d.TY = 32; d.TZ = 1; }
    else { d.TX = 16; d.TY = 8; d.TZ = 8; }
    d.PX = round_up(X, d.TX); d.PY = round_up(Y, d.TY); d.PZ = round_up(Z, d.TZ);
    d.ntx = d.PX / d.TX; d.nty = d.PY / d.TY; d.ntz = d.PZ / d.TZ;
    d.wpr = (d.PX + 31) / 32;
    const u64 nt = (u64)d.ntx * d.nty * d.ntz;
    d.pcells = (u64)d.PX * d.PY * d.PZ;
    if (nt >= 0xFFFFFFFFull || d.pcells >= 0xFFFFFFFFull) {
        pmp_set_error("grid too large: %llu padded cells (limit 2^32-2)", (unsigned long long)d.pcells);
        delete g;
        return -1;
    }
    d.ntiles = (u32)nt;
    if (cudaGetDevice(&g->device) != cudaSuccess) {
        delete g;
        pmp_set_error("no CUDA device: %s", cudaGetErrorString(cudaGetLastError()));
        return -2;
    }
    const size_t n_words = (size_t)d.wpr * d.PY * d.PZ;
    cudaError_t e = cudaMalloc(&g->bits, n_words * sizeof(u32));
    if (e != cudaSuccess) {
        delete g;
        pmp_set_error("cudaMalloc(bitmask %zu B): %s", n_words * 4, cudaGetErrorString(e));
        return -2;
    }
    d.bits = g->bits;
    k_bits_reset<<<(unsigned)((n_words + 255) / 256 > 4096 ? 4096 : (n_words + 255) / 256), 256>>>(d, g->bits);
    e = cudaGetLastError();
    if (e != cudaSuccess) {
        cudaFree(g->bits);
        delete g;
        pmp_set_error("k_bits_reset launch: %s", cudaGetErrorString(e));
        return -2;
    }
    // default motion tables in the reference's order (env.py:52-55, env3d.py:56-70)
    static const int32_t m2[8][3] = {{-1, 0, 0}, {-1, 1, 0}, {0, 1, 0}, {1, 1, 0},
                                     {1, 0, 0},  {1, -1, 0}, {0, -1, 0}, {-1, -1, 0}};
    static const int32_t m3[26][3] = {
        {-1, 0, 0}, {-1, 1, 0}, {0, 1, 0}, {1, 1, 0}, {1, 0, 0}, {1, -1, 0}, {0, -1, 0}, {-1, -1, 0},
        {0, 0, 1}, {0, 0, -1},
        {-1, 0, 1}, {-1, 1, 1}, {0, 1, 1}, {1, 1, 1}, {1, 0, 1}, {1, -1, 1}, {0, -1, 1}, {-1, -1, 1},
        {-1, 0, -1}, {-1, 1, -1}, {0, 1, -1}, {1, 1, -1}, {1, 0, -1}, {1, -1, -1}, {0, -1, -1}, {-1, -1, -1}};
    const int k = ndim == 2 ? 8 : 26;
    const int32_t *off = ndim == 2 ? &m2[0][0] : &m3[0][0];
    double cost[26];
    for (int m = 0; m < k; ++m) {
        const int nz = (off[3 * m] != 0) + (off[3 * m + 1] != 0) + (off[3 * m + 2] != 0);
        cost[m] = sqrt((double)nz);
    }
    if (install_motions(g, k, off, cost)) {
        cudaFree(g->bits);
        delete g;
        return -1;
    }
    *out = g;
    return 0;
}

extern "C" int pmp_grid_set_corner_cutting(pmp_grid *g, int allow) {
    if (!g) { pmp_set_error("pmp_grid_set_corner_cutting: null"); return -1; }
    if (g->cut == (allow != 0)) return 0;
    g->cut = allow != 0;
    int32_t off[3 * PMP_MAXM];
    double cost[PMP_MAXM];
    const int k = g->mot.k;
    for (int m = 0; m < k; ++m) {
        off[3 * m] = g->mot.dx[m]; off[3 * m + 1] = g->mot.dy[m]; off[3 * m + 2] = g->mot.dz[m];
        cost[m] = g->mot.w[m];
    }
    return install_motions(g, k, off, cost);
}

extern "C" int pmp_grid_destroy(pmp_grid *g) {
    if (!g) return 0;
    if (g->ready_dev) cudaFree(g->ready_dev);
    if (g->ones_host) cudaFreeHost(g->ones_host);
    if (g->ev_reset) { cudaEventDestroy(g->ev_reset); cudaEventDestroy(g->ev_uploaded); }
    cudaFree(g->bits);
    delete g;
    return 0;
}

extern "C" int pmp_grid_layout(const pmp_grid *g, int64_t *o) {
    if (!g || !o) { pmp_set_error("pmp_grid_layout: null"); return -1; }
    o[0] = g->d.PX; o[1] = g->d.PY; o[2] = g->d.PZ;
    o[3] = g->d.TX; o[4] = g->d.TY; o[5] = g->d.TZ;
    o[6] = g->d.ntiles; o[7] = g->d.wpr;
    return 0;
}

static unsigned blocks_for(size_t n, int per) {
    size_t b = (n + per - 1) / per;
    if (b < 1) b = 1;
    if (b > 148 * 16) b = 148 * 16;
    return (unsigned)b;
}

extern "C" int pmp_grid_set_cells(pmp_grid *g, const int32_t *coords_dev, int64_t n, void *stream) {
    if (g && pmp_grid_upload_fence((pmp_grid *)g, (cudaStream_t)stream)) return -2; // a streamed occupancy upload must have landed
    if (!g || (n > 0 && !coords_dev) || n < 0) { pmp_set_error("pmp_grid_set_cells: bad arguments"); return -1; }
    cudaStream_t s = (cudaStream_t)stream;
    const size_t n_words = (size_t)g->d.wpr * g->d.PY * g->d.PZ;
    k_bits_reset<<<blocks_for(n_words, 256), 256, 0, s>>>(g->d, g->bits);
    if (n > 0) k_bits_scatter<<<blocks_for((size_t)n, 256), 256, 0, s>>>(g->d, g->bits, coords_dev, n, 1);
    PMP_CHECK_CUDA(cudaGetLastError());
    return 0;
}

extern "C" int pmp_grid_edit_cells(pmp_grid *g, const int32_t *coords_dev, int64_t n, int value, void *stream) {
    if (g && pmp_grid_upload_fence((pmp_grid *)g, (cudaStream_t)stream)) return -2; // a streamed occupancy upload must have landed
    if (!g || (n > 0 && !coords_dev) || n < 0) { pmp_set_error("pmp_grid_edit_cells: bad arguments"); return -1; }
    if (n > 0)
        k_bits_scatter<<<blocks_for((size_t)n, 256), 256, 0, (cudaStream_t)stream>>>(g->d, g->bits, coords_dev, n, value ? 1 : 0);
    PMP_CHECK_CUDA(cudaGetLastError());
    return 0;
}

extern "C" int pmp_grid_set_dense(pmp_grid *g, const uint8_t *occ_dev, void *stream) {
    if (g && pmp_grid_upload_fence((pmp_grid *)g, (cudaStream_t)stream)) return -2; // a streamed occupancy upload must have landed
    if (!g || !occ_dev) { pmp_set_error("pmp_grid_set_dense: bad arguments"); return -1; }
    const size_t n_words = (size_t)g->d.wpr * g->d.PY * g->d.PZ;
    k_bits_pack_dense<<<blocks_for(n_words, 8), 256, 0, (cudaStream_t)stream>>>(g->d, g->bits, occ_dev);
    PMP_CHECK_CUDA(cudaGetLastError());
    return 0;
}

extern "C" int pmp_grid_set_dense_xyz(pmp_grid *g, const uint8_t *occ_dev, void *stream) {
    if (g && pmp_grid_upload_fence((pmp_grid *)g, (cudaStream_t)stream)) return -2; // a streamed occupancy upload must have landed
    if (!g || !occ_dev) { pmp_set_error("pmp_grid_set_dense_xyz: bad arguments"); return -1; }
    if (g->d.Y > 65535 || g->d.wpr > 65535) { pmp_set_error("pmp_grid_set_dense_xyz: grid too wide"); return -1; }
    cudaStream_t s = (cudaStream_t)stream;
    const size_t n_words = (size_t)g->d.wpr * g->d.PY * g->d.PZ;
    k_bits_reset<<<blocks_for(n_words, 256), 256, 0, s>>>(g->d, g->bits); // padding rows / planes: blocked
    dim3 grid((unsigned)((g->d.Z + 127) / 128), (unsigned)g->d.Y, (unsigned)g->d.wpr), block(32, 8);
    const int vec = (g->d.Z % 4 == 0) && (((uintptr_t)occ_dev & 3) == 0); // 4-byte loads stay aligned
    k_bits_pack_xyz<<<grid, block, 0, s>>>(g->d, g->bits, occ_dev, vec);
    PMP_CHECK_CUDA(cudaGetLastError());
    return 0;
}

extern "C" int pmp_grid_set_packed(pmp_grid *g, const uint32_t *rows_dev, void *stream) {
    if (g && pmp_grid_upload_fence((pmp_grid *)g, (cudaStream_t)stream)) return -2; // a streamed occupancy upload must have landed
    if (!g || !rows_dev) { pmp_set_error("pmp_grid_set_packed: bad arguments"); return -1; }
    const size_t n_words = (size_t)g->d.wpr * g->d.PY * g->d.PZ;
    k_bits_from_packed<<<blocks_for(n_words, 256), 256, 0, (cudaStream_t)stream>>>(g->d, g->bits, rows_dev, (g->d.X + 31) / 32);
    PMP_CHECK_CUDA(cudaGetLastError());
    return 0;
}

// ---- streamed upload: the search may start while the occupancy is still arriving -----------------------
int pmp_grid_upload_fence(pmp_grid *g, cudaStream_t stream) {
    if (!g || !g->upload_pending) return 0;
    PMP_CHECK_CUDA(cudaStreamWaitEvent(stream, g->ev_uploaded, 0));
    g->upload_pending = false;
    g->d.ready = nullptr;
    return 0;
}

extern "C" int pmp_grid_set_packed_host_streamed(pmp_grid *g, const uint32_t *rows_host_pinned, int planes_per_chunk,
                                                 void *search_stream, void *copy_stream) {
    if (!g || !rows_host_pinned || !copy_stream || copy_stream == search_stream) {
        pmp_set_error("pmp_grid_set_packed_host_streamed: bad arguments (needs a copy stream of its own)");
        return -1;
    }
    const GridDev &d = g->d;
    if (d.ndim != 3 || d.X % 32 || d.PY != d.Y || d.PZ != d.Z) {
        pmp_set_error("pmp_grid_set_packed_host_streamed: needs X %% 32 == 0, Y %% 8 == 0, Z %% 8 == 0 (the rows must be the "
                      "grid's own layout); use pmp_grid_set_packed");
        return -1;
    }
    int shift = 3;
    while ((1 << (shift + 1)) <= planes_per_chunk && shift < 10) ++shift; // chunks of 2^shift >= 8 planes
    const int cp = 1 << shift, nchunks = (d.Z + cp - 1) / cp;
    if (g->n_ready < nchunks) {
        if (g->ready_dev) cudaFree(g->ready_dev);
        if (g->ones_host) cudaFreeHost(g->ones_host);
        g->ready_dev = nullptr; g->ones_host = nullptr; g->n_ready = 0;
        PMP_CHECK_CUDA(cudaMalloc(&g->ready_dev, (size_t)nchunks * 4));
        PMP_CHECK_CUDA(cudaHostAlloc(&g->ones_host, (size_t)nchunks * 4, cudaHostAllocDefault));
        for (int i = 0; i < nchunks; ++i) g->ones_host[i] = 1u;
        g->n_ready = nchunks;
    }
    if (!g->ev_reset) {
        PMP_CHECK_CUDA(cudaEventCreateWithFlags(&g->ev_reset, cudaEventDisableTiming));
        PMP_CHECK_CUDA(cudaEventCreateWithFlags(&g->ev_uploaded, cudaEventDisableTiming));
    }
    cudaStream_t ss = (cudaStream_t)search_stream, cs = (cudaStream_t)copy_stream;
    // the flags go back to zero in search-stream order (after earlier searches, before the next one); the
    // copies start after that
    PMP_CHECK_CUDA(cudaMemsetAsync(g->ready_dev, 0, (size_t)nchunks * 4, ss));
    PMP_CHECK_CUDA(cudaEventRecord(g->ev_reset, ss));
    PMP_CHECK_CUDA(cudaStreamWaitEvent(cs, g->ev_reset, 0));
    const size_t plane_words = (size_t)d.wpr * d.PY;
    for (int c = 0; c < nchunks; ++c) {
        const int z0 = c * cp, z1 = z0 + cp < d.Z ? z0 + cp : d.Z;
        PMP_CHECK_CUDA(cudaMemcpyAsync(g->bits + (size_t)z0 * plane_words, rows_host_pinned + (size_t)z0 * plane_words,
                                       (size_t)(z1 - z0) * plane_words * 4, cudaMemcpyHostToDevice, cs));
        PMP_CHECK_CUDA(cudaMemcpyAsync(g->ready_dev + c, g->ones_host + c, 4, cudaMemcpyHostToDevice, cs));
    }
    PMP_CHECK_CUDA(cudaEventRecord(g->ev_uploaded, cs));
    g->d.ready = g->ready_dev;
    g->d.ready_shift = shift;
    g->upload_pending = true;
    return 0;
}

extern "C" int pmp_grid_read_bits(const pmp_grid *g, uint32_t *bits_host, int64_t n_words) {
    if (!g || !bits_host) { pmp_set_error("pmp_grid_read_bits: bad arguments"); return -1; }
    if (g->upload_pending) { cudaEventSynchronize(g->ev_uploaded); }
    const int64_t have = (int64_t)g->d.wpr * g->d.PY * g->d.PZ;
    if (n_words != have) { pmp_set_error("pmp_grid_read_bits: expected %lld words", (long long)have); return -1; }
    PMP_CHECK_CUDA(cudaMemcpy(bits_host, g->bits, (size_t)have * 4, cudaMemcpyDeviceToHost));
    return 0;
}

extern "C" int pmp_grid_count_free(const pmp_grid *g, int64_t *n_free) {
    if (!g || !n_free) { pmp_set_error("pmp_grid_count_free: bad arguments"); return -1; }
    if (g->upload_pending) cudaEventSynchronize(g->ev_uploaded);
    u64 *d_out = nullptr;
    PMP_CHECK_CUDA(cudaMalloc(&d_out, 8));
    cudaMemset(d_out, 0, 8);
    const size_t n_words = (size_t)g->d.wpr * g->d.PY * g->d.PZ;
    k_count_free<<<blocks_for(n_words, 256), 256>>>(g->d, g->bits, d_out);
    u64 h = 0;
    cudaError_t e = cudaMemcpy(&h, d_out, 8, cudaMemcpyDeviceToHost);
    cudaFree(d_out);
    if (e != cudaSuccess) { pmp_set_error("pmp_grid_count_free: %s", cudaGetErrorString(e)); return -2; }
    // words past PX inside the last row word are blocked, so popcounts are exact
    *n_free = (int64_t)h;
    return 0;
}

extern "C" int pmp_grid_set_motions(pmp_grid *g, int k, const int32_t *offsets_host, const double *cost_host) {
    if (!g || !offsets_host || !cost_host) { pmp_set_error("pmp_grid_set_motions: bad arguments"); return -1; }
    return install_motions(g, k, offsets_host, cost_host);
}
