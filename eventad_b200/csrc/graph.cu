// eventad_b200/csrc/graph.cu -- event-graph construction (synchronous / reset mode) for sm_100a.
//
// Reference behaviour reproduced bit-exactly (paths relative to /root/reference):
//   src/dagr/graph/ev_graph.cu:15-80   fill_edges_cuda_kernel  (self loop first, spiral over (2r+1)^2 pixels,
//                                      newest-first walk of the per-pixel FIFO, causal filter idx_j < idx_i,
//                                      `continue` on the time gate, cap K including the self loop)
//   src/dagr/graph/ev_graph.cu:169-212 insert_in_queue_cuda_kernel (FIFO keeps the newest Q indices per pixel,
//                                      INCLUDING later events of the same call)
//   src/dagr/graph/utils.py:6-23       host glue (stable sort by pixel key, mask compaction)
//
// B200 design: no [B,Q,H,W] FIFO tensor (39 MB per sample to clear and walk with stride H*W).  Events are bucketed
// per pixel with a stable radix sort (primitives.cu); the bucket of a pixel, read from its tail, IS the FIFO.
// The search runs one thread per event in pixel-sorted order, so the 81 cell lookups of neighbouring threads hit
// the same L1 lines; candidate lists (index, timestamp) are contiguous per pixel row.
#include "common.cuh"

#define G_THREADS 128

__global__ void __launch_bounds__(256) g_keys_kernel(const int32_t* __restrict__ batch, const int32_t* __restrict__ pos,
                                                     int64_t n, int B, int H, int W, uint32_t* __restrict__ keys,
                                                     uint32_t* __restrict__ status) {
    int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    int b = batch[i], x = pos[3 * i], y = pos[3 * i + 1];
    if (x < 0 || y < 0 || x >= W || y >= H || b < 0 || b >= B) {
        atomicOr(status, EVAD_STATUS_EVENT_OOB);
        x = min(max(x, 0), W - 1);
        y = min(max(y, 0), H - 1);
        b = min(max(b, 0), B - 1);
    }
    keys[i] = (uint32_t)(x + W * (y + H * b));
}

// cell_start[c] = first sorted position whose key >= c  (c = 0..ncells), plus sorted-order timestamps.
__global__ void __launch_bounds__(256) g_cells_kernel(const uint32_t* __restrict__ skey, const uint32_t* __restrict__ sidx,
                                                      const int32_t* __restrict__ pos, int64_t n, int64_t ncells,
                                                      int32_t* __restrict__ cell_start, int2* __restrict__ rec,
                                                      int32_t* __restrict__ cell_perm) {
    int64_t t = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (t <= ncells) {
        int64_t lo = 0, hi = n;                               // lower_bound(skey, t)
        while (lo < hi) {
            int64_t mid = (lo + hi) >> 1;
            if ((int64_t)skey[mid] < t) lo = mid + 1; else hi = mid;
        }
        cell_start[t] = (int32_t)lo;
    }
    if (t < n) {
        uint32_t i = sidx[t];
        rec[t] = make_int2((int)i, pos[3 * (int64_t)i + 2]);  // (event index, timestamp): one 64-bit load per candidate
        if (cell_perm) cell_perm[t] = (int32_t)i;
    }
}

// one thread per event, in pixel-sorted order
__global__ void __launch_bounds__(G_THREADS) g_search_kernel(const uint32_t* __restrict__ skey,
                                                             const int2* __restrict__ rec,
                                                             const int32_t* __restrict__ cell_start, int64_t n, int H,
                                                             int W, int radius, float delta_t, int Q, int K,
                                                             int32_t* __restrict__ nbr /*[n][K]*/,
                                                             uint8_t* __restrict__ nbr_cls /*[n][K] or null*/,
                                                             uint32_t* __restrict__ count /*[n]*/) {
    // spiral walk of spiral.h:1-15 (layer starts at 1, leg 0), tabulated once per CTA: the per-thread state machine cost
    // more instructions per cell visit than the visit itself
    extern __shared__ int8_t g_spiral[];                      // [nsp] dx, then [nsp] dy
    const int nsp = (2 * radius + 1) * (2 * radius + 1);
    if (threadIdx.x == 0) {
        int sx = 0, sy = 0, layer = 1, leg = 0;
        for (int s = 0; s < nsp; s++) {
            g_spiral[s] = (int8_t)sx;
            g_spiral[nsp + s] = (int8_t)sy;
            switch (leg) {
            case 0: ++sx; if (sx == layer) ++leg; break;
            case 1: ++sy; if (sy == layer) ++leg; break;
            case 2: --sx; if (-sx == layer) ++leg; break;
            default: --sy; if (-sy == layer) { leg = 0; ++layer; } break;
            }
        }
    }
    __syncthreads();
    int64_t p = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (p >= n) return;
    const int2 me = rec[p];
    const int i = me.x;
    const int c = (int)skey[p];
    const int ti = me.y;
    const int rem = c % (W * H);
    const int y = rem / W, x = rem - y * W;
    int32_t* out = nbr + (int64_t)i * K;
    uint8_t* ocls = nbr_cls ? nbr_cls + (int64_t)i * K : nullptr;   // spiral offset class (dx + r) * (2r + 1) + (dy + r)
    const int side = 2 * radius + 1;
    out[0] = i;                                               // self edge first (ev_graph.cu:43-46)
    if (ocls) ocls[0] = (uint8_t)(radius * side + radius);
    int cnt = 1;
    for (int s = 0; s < nsp && cnt < K; s++) {
        const int sx = g_spiral[s], sy = g_spiral[nsp + s];
        const int xn = x + sx, yn = y + sy;
        if ((unsigned)xn < (unsigned)W && (unsigned)yn < (unsigned)H) {
            const int c2 = c + sy * W + sx;
            const int s0 = cell_start[c2], s1 = cell_start[c2 + 1];
            const int lo = max(s0, s1 - Q);                   // the FIFO holds the newest Q indices of the pixel
            int q = s1 - 1;
            if (s1 - lo > 8) {                                // long bucket: skip the entries that are not older than i
                int a = lo, b = s1;
                while (a < b) {
                    const int m = (a + b) >> 1;
                    if (rec[m].x < i) a = m + 1; else b = m;
                }
                q = a - 1;
            }
            for (; q >= lo; q--) {                            // newest first
                const int2 r = rec[q];
                if (r.x >= i) continue;                       // later events occupy FIFO slots but are skipped
                if ((float)(ti - r.y) > delta_t) continue;    // NOT break (ev_graph.cu:69)
                if (ocls) ocls[cnt] = (uint8_t)((sx + radius) * side + (sy + radius));
                out[cnt++] = r.x;
                if (cnt >= K) break;
            }
        }
    }
    count[i] = (uint32_t)cnt;
}

// scatter the per-event neighbour slots into the compact CSR (+ optional int64 edge_index rows)
__global__ void __launch_bounds__(256) g_compact_kernel(const int32_t* __restrict__ nbr, const uint32_t* __restrict__ rowptr_u,
                                                        int64_t n, int K, int32_t* __restrict__ rowptr,
                                                        int32_t* __restrict__ src, int32_t* __restrict__ num_edges,
                                                        int64_t* __restrict__ edge_index, int64_t stride,
                                                        const uint8_t* __restrict__ nbr_cls,
                                                        uint8_t* __restrict__ edge_cls) {
    int64_t t = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (t >= n * K) return;
    const int64_t i = t / K;
    const int k = (int)(t - i * K);
    const uint32_t r0 = rowptr_u[i], r1 = rowptr_u[i + 1];
    if (k == 0) {
        rowptr[i] = (int32_t)r0;
        if (i == n - 1) { rowptr[n] = (int32_t)r1; *num_edges = (int32_t)r1; }
    }
    if ((uint32_t)k < r1 - r0) {
        const int32_t j = nbr[t];
        src[r0 + k] = j;
        if (edge_index) { edge_index[r0 + k] = j; edge_index[stride + r0 + k] = i; }
        if (edge_cls) edge_cls[r0 + k] = nbr_cls[t];
    }
}

__global__ void g_empty_kernel(int32_t* rowptr, int32_t* num_edges) { rowptr[0] = 0; *num_edges = 0; }

struct GraphWs {
    uint32_t *k0, *v0, *k1, *v1, *count, *rowptr_u, *sort_tmp, *scan_tmp;
    int32_t *nbr, *cell_start;
    int2* rec;
    uint8_t* nbr_cls;
};

static size_t graph_ws_layout(EvadArena& a, int64_t n, int64_t ncells, int K, GraphWs* w) {
    w->k0 = a.take<uint32_t>(n); w->v0 = a.take<uint32_t>(n);
    w->k1 = a.take<uint32_t>(n); w->v1 = a.take<uint32_t>(n);
    w->rec = a.take<int2>(n);
    w->nbr = a.take<int32_t>(n * K);
    w->nbr_cls = a.take<uint8_t>(n * K);
    w->count = a.take<uint32_t>(n + 1);
    w->rowptr_u = a.take<uint32_t>(n + 1);
    w->cell_start = a.take<int32_t>(ncells + 2);
    w->sort_tmp = a.take<uint32_t>(evad_sort_tmp_elems(n));
    w->scan_tmp = a.take<uint32_t>(evad_scan_tmp_elems(n + 1));
    return a.off;
}

extern "C" size_t evad_graph_ws_bytes(int64_t n, int B, int H, int W, int K) {
    EvadArena a(nullptr, 0);
    GraphWs w;
    return graph_ws_layout(a, n > 0 ? n : 1, (int64_t)B * H * W, K, &w) + 256;
}

// keys -> pixel sort -> cell table -> spiral search into fixed-stride rows nbr[n][K] / nbr_cls[n][K] / count[n]
static int graph_rows(const int32_t* batch, const int32_t* pos, int64_t n, int B, int H, int W, int radius, int delta_t_us,
                      int Q, int K, const GraphWs& w, int32_t* cell_start, int32_t* cell_perm, int32_t* nbr,
                      uint8_t* nbr_cls, uint32_t* count, uint32_t* status, cudaStream_t st) {
    const int64_t ncells = (int64_t)B * H * W;
    g_keys_kernel<<<evad_cdiv(n, 256), 256, 0, st>>>(batch, pos, n, B, H, W, w.k0, status);
    EVAD_CHECK_LAUNCH();
    int key_bits = 1;
    while ((1ll << key_bits) < ncells) key_bits++;
    uint32_t *skey, *sidx;
    EVAD_TRY(evad_sort_pairs_u32(w.k0, w.v0, w.k1, w.v1, 1, n, key_bits, w.sort_tmp, &skey, &sidx, st));
    const int64_t m = (ncells + 1 > n) ? ncells + 1 : n;
    g_cells_kernel<<<evad_cdiv(m, 256), 256, 0, st>>>(skey, sidx, pos, n, ncells, cell_start, w.rec, cell_perm);
    EVAD_CHECK_LAUNCH();
    g_search_kernel<<<evad_cdiv(n, G_THREADS), G_THREADS, 2 * (2 * radius + 1) * (2 * radius + 1), st>>>(skey, w.rec, cell_start, n, H, W, radius,
                                                                  (float)delta_t_us, Q, K, nbr, nbr_cls, count);
    EVAD_CHECK_LAUNCH();
    return EVAD_OK;
}

extern "C" int evad_graph_build(const int32_t* batch, const int32_t* pos, int64_t n, int B, int H, int W, int radius,
                                int delta_t_us, int Q, int K, int32_t* rowptr, int32_t* src, int32_t* num_edges,
                                int64_t* edge_index, int64_t edge_index_stride, int32_t* cell_perm,
                                int32_t* cell_start_out, uint8_t* edge_cls, uint32_t* status, void* ws, size_t ws_bytes,
                                void* stream) {
    cudaStream_t st = (cudaStream_t)stream;
    if (n < 0 || B < 1 || H < 1 || W < 1 || radius < 0 || Q < 1 || K < 1 || K > 32) return EVAD_ERR_ARG;
    if (edge_cls && (2 * radius + 1) * (2 * radius + 1) > 256) return EVAD_ERR_UNSUPPORTED;
    if (radius > 64) return EVAD_ERR_UNSUPPORTED;            // spiral table of the search kernel: 2 (2r+1)^2 bytes of smem
    if (!rowptr || !src || !num_edges || !status) return EVAD_ERR_ARG;
    const int64_t ncells = (int64_t)B * H * W;
    if (ncells >= (1ll << 31) || n * K >= (1ll << 31)) return EVAD_ERR_UNSUPPORTED;
    if (n == 0) {
        g_empty_kernel<<<1, 1, 0, st>>>(rowptr, num_edges);
        EVAD_CHECK_LAUNCH();
        return EVAD_OK;
    }
    if (!batch || !pos || !ws) return EVAD_ERR_ARG;
    EvadArena a(ws, ws_bytes);
    GraphWs w;
    graph_ws_layout(a, n, ncells, K, &w);
    if (!a.ok) return EVAD_ERR_WORKSPACE;
    int32_t* cell_start = cell_start_out ? cell_start_out : w.cell_start;
    EVAD_TRY(graph_rows(batch, pos, n, B, H, W, radius, delta_t_us, Q, K, w, cell_start, cell_perm, w.nbr,
                        edge_cls ? w.nbr_cls : nullptr, w.count, status, st));
    EVAD_TRY(evad_exclusive_scan_u32(w.count, w.rowptr_u, n, w.rowptr_u + n, w.scan_tmp, st));
    g_compact_kernel<<<evad_cdiv(n * K, 256), 256, 0, st>>>(w.nbr, w.rowptr_u, n, K, rowptr, src, num_edges,
                                                          edge_index, edge_index_stride, w.nbr_cls, edge_cls);
    EVAD_CHECK_LAUNCH();
    return EVAD_OK;
}

// The same graph as fixed-stride rows, without the scan + compaction into a CSR: row i = nbr[i*K .. i*K + count[i]).
extern "C" int evad_graph_build_rows(const int32_t* batch, const int32_t* pos, int64_t n, int B, int H, int W, int radius,
                                     int delta_t_us, int Q, int K, int32_t* nbr, int32_t* count, uint8_t* nbr_cls,
                                     int32_t* cell_perm, int32_t* cell_start_out, uint32_t* status, void* ws,
                                     size_t ws_bytes, void* stream) {
    cudaStream_t st = (cudaStream_t)stream;
    if (n < 0 || B < 1 || H < 1 || W < 1 || radius < 0 || Q < 1 || K < 1 || K > 32) return EVAD_ERR_ARG;
    if (nbr_cls && (2 * radius + 1) * (2 * radius + 1) > 256) return EVAD_ERR_UNSUPPORTED;
    if (radius > 64) return EVAD_ERR_UNSUPPORTED;
    if (!nbr || !count || !status) return EVAD_ERR_ARG;
    const int64_t ncells = (int64_t)B * H * W;
    if (ncells >= (1ll << 31) || n * K >= (1ll << 31)) return EVAD_ERR_UNSUPPORTED;
    if (n == 0) return EVAD_OK;
    if (!batch || !pos || !ws) return EVAD_ERR_ARG;
    EvadArena a(ws, ws_bytes);
    GraphWs w;
    graph_ws_layout(a, n, ncells, K, &w);
    if (!a.ok) return EVAD_ERR_WORKSPACE;
    int32_t* cell_start = cell_start_out ? cell_start_out : w.cell_start;
    return graph_rows(batch, pos, n, B, H, W, radius, delta_t_us, Q, K, w, cell_start, cell_perm, nbr, nbr_cls,
                      reinterpret_cast<uint32_t*>(count), status, st);
}
