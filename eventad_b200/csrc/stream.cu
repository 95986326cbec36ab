// eventad_b200/csrc/stream.cu -- event-graph construction, streaming ("append") mode.
//
// Reference behaviour reproduced bit-exactly (paths relative to /root/reference):
//   src/dagr/model/layers/ev_tgn.py:45-56       EV_TGN.forward(reset=False): no reset, new events appended
//   src/dagr/graph/ev_graph.py:75-103           indices = max_index + arange(n); edges only for the new events
//   src/dagr/graph/ev_graph.cu:169-212,130-166  FIFO shift/insert per active pixel (single-event variant forces b = 0)
//   src/dagr/graph/ev_graph.cu:15-80            search; entries < min_index terminate a pixel's walk (:61)
//
// B200 design: the FIFO is a per-pixel RING [cell][Q] (Q contiguous slots per pixel instead of the reference's
// [B,Q,H,W] planes, so a walk touches one or two 128 B lines per pixel) with a per-pixel insert counter; nothing is
// shifted, nothing but the counters is cleared, and timestamps ride along in a second ring so the search never
// gathers from the growing all_timestamps array.
#include "common.cuh"

struct StreamView {
    int32_t* count;   // [cells]   events ever inserted at the pixel
    int32_t* ring;    // [cells*Q] event index of insert number k at slot k % Q
    int32_t* tsr;     // [cells*Q] its timestamp
};

static StreamView stream_view(void* state, int64_t cells, int Q) {
    StreamView v;
    char* p = (char*)state;
    v.count = (int32_t*)p;
    p += evad_align((size_t)cells * 4);
    v.ring = (int32_t*)p;
    p += evad_align((size_t)cells * Q * 4);
    v.tsr = (int32_t*)p;
    return v;
}

extern "C" size_t evad_stream_state_bytes(int B, int H, int W, int Q) {
    const int64_t cells = (int64_t)B * H * W;
    return evad_align((size_t)cells * 4) + 2 * evad_align((size_t)cells * Q * 4) + 256;
}

extern "C" int evad_stream_reset(void* state, int B, int H, int W, int Q, void* stream) {
    if (!state) return EVAD_ERR_ARG;
    const int64_t cells = (int64_t)B * H * W;
    if (cudaMemsetAsync(state, 0, (size_t)cells * 4, (cudaStream_t)stream) != cudaSuccess) return EVAD_ERR_LAUNCH;
    (void)Q;
    return EVAD_OK;
}

__global__ void __launch_bounds__(256) st_keys_kernel(const int32_t* __restrict__ batch, const int32_t* __restrict__ pos,
                                                      int64_t n, int B, int H, int W, int force_b0,
                                                      uint32_t* __restrict__ keys, uint32_t* __restrict__ status) {
    int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    int b = force_b0 ? 0 : batch[i], x = pos[3 * i], y = pos[3 * i + 1];
    if (x < 0 || y < 0 || x >= W || y >= H || b < 0 || b >= B) {
        atomicOr(status, EVAD_STATUS_EVENT_OOB);
        x = min(max(x, 0), W - 1); y = min(max(y, 0), H - 1); b = min(max(b, 0), B - 1);
    }
    keys[i] = (uint32_t)(x + W * (y + H * b));
}

// one thread per new event in pixel-sorted order: write it into the ring (only the newest Q of a run survive)
__global__ void __launch_bounds__(256) st_insert_kernel(const uint32_t* __restrict__ skey, const uint32_t* __restrict__ sidx,
                                                        const int32_t* __restrict__ pos, int64_t n, int base_index,
                                                        int Q, StreamView S, int32_t* __restrict__ run_len) {
    int64_t p = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (p >= n) return;
    const uint32_t k = skey[p];
    int64_t lo = 0, hi = p;                                   // first position of the run of k
    while (lo < hi) { int64_t m = (lo + hi) >> 1; if (skey[m] < k) lo = m + 1; else hi = m; }
    const int64_t rs = lo;
    lo = p; hi = n;                                           // one past the last position of the run
    while (lo < hi) { int64_t m = (lo + hi) >> 1; if (skey[m] <= k) lo = m + 1; else hi = m; }
    const int c = (int)(lo - rs), r = (int)(p - rs);
    run_len[p] = (r == c - 1) ? c : 0;                        // the last event of a run carries the run length
    if (r < c - Q) return;                                    // would be overwritten by a newer event of this call
    const int cnt = S.count[k];
    const int slot = (int)(((int64_t)cnt + r) % Q);
    const uint32_t i = sidx[p];
    S.ring[(int64_t)k * Q + slot] = base_index + (int)i;
    S.tsr[(int64_t)k * Q + slot] = pos[3 * (int64_t)i + 2];
}

__global__ void __launch_bounds__(256) st_bump_kernel(const uint32_t* __restrict__ skey, const int32_t* __restrict__ run_len,
                                                      int64_t n, StreamView S) {
    int64_t p = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (p >= n) return;
    const int c = run_len[p];
    if (c > 0) S.count[skey[p]] += c;                         // one writer per pixel
}

// ev_graph.cu:15-80 on the ring; one thread per new event
__global__ void __launch_bounds__(128) st_search_kernel(const int32_t* __restrict__ batch, const int32_t* __restrict__ pos,
                                                        int64_t n, int base_index, int min_index, int B, int H, int W,
                                                        int radius, float delta_t, int Q, int K, StreamView S,
                                                        int32_t* __restrict__ nbr, uint32_t* __restrict__ count) {
    int64_t e = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (e >= n) return;
    const int i = base_index + (int)e;
    const int x = pos[3 * e], y = pos[3 * e + 1], ti = pos[3 * e + 2];
    const int b = min(max(batch[e], 0), B - 1);
    int32_t* out = nbr + e * K;
    out[0] = i - min_index;
    int cnt = 1;
    int sx = 0, sy = 0, layer = 1, leg = 0;
    const int nsp = (2 * radius + 1) * (2 * radius + 1);
    for (int s = 0; s < nsp && cnt < K; s++) {
        const int xn = x + sx, yn = y + sy;
        if (xn >= 0 && yn >= 0 && xn < W && yn < H) {
            const int64_t c2 = xn + (int64_t)W * (yn + (int64_t)H * b);
            const int tot = S.count[c2];
            const int depth = min(tot, Q);
            for (int q = 0; q < depth; q++) {                 // q = 0 is the newest entry
                const int slot = (int)(((int64_t)tot - 1 - q) % Q);
                const int j = S.ring[c2 * Q + slot];
                if (j < min_index) break;                     // deleted node: stop this pixel (ev_graph.cu:61)
                if (i > j) {
                    const int dt = ti - S.tsr[c2 * Q + slot];
                    if ((float)dt > delta_t) continue;
                    out[cnt++] = j - min_index;
                    if (cnt >= K) break;
                }
            }
        }
        switch (leg) {
        case 0: ++sx; if (sx == layer) ++leg; break;
        case 1: ++sy; if (sy == layer) ++leg; break;
        case 2: --sx; if (-sx == layer) ++leg; break;
        default: --sy; if (-sy == layer) { leg = 0; ++layer; } break;
        }
    }
    count[e] = (uint32_t)cnt;
}

__global__ void __launch_bounds__(256) st_compact_kernel(const int32_t* __restrict__ nbr, const uint32_t* __restrict__ rowptr_u,
                                                         int64_t n, int K, int dst_base, int32_t* __restrict__ rowptr,
                                                         int32_t* __restrict__ src, int32_t* __restrict__ num_edges,
                                                         int64_t* __restrict__ edge_index, int64_t stride) {
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
        if (edge_index) { edge_index[r0 + k] = j; edge_index[stride + r0 + k] = dst_base + i; }
    }
}

__global__ void st_empty_kernel(int32_t* rowptr, int32_t* num_edges) { rowptr[0] = 0; *num_edges = 0; }

struct StreamWs {
    uint32_t *k0, *v0, *k1, *v1, *count, *rowptr_u, *sort_tmp, *scan_tmp;
    int32_t *nbr, *run_len;
};

static size_t stream_ws_layout(EvadArena& a, int64_t n, int K, StreamWs* w) {
    w->k0 = a.take<uint32_t>(n); w->v0 = a.take<uint32_t>(n);
    w->k1 = a.take<uint32_t>(n); w->v1 = a.take<uint32_t>(n);
    w->run_len = a.take<int32_t>(n);
    w->nbr = a.take<int32_t>(n * K);
    w->count = a.take<uint32_t>(n + 1);
    w->rowptr_u = a.take<uint32_t>(n + 1);
    w->sort_tmp = a.take<uint32_t>(evad_sort_tmp_elems(n));
    w->scan_tmp = a.take<uint32_t>(evad_scan_tmp_elems(n + 1));
    return a.off;
}

extern "C" size_t evad_stream_ws_bytes(int64_t n, int K) {
    EvadArena a(nullptr, 0);
    StreamWs w;
    return stream_ws_layout(a, n > 0 ? n : 1, K, &w) + 256;
}

extern "C" int evad_stream_append(void* state, int B, int H, int W, int Q, const int32_t* batch, const int32_t* pos,
                                  int64_t n, int base_index, int min_index, int radius, int delta_t_us, int K,
                                  int do_search, int32_t* rowptr, int32_t* src, int32_t* num_edges,
                                  int64_t* edge_index, int64_t edge_index_stride, uint32_t* status, void* ws,
                                  size_t ws_bytes, void* stream) {
    cudaStream_t st = (cudaStream_t)stream;
    if (!state || n < 0 || B < 1 || H < 1 || W < 1 || Q < 1 || K < 1 || K > 32 || !status) return EVAD_ERR_ARG;
    if (do_search && (!rowptr || !src || !num_edges)) return EVAD_ERR_ARG;
    const int64_t cells = (int64_t)B * H * W;
    if (cells >= (1ll << 31) || n * K >= (1ll << 31)) return EVAD_ERR_UNSUPPORTED;
    if (n == 0) {
        if (do_search) { st_empty_kernel<<<1, 1, 0, st>>>(rowptr, num_edges); EVAD_CHECK_LAUNCH(); }
        return EVAD_OK;
    }
    if (!batch || !pos || !ws) return EVAD_ERR_ARG;
    EvadArena a(ws, ws_bytes);
    StreamWs w;
    stream_ws_layout(a, n, K, &w);
    if (!a.ok) return EVAD_ERR_WORKSPACE;
    StreamView S = stream_view(state, cells, Q);
    // insert_in_queue_single (one event) ignores `batch` and writes into sample 0 (ev_graph.cu:152)
    st_keys_kernel<<<evad_cdiv(n, 256), 256, 0, st>>>(batch, pos, n, B, H, W, n == 1 ? 1 : 0, w.k0, status);
    EVAD_CHECK_LAUNCH();
    int key_bits = 1;
    while ((1ll << key_bits) < cells) key_bits++;
    uint32_t *skey, *sidx;
    EVAD_TRY(evad_sort_pairs_u32(w.k0, w.v0, w.k1, w.v1, 1, n, key_bits, w.sort_tmp, &skey, &sidx, st));
    st_insert_kernel<<<evad_cdiv(n, 256), 256, 0, st>>>(skey, sidx, pos, n, base_index, Q, S, w.run_len);
    EVAD_CHECK_LAUNCH();
    st_bump_kernel<<<evad_cdiv(n, 256), 256, 0, st>>>(skey, w.run_len, n, S);
    EVAD_CHECK_LAUNCH();
    if (!do_search) return EVAD_OK;
    st_search_kernel<<<evad_cdiv(n, 128), 128, 0, st>>>(batch, pos, n, base_index, min_index, B, H, W, radius,
                                                       (float)delta_t_us, Q, K, S, w.nbr, w.count);
    EVAD_CHECK_LAUNCH();
    EVAD_TRY(evad_exclusive_scan_u32(w.count, w.rowptr_u, n, w.rowptr_u + n, w.scan_tmp, st));
    st_compact_kernel<<<evad_cdiv(n * K, 256), 256, 0, st>>>(w.nbr, w.rowptr_u, n, K, base_index - min_index, rowptr, src,
                                                           num_edges, edge_index, edge_index_stride);
    EVAD_CHECK_LAUNCH();
    return EVAD_OK;
}
