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

// Warp-cooperative form of the search for micro-batches: one warp per new event, the (2r+1)^2 spiral cells spread over the
// lanes (cell s = 32 * round + lane), so the 81 dependent ring look-ups of the serial walk become three rounds of 32
// independent ones -- a micro-batch of 1..10^4 events is latency bound, not throughput bound (for the 6.4 M-event
// reset-mode build the same layout was measured 2.9x SLOWER than thread-per-event, whose neighbouring threads coalesce;
// profiles/r02_summary.md).  Per round: every lane walks its pixel's ring newest-first counting the entries that pass
// the deleted-node stop (j < min_index: break, ev_graph.cu:61), the causal filter and the time gate (continue, :69); a
// prefix sum over the lanes (= spiral order) assigns the output slots; lanes write until slot K.
__device__ __forceinline__ void st_search_one(int e, int lane, const int32_t* __restrict__ batch,
                                              const int32_t* __restrict__ pos, int base_index, int min_index, int B, int H,
                                              int W, int radius, float delta_t, int Q, int K, const StreamView& S,
                                              const int8_t* __restrict__ spiral, int32_t* __restrict__ out,
                                              uint32_t* __restrict__ count_out) {
    const int side = 2 * radius + 1, nsp = side * side;
    const int rounds = (nsp + 31) >> 5;
    const unsigned lt = (1u << lane) - 1u;
    const int i = base_index + e;
    const int x = pos[3 * (int64_t)e], y = pos[3 * (int64_t)e + 1], ti = pos[3 * (int64_t)e + 2];
    const int b = min(max(batch[e], 0), B - 1);
    if (lane == 0) out[0] = i - min_index;
    int total = 1;
    for (int rd = 0; rd < rounds && total < K; rd++) {
        const int s = rd * 32 + lane;
        int cnt = 0, tot = 0, depth = 0, v0 = 0, v1 = 0;
        int64_t c2 = 0;
        if (s < nsp) {
            const int xn = x + spiral[s], yn = y + spiral[nsp + s];
            if (xn >= 0 && yn >= 0 && xn < W && yn < H) {
                c2 = xn + (int64_t)W * (yn + (int64_t)H * b);
                tot = S.count[c2];
                depth = min(tot, Q);
                for (int q = 0; q < depth; q++) {             // q = 0 is the newest entry
                    const int slot = (int)(((int64_t)tot - 1 - q) % Q);
                    const int j = S.ring[c2 * Q + slot];
                    if (j < min_index) break;                 // deleted node: stop this pixel
                    if (i > j) {
                        if ((float)(ti - S.tsr[c2 * Q + slot]) > delta_t) continue;
                        if (cnt == 0) v0 = j; else if (cnt == 1) v1 = j;
                        if (++cnt >= K - 1) break;
                    }
                }
            }
        }
        const unsigned m1 = __ballot_sync(0xffffffffu, cnt > 0);
        if (m1 == 0u) continue;
        int incl = cnt;
#pragma unroll
        for (int d = 1; d < 32; d <<= 1) {
            const int v = __shfl_up_sync(0xffffffffu, incl, d);
            if (lane >= d) incl += v;
        }
        int slot_out = total + incl - cnt;
        const int sum = __shfl_sync(0xffffffffu, incl, 31);
        if (cnt > 0 && slot_out < K) {
            if (cnt <= 2) {
                out[slot_out] = v0 - min_index;
                if (cnt == 2 && slot_out + 1 < K) out[slot_out + 1] = v1 - min_index;
            } else {
                for (int q = 0; q < depth && slot_out < K; q++) {
                    const int slot = (int)(((int64_t)tot - 1 - q) % Q);
                    const int j = S.ring[c2 * Q + slot];
                    if (j < min_index) break;
                    if (i > j) {
                        if ((float)(ti - S.tsr[c2 * Q + slot]) > delta_t) continue;
                        out[slot_out++] = j - min_index;
                    }
                }
            }
        }
        total += sum;
    }
    if (lane == 0) *count_out = (uint32_t)min(total, K);
}

__device__ __forceinline__ void st_spiral_table(int radius, int8_t* spiral) {
    const int nsp = (2 * radius + 1) * (2 * radius + 1);
    int sx = 0, sy = 0, layer = 1, leg = 0;
    for (int s = 0; s < nsp; s++) {
        spiral[s] = (int8_t)sx;
        spiral[nsp + s] = (int8_t)sy;
        switch (leg) {
        case 0: ++sx; if (sx == layer) ++leg; break;
        case 1: ++sy; if (sy == layer) ++leg; break;
        case 2: --sx; if (-sx == layer) ++leg; break;
        default: --sy; if (-sy == layer) { leg = 0; ++layer; } break;
        }
    }
}

__global__ void __launch_bounds__(256) st_search_warp_kernel(const int32_t* __restrict__ batch, const int32_t* __restrict__ pos,
                                                             int64_t n, int base_index, int min_index, int B, int H, int W,
                                                             int radius, float delta_t, int Q, int K, StreamView S,
                                                             int32_t* __restrict__ nbr, uint32_t* __restrict__ count) {
    extern __shared__ int8_t st_spiral[];
    if (threadIdx.x == 0) st_spiral_table(radius, st_spiral);
    __syncthreads();
    const int64_t e = (int64_t)blockIdx.x * 8 + (threadIdx.x >> 5);
    if (e >= n) return;
    st_search_one((int)e, threadIdx.x & 31, batch, pos, base_index, min_index, B, H, W, radius, delta_t, Q, K, S, st_spiral,
                  nbr + e * K, count + e);
}

// The whole append for a micro-batch of at most ST_SMALL_MAX events in ONE launch of one CTA (a micro-batch is pure
// latency: the multi-kernel path below is ~15 launches): keys -> stable sort in shared memory (bitonic on key << 32 |
// position) -> ring insert + counter bump -> warp-cooperative search -> block scan -> CSR / edge_index + edge count.
#define ST_SMALL_MAX 1024
#define ST_SMALL_THREADS 1024
#define ST_SMALL_ONE_CTA 96               // beyond this a single CTA's 32 warps serialise the searches: multi-kernel path

__global__ void __launch_bounds__(ST_SMALL_THREADS) st_small_kernel(const int32_t* __restrict__ batch, const int32_t* __restrict__ pos,
                                                                    int n, int base_index, int min_index, int B, int H, int W,
                                                                    int radius, float delta_t, int Q, int K, int do_search,
                                                                    StreamView S, int32_t* __restrict__ nbr_ws,
                                                                    int32_t* __restrict__ rowptr, int32_t* __restrict__ src,
                                                                    int32_t* __restrict__ num_edges, int64_t* __restrict__ edge_index,
                                                                    int64_t stride, uint32_t* __restrict__ status) {
    __shared__ unsigned long long skv[ST_SMALL_MAX];          // key << 32 | position in the micro-batch
    __shared__ uint32_t scnt[ST_SMALL_MAX + 1];
    __shared__ uint32_t swarp[32];
    extern __shared__ int8_t st_spiral[];
    const int tid = threadIdx.x;
    int np2 = 1;
    while (np2 < n) np2 <<= 1;
    if (tid == 0) st_spiral_table(radius, st_spiral);
    for (int t = tid; t < np2; t += ST_SMALL_THREADS) {
        unsigned long long kv = ~0ull;
        if (t < n) {
            int b = n == 1 ? 0 : batch[t], x = pos[3 * t], y = pos[3 * t + 1];   // single event: b forced to 0 (ev_graph.cu:152)
            if (x < 0 || y < 0 || x >= W || y >= H || b < 0 || b >= B) {
                atomicOr(status, EVAD_STATUS_EVENT_OOB);
                x = min(max(x, 0), W - 1); y = min(max(y, 0), H - 1); b = min(max(b, 0), B - 1);
            }
            kv = ((unsigned long long)(uint32_t)(x + W * (y + H * b)) << 32) | (uint32_t)t;
        }
        skv[t] = kv;
    }
    __syncthreads();
    for (int k = 2; k <= np2; k <<= 1)                        // bitonic sort; unique keys, so the result is the stable order
        for (int j = k >> 1; j > 0; j >>= 1) {
            for (int t = tid; t < np2; t += ST_SMALL_THREADS) {
                const int l = t ^ j;
                if (l > t) {
                    const unsigned long long a = skv[t], c = skv[l];
                    const bool up = (t & k) == 0;
                    if ((a > c) == up) { skv[t] = c; skv[l] = a; }
                }
            }
            __syncthreads();
        }
    // ring insert: position p of a run of c events at one pixel, r-th of the run (only the newest Q of a run survive)
    for (int p = tid; p < n; p += ST_SMALL_THREADS) {
        const uint32_t k = (uint32_t)(skv[p] >> 32), i = (uint32_t)skv[p];
        int rs = p, re = p + 1;
        while (rs > 0 && (uint32_t)(skv[rs - 1] >> 32) == k) rs--;
        while (re < n && (uint32_t)(skv[re] >> 32) == k) re++;
        const int c = re - rs, r = p - rs;
        const int cnt = S.count[k];
        if (r >= c - Q) {
            const int slot = (int)(((int64_t)cnt + r) % Q);
            S.ring[(int64_t)k * Q + slot] = base_index + (int)i;
            S.tsr[(int64_t)k * Q + slot] = pos[3 * (int64_t)i + 2];
        }
        scnt[p] = (r == c - 1) ? (uint32_t)(cnt + c) : 0u;    // the last event of a run publishes the new counter
    }
    __syncthreads();
    for (int p = tid; p < n; p += ST_SMALL_THREADS)
        if (scnt[p]) S.count[(uint32_t)(skv[p] >> 32)] = (int32_t)scnt[p];
    __threadfence_block();
    __syncthreads();
    if (!do_search) return;
    // search: one warp per new event, in arrival order
    for (int e = tid >> 5; e < n; e += ST_SMALL_THREADS / 32)
        st_search_one(e, tid & 31, batch, pos, base_index, min_index, B, H, W, radius, delta_t, Q, K, S, st_spiral,
                      nbr_ws + (int64_t)e * K, scnt + e);
    __syncthreads();
    // exclusive scan of the row lengths (n <= 1024: one element per thread)
    const uint32_t mine = tid < n ? scnt[tid] : 0u;
    uint32_t incl = mine;
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) {
        const uint32_t v = __shfl_up_sync(0xffffffffu, incl, d);
        if ((tid & 31) >= d) incl += v;
    }
    if ((tid & 31) == 31) swarp[tid >> 5] = incl;
    __syncthreads();
    if (tid < 32) {
        uint32_t w = swarp[tid], wi = w;
#pragma unroll
        for (int d = 1; d < 32; d <<= 1) {
            const uint32_t v = __shfl_up_sync(0xffffffffu, wi, d);
            if (tid >= d) wi += v;
        }
        swarp[tid] = wi - w;
    }
    __syncthreads();
    const uint32_t r0 = swarp[tid >> 5] + incl - mine;
    if (tid < n) {
        rowptr[tid] = (int32_t)r0;
        if (tid == n - 1) { rowptr[n] = (int32_t)(r0 + mine); *num_edges = (int32_t)(r0 + mine); }
        const int dst = base_index - min_index + tid;
        for (uint32_t k = 0; k < mine; k++) {
            const int32_t j = nbr_ws[(int64_t)tid * K + k];
            src[r0 + k] = j;
            if (edge_index) { edge_index[r0 + k] = j; edge_index[stride + r0 + k] = dst; }
        }
    }
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
    const int nsp = (2 * radius + 1) * (2 * radius + 1);
    if (n <= ST_SMALL_ONE_CTA && radius <= 64) {              // micro-batch: the whole append in one launch
        st_small_kernel<<<1, ST_SMALL_THREADS, 2 * nsp, st>>>(batch, pos, (int)n, base_index, min_index, B, H, W, radius,
                                                              (float)delta_t_us, Q, K, do_search, S, w.nbr, rowptr, src,
                                                              num_edges, edge_index, edge_index_stride, status);
        EVAD_CHECK_LAUNCH();
        return EVAD_OK;
    }
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
    if (n <= 200000 && radius <= 64)                          // latency-bound sizes: lanes over cells
        st_search_warp_kernel<<<evad_cdiv(n, 8), 256, 2 * nsp, st>>>(batch, pos, n, base_index, min_index, B, H, W, radius,
                                                                     (float)delta_t_us, Q, K, S, w.nbr, w.count);
    else
        st_search_kernel<<<evad_cdiv(n, 128), 128, 0, st>>>(batch, pos, n, base_index, min_index, B, H, W, radius,
                                                           (float)delta_t_us, Q, K, S, w.nbr, w.count);
    EVAD_CHECK_LAUNCH();
    EVAD_TRY(evad_exclusive_scan_u32(w.count, w.rowptr_u, n, w.rowptr_u + n, w.scan_tmp, st));
    st_compact_kernel<<<evad_cdiv(n * K, 256), 256, 0, st>>>(w.nbr, w.rowptr_u, n, K, base_index - min_index, rowptr, src,
                                                           num_edges, edge_index, edge_index_stride);
    EVAD_CHECK_LAUNCH();
    return EVAD_OK;
}

// ------------------------------------------------------------------------------------------------ rows of appended events
// Incremental layer updates (reference: the asynchronous convolution recomputes new nodes from their
// in-edges, src/dagr/asynchronous/conv.py:110-127,183-193): the CSR rows evad_stream_append returns for the n new events
// are written into the GLOBAL fixed-stride row arrays nbr[N][K] / count[N] / nbr_cls[N][K] at node ids base_index + p,
// the layout evad_spline_conv_l1_h16 and evad_voxel_pool read (ell_stride = K).  The class byte of an edge is its pixel
// offset (dx + r) * (2r + 1) + (dy + r), as evad_graph_build_rows writes it; unused slots hold the node itself and 255.
__global__ void __launch_bounds__(256) st_rows_kernel(const int32_t* __restrict__ rowptr, const int32_t* __restrict__ src,
                                                      int64_t n, int base_index, int min_index, const int32_t* __restrict__ pos,
                                                      int radius, int K, int32_t* __restrict__ nbr, int32_t* __restrict__ count,
                                                      uint8_t* __restrict__ nbr_cls, uint32_t* __restrict__ status) {
    const int64_t t = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    const int64_t p = t / K;
    const int s = (int)(t - p * K);
    if (p >= n) return;
    const int i = base_index + (int)p;
    const int r0 = rowptr[p], deg = min(rowptr[p + 1] - r0, K);
    if (s == 0) count[i] = deg;
    int j = i;
    uint32_t c = 255u;
    if (s < deg) {
        j = src[r0 + s] + min_index;
        const int dx = pos[3 * (int64_t)j] - pos[3 * (int64_t)i], dy = pos[3 * (int64_t)j + 1] - pos[3 * (int64_t)i + 1];
        if (abs(dx) > radius || abs(dy) > radius) {
            atomicOr(status, EVAD_STATUS_LUT_OOB);
        } else {
            c = (uint32_t)((dx + radius) * (2 * radius + 1) + (dy + radius));
        }
    }
    nbr[(int64_t)i * K + s] = j;
    nbr_cls[(int64_t)i * K + s] = (uint8_t)c;
}

extern "C" int evad_stream_rows(const int32_t* rowptr, const int32_t* src, int64_t n_new, int base_index, int min_index,
                                const int32_t* pos_int, int radius, int max_neighbors, int32_t* nbr, int32_t* count,
                                uint8_t* nbr_cls, uint32_t* status, void* stream) {
    if (!rowptr || !src || !pos_int || !nbr || !count || !nbr_cls || !status || n_new < 0 || max_neighbors < 1)
        return EVAD_ERR_ARG;
    if ((2 * radius + 1) * (2 * radius + 1) > 255) return EVAD_ERR_UNSUPPORTED;
    if (n_new == 0) return EVAD_OK;
    st_rows_kernel<<<evad_cdiv(n_new * max_neighbors, 256), 256, 0, (cudaStream_t)stream>>>(
        rowptr, src, n_new, base_index, min_index, pos_int, radius, max_neighbors, nbr, count, nbr_cls, status);
    EVAD_CHECK_LAUNCH();
    return EVAD_OK;
}
