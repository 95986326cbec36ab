/*
 * oracle/evg_oracle.c -- CPU restatement of the reference event-graph builder.
 *
 * TEST INFRASTRUCTURE ONLY.  Nothing in the product path (eventad_b200/) may link, import
 * or call this file; only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline /
 * --impl reference legs use it, as the checker or as the timed CPU baseline.
 *
 * What it restates (all paths relative to /root/reference):
 *   - the per-pixel FIFO of event indices  queue[B][Q][H][W] int32, -1 = empty, slot 0 = newest
 *     (src/dagr/graph/ev_graph.py:45-50 allocation, :52-61 reset)
 *   - insertion of a batch of events into the FIFO
 *     (src/dagr/graph/utils.py:6-18 host glue: key = x + W*y + W*H*b, STABLE sort, run lengths;
 *      src/dagr/graph/ev_graph.cu:169-212 per-active-pixel shift/insert kernel;
 *      src/dagr/graph/ev_graph.cu:130-166 single-event variant that forces b = 0)
 *   - the spiral neighbour search  (src/dagr/graph/ev_graph.cu:15-80, spiral walk spiral.h:1-15)
 *   - compaction of the -1 filled [2, K*N] edge buffer (src/dagr/graph/utils.py:20-23)
 *   - index bookkeeping of AsyncGraph.forward / SlidingWindowGraph.delete_nodes
 *     (src/dagr/graph/ev_graph.py:63-103, :121-136): max_index, min_index, all_timestamps.
 *
 * Parity status: pinned on the GPU box against the UNMODIFIED reference kernel built by
 * oracle/build_ref.py (tests/test_ref_kernel_gpu.py); the reference ships no golden vectors
 * of its own.
 *
 * The GPU kernel's "one thread per event" becomes a loop over events (OpenMP when available:
 * every event writes only its own K slots, every active pixel only its own FIFO column, exactly
 * the independence the reference kernels rely on).
 */
#include <stdint.h>
#include <stdlib.h>
#include <string.h>

typedef struct {
    int W, H, B, K, Q;
    float radius;      /* the reference passes radius and delta_t_us as float (ev_graph.cu:24) */
    float delta_t_us;
    int32_t *queue;    /* [B][Q][H][W] */
    int32_t *all_ts;   /* timestamps of resident events, index j - min_index */
    int64_t n_ts, cap_ts;
    int64_t max_index, min_index;
} evgo_t;

/* ---- the spiral of spiral.h:1-15: layer starts at 1, leg 0 ---- */
typedef struct { unsigned layer, leg; int x, y; } spiral_t;
static void spiral_init(spiral_t *s) { s->layer = 1; s->leg = 0; s->x = 0; s->y = 0; }
static void spiral_next(spiral_t *s) {
    switch (s->leg) {
    case 0: ++s->x; if (s->x == (int)s->layer) ++s->leg; break;
    case 1: ++s->y; if (s->y == (int)s->layer) ++s->leg; break;
    case 2: --s->x; if (-s->x == (int)s->layer) ++s->leg; break;
    case 3: --s->y; if (-s->y == (int)s->layer) { s->leg = 0; ++s->layer; } break;
    }
}

evgo_t *evgo_create(int W, int H, int B, int K, int Q, float radius, float delta_t_us) {
    evgo_t *g = (evgo_t *)calloc(1, sizeof(evgo_t));
    g->W = W; g->H = H; g->B = B; g->K = K; g->Q = Q;
    g->radius = radius; g->delta_t_us = delta_t_us;
    size_t n = (size_t)B * Q * H * W;
    g->queue = (int32_t *)malloc(n * sizeof(int32_t));
    memset(g->queue, 0xff, n * sizeof(int32_t));
    g->cap_ts = 1 << 16;
    g->all_ts = (int32_t *)malloc(g->cap_ts * sizeof(int32_t));
    return g;
}

void evgo_destroy(evgo_t *g) { if (g) { free(g->queue); free(g->all_ts); free(g); } }

/* AsyncGraph.reset (ev_graph.py:52-61): queue.fill_(-1), indices back to 0 */
void evgo_reset(evgo_t *g) {
    size_t n = (size_t)g->B * g->Q * g->H * g->W;
    memset(g->queue, 0xff, n * sizeof(int32_t));
    g->n_ts = 0; g->max_index = 0; g->min_index = 0;
}

int64_t evgo_max_index(const evgo_t *g) { return g->max_index; }
int64_t evgo_min_index(const evgo_t *g) { return g->min_index; }
int64_t evgo_num_timestamps(const evgo_t *g) { return g->n_ts; }
const int32_t *evgo_queue(const evgo_t *g) { return g->queue; }

/* SlidingWindowGraph.delete_nodes (ev_graph.py:121-136), node part only:
 * all_timestamps = all_timestamps[n:], min_index += n.  (Edge renumbering is host-side python.) */
void evgo_delete_nodes(evgo_t *g, int64_t n) {
    if (n > g->n_ts) n = g->n_ts;
    memmove(g->all_ts, g->all_ts + n, (size_t)(g->n_ts - n) * sizeof(int32_t));
    g->n_ts -= n;
    g->min_index += n;
}

/* stable merge sort of (key, idx) pairs by key == torch.sort(stable=True) (graph/utils.py:10) */
typedef struct { int64_t key; int32_t idx; } kv_t;
static void merge_sort(kv_t *a, kv_t *tmp, int64_t n) {
    if (n < 2) return;
    int64_t h = n / 2;
    merge_sort(a, tmp, h); merge_sort(a + h, tmp, n - h);
    int64_t i = 0, j = h, k = 0;
    while (i < h && j < n) tmp[k++] = (a[j].key < a[i].key) ? a[j++] : a[i++];
    while (i < h) tmp[k++] = a[i++];
    while (j < n) tmp[k++] = a[j++];
    memcpy(a, tmp, (size_t)n * sizeof(kv_t));
}

/* per-active-pixel FIFO shift/insert: ev_graph.cu:196-211 (same body in the single-event kernel :152-165) */
static void fifo_insert(evgo_t *g, int b, int y, int x, const int32_t *sorted_idx, int64_t offset, int counts) {
    const int W = g->W, H = g->H, Q = g->Q;
    for (int q = Q - 1; q >= 0; q--) {
        int64_t index = (int64_t)b * H * W * Q + (int64_t)q * H * W + (int64_t)y * W + x;
        if (q >= counts) {
            int64_t shifted = (int64_t)b * H * W * Q + (int64_t)(q - counts) * H * W + (int64_t)y * W + x;
            g->queue[index] = g->queue[shifted];
        } else {
            g->queue[index] = sorted_idx[offset + counts - 1 - q];
        }
    }
}

/*
 * AsyncGraph.forward (ev_graph.py:63-103) for one call.
 *   batch[n] int32, pos[n][3] int32 (x px, y px, t us)
 *   src_out / dst_out: capacity K*n each; returns the number of edges E (compacted, in buffer order).
 */
int64_t evgo_forward(evgo_t *g, const int32_t *batch, const int32_t *pos, int64_t n,
                     int64_t *src_out, int64_t *dst_out) {
    if (n == 0) return 0;
    const int W = g->W, H = g->H, Q = g->Q, K = g->K;

    /* all_timestamps = cat(all_timestamps, pos[:,2])   (ev_graph.py:75) */
    if (g->n_ts + n > g->cap_ts) {
        while (g->n_ts + n > g->cap_ts) g->cap_ts *= 2;
        g->all_ts = (int32_t *)realloc(g->all_ts, (size_t)g->cap_ts * sizeof(int32_t));
    }
    for (int64_t i = 0; i < n; i++) g->all_ts[g->n_ts + i] = pos[3 * i + 2];
    g->n_ts += n;

    /* indices = max_index + arange(n); max_index += n   (ev_graph.py:82-83) */
    int32_t *indices = (int32_t *)malloc((size_t)n * sizeof(int32_t));
    for (int64_t i = 0; i < n; i++) indices[i] = (int32_t)(g->max_index + i);
    g->max_index += n;

    /* ---- insert (graph/utils.py:6-18) ---- */
    if (n > 1) {
        kv_t *kv = (kv_t *)malloc((size_t)n * sizeof(kv_t));
        kv_t *tmp = (kv_t *)malloc((size_t)n * sizeof(kv_t));
        for (int64_t i = 0; i < n; i++) {
            kv[i].key = (int64_t)pos[3 * i + 0] + (int64_t)W * pos[3 * i + 1] + (int64_t)W * H * batch[i];
            kv[i].idx = indices[i];
        }
        merge_sort(kv, tmp, n);
        int32_t *sorted_idx = (int32_t *)tmp;               /* reuse */
        int64_t *run_start = (int64_t *)malloc((size_t)(n + 1) * sizeof(int64_t));
        int64_t n_runs = 0;
        for (int64_t i = 0; i < n; i++) if (i == 0 || kv[i].key != kv[i - 1].key) run_start[n_runs++] = i;
        run_start[n_runs] = n;
        int64_t *keys = (int64_t *)malloc((size_t)n_runs * sizeof(int64_t));
        for (int64_t r = 0; r < n_runs; r++) keys[r] = kv[run_start[r]].key;
        for (int64_t i = 0; i < n; i++) sorted_idx[i] = kv[i].idx;
        #pragma omp parallel for schedule(static)
        for (int64_t r = 0; r < n_runs; r++) {
            /* coordinate decode of ev_graph.cu:191-193 */
            int64_t c = keys[r];
            int x = (int)(c % W);
            int y = (int)(((c - x) / W) % H);
            int b = (int)(c / ((int64_t)W * H));
            fifo_insert(g, b, y, x, sorted_idx, run_start[r], (int)(run_start[r + 1] - run_start[r]));
        }
        free(keys); free(run_start); free(kv); free(tmp);
    } else {
        /* insert_in_queue_single (ev_graph.cu:130-166): counts = 1, batch index forced to 0 */
        fifo_insert(g, 0, pos[1], pos[0], indices, 0, 1);
    }

    /* ---- search (ev_graph.cu:15-80) into a -1 filled [2][K*n] buffer (ev_graph.py:89) ---- */
    int64_t cap = (int64_t)K * n;
    int64_t *buf_src = (int64_t *)malloc((size_t)cap * sizeof(int64_t));
    int64_t *buf_dst = (int64_t *)malloc((size_t)cap * sizeof(int64_t));
    memset(buf_src, 0xff, (size_t)cap * sizeof(int64_t));
    memset(buf_dst, 0xff, (size_t)cap * sizeof(int64_t));
    const int radius_int = (int)g->radius;                     /* ev_graph.cu:33 */
    const int n_spiral = (2 * radius_int + 1) * (2 * radius_int + 1);
    const int min_index = (int)g->min_index;
    const int max_nb = K;

    #pragma omp parallel for schedule(dynamic, 256)
    for (int64_t e = 0; e < n; e++) {
        int num = 0;
        int64_t off = e * max_nb;
        int b = batch[e], x = pos[3 * e], y = pos[3 * e + 1], ts = pos[3 * e + 2];
        buf_src[off] = indices[e] - min_index;                 /* self edge first (ev_graph.cu:43-46) */
        buf_dst[off] = indices[e] - min_index;
        num++;
        spiral_t sp; spiral_init(&sp);
        for (int i = 0; i < n_spiral; i++) {
            if (num >= max_nb) break;
            for (int q = 0; q < Q; q++) {
                int xn = x + sp.x, yn = y + sp.y;
                if (!(xn >= 0 && yn >= 0 && xn < W && yn < H)) break;          /* out of FOV: leaves q-loop only */
                int64_t qi = (int64_t)xn + (int64_t)W * yn + (int64_t)H * W * q + (int64_t)H * W * Q * b;
                int idx = g->queue[qi];
                if (idx < min_index) break;                                    /* empty or deleted */
                if (indices[e] > idx) {
                    int32_t dt = ts - g->all_ts[idx - min_index];
                    if ((float)dt > g->delta_t_us) continue;                    /* NOT break (ev_graph.cu:69) */
                    buf_src[off + num] = idx - min_index;
                    buf_dst[off + num] = indices[e] - min_index;
                    num++;
                    if (num >= max_nb) break;
                }
            }
            spiral_next(&sp);
        }
    }

    /* edges[:, edges[1] >= 0]  (graph/utils.py:22) */
    int64_t E = 0;
    for (int64_t i = 0; i < cap; i++) {
        if (buf_dst[i] >= 0) { src_out[E] = buf_src[i]; dst_out[E] = buf_dst[i]; E++; }
    }
    free(buf_src); free(buf_dst); free(indices);
    return E;
}


/* ------------------------------------------------------------------------------------------------------------
 * Kernel-level entry points with the pybind signatures of ev_graph.cu:279-283, on caller-owned CPU arrays.
 * oracle/ref_py.py installs them as a CPU stand-in for the `ev_graph_cuda` module, so that the reference's OWN
 * Python glue (graph/utils.py:6-23, graph/ev_graph.py:18-166) runs unmodified above them.
 * ------------------------------------------------------------------------------------------------------------ */
static void fifo_insert_q(int32_t *queue, int W, int H, int Q, int b, int y, int x, const int32_t *sorted_idx,
                          int64_t offset, int counts) {
    for (int q = Q - 1; q >= 0; q--) {                          /* ev_graph.cu:196-211 */
        int64_t index = (int64_t)b * H * W * Q + (int64_t)q * H * W + (int64_t)y * W + x;
        if (q >= counts) {
            int64_t shifted = (int64_t)b * H * W * Q + (int64_t)(q - counts) * H * W + (int64_t)y * W + x;
            queue[index] = queue[shifted];
        } else {
            queue[index] = sorted_idx[offset + counts - 1 - q];
        }
    }
}

/* insert_in_queue_cuda (ev_graph.cu:169-212, launcher :241-276): one "thread" per active pixel */
void evgo_k_insert(int32_t *queue, int B, int Q, int H, int W, const int32_t *sorted_idx, const int32_t *unique_coords,
                   const int32_t *cumsum_counts, int64_t n_runs) {
    (void)B;
    #pragma omp parallel for schedule(static)
    for (int64_t r = 0; r < n_runs; r++) {
        int32_t c = unique_coords[r];
        int x = c % W;
        int y = ((c - x) / W) % H;
        int b = c / (W * H);
        int64_t off = r > 0 ? cumsum_counts[r - 1] : 0;
        int counts = (int)(cumsum_counts[r] - off);
        fifo_insert_q(queue, W, H, Q, b, y, x, sorted_idx, off, counts);
    }
}

/* insert_in_queue_single_cuda (ev_graph.cu:130-166): exactly one event, batch index forced to 0 */
void evgo_k_insert_single(int32_t *queue, int Q, int H, int W, const int32_t *indices, const int32_t *events) {
    fifo_insert_q(queue, W, H, Q, 0, events[1], events[0], indices, 0, 1);
}

/* fill_edges_cuda (ev_graph.cu:15-80, launcher :82-128): writes (src,dst) into slot e*K+num of edges[2][cap] */
void evgo_k_fill_edges(const int32_t *batch, const int32_t *pos, const int32_t *all_ts, const int32_t *queue,
                       const int32_t *indices, int64_t n, int Q, int H, int W, int max_nb, float radius,
                       float delta_t_us, int64_t *edges, int64_t cap, int min_index) {
    const int radius_int = (int)radius;
    const int n_spiral = (2 * radius_int + 1) * (2 * radius_int + 1);
    #pragma omp parallel for schedule(dynamic, 256)
    for (int64_t e = 0; e < n; e++) {
        int num = 0;
        int64_t off = e * max_nb;
        int b = batch[e], x = pos[3 * e], y = pos[3 * e + 1], ts = pos[3 * e + 2];
        edges[off] = indices[e] - min_index;
        edges[cap + off] = indices[e] - min_index;
        num++;
        spiral_t sp; spiral_init(&sp);
        for (int i = 0; i < n_spiral; i++) {
            if (num >= max_nb) break;
            for (int q = 0; q < Q; q++) {
                int xn = x + sp.x, yn = y + sp.y;
                if (!(xn >= 0 && yn >= 0 && xn < W && yn < H)) break;
                int64_t qi = (int64_t)xn + (int64_t)W * yn + (int64_t)H * W * q + (int64_t)H * W * Q * b;
                int idx = queue[qi];
                if (idx < min_index) break;
                if (indices[e] > idx) {
                    int32_t dt = ts - all_ts[idx - min_index];
                    if ((float)dt > delta_t_us) continue;
                    edges[off + num] = idx - min_index;
                    edges[cap + off + num] = indices[e] - min_index;
                    num++;
                    if (num >= max_nb) break;
                }
            }
            spiral_next(&sp);
        }
    }
}

/* first `count` spiral offsets, for tests of the ordering (spiral.h) */
void evgo_spiral(int count, int32_t *dx, int32_t *dy) {
    spiral_t sp; spiral_init(&sp);
    for (int i = 0; i < count; i++) { dx[i] = sp.x; dy[i] = sp.y; spiral_next(&sp); }
}
