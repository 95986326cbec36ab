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
// The search runs one thread per event in pixel-sorted order, so neighbouring threads read the same records; the
// records (index, timestamp) of the 2r+1 pixels of an image row are one contiguous span, which is what the default search
// (g_search_rows_kernel) reads instead of (2r+1)^2 buckets; the reference's per-cell walk (g_walk_event) remains for
// crowded neighbourhoods, overflowed FIFOs and dense scenes.  Kernel order of one build:
//   g_keys -> 3 x (rs_hist, scan, rs_scatter) -> g_cells (+ g_cells_big) -> g_search_rows -> g_search_list -> g_search
// of which the last three each decide on the device whether they have work (flags[], g_dense_scene).
#include "common.cuh"
#include <stdlib.h>
#include <string.h>

#define G_THREADS 128
#define G_GAP_BIG 4096           // runs of empty cells longer than this are filled by the whole grid (g_cells_big_kernel)

// flags[0]: entries of g_cells_kernel's big-gap list; flags[1]: entries of the walk list; flags[2]: events with at least
// three later events on their pixel.  A scene counts as dense when more than 1/5 of its events are such (about three
// events per pixel): every event is then walked by g_search_kernel, which stops at K neighbours after a few pixels,
// while the row-span search would read all (2r+1)^2 buckets of every event.
__device__ __forceinline__ bool g_dense_scene(const uint32_t* __restrict__ flags, int64_t n) {
    return (unsigned long long)flags[2] * 5ull > (unsigned long long)n;
}
static int g_span_limit = 32;
// tuning knob (tests force the per-cell walk with it): an event whose (2r+1)-pixel row spans all hold at most
// min(limit, Q, 32) records is searched over row spans, any other one by the per-cell walk; 0 = always walk.
// Returns the previous value; a negative argument only reads it.
extern "C" int evad_graph_set_span_limit(int v) {
    const int old = g_span_limit;
    if (v >= 0) g_span_limit = v > 32 ? 32 : v;
    return old;
}

__global__ void __launch_bounds__(256) g_keys_kernel(const int32_t* __restrict__ batch, const int32_t* __restrict__ pos,
                                                     int64_t n, int B, int H, int W, uint32_t* __restrict__ keys,
                                                     uint32_t* __restrict__ status, uint32_t* __restrict__ flags) {
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
    if (i == 0) flags[0] = flags[1] = flags[2] = 0u;
}

// cell_start[c] = first sorted position whose key >= c  (c = 0..ncells), plus sorted-order (index, timestamp) records.
__global__ void __launch_bounds__(256) g_cells_kernel(const uint32_t* __restrict__ skey, const uint32_t* __restrict__ sidx,
                                                      const int32_t* __restrict__ pos, int64_t n, int64_t ncells,
                                                      int32_t* __restrict__ cell_start, int2* __restrict__ rec,
                                                      int32_t* __restrict__ cell_perm, uint32_t* __restrict__ flags,
                                                      int4* __restrict__ gaps, int W, int xs) {
    int64_t t = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    {
        // position t (0..n) is the start of every cell in (skey[t-1], skey[t]] (keys -1 before the first and ncells behind
        // the last event): short runs are written by the thread, long ones (empty image regions) by its whole warp
        const bool has = t <= n;
        const int64_t prev = (has && t > 0) ? (int64_t)skey[t - 1] : -1;
        const int64_t cur = !has ? prev : t < n ? (int64_t)skey[t] : ncells;
        const int64_t first = prev + 1;
        int64_t cnt = cur - prev;
        if (cnt > G_GAP_BIG) {                                // e.g. the images of a batch that hold no events
            gaps[atomicAdd(&flags[0], 1u)] = make_int4((int)first, (int)cnt, (int)t, 0);
            cnt = 0;
        }
        if (cnt <= 4)
            for (int64_t k = 0; k < cnt; k++) cell_start[first + k] = (int32_t)t;
        const int lane = threadIdx.x & 31;
        for (uint32_t big = __ballot_sync(0xffffffffu, cnt > 4); big; big &= big - 1) {
            const int l = __ffs((int)big) - 1;
            const int64_t f = __shfl_sync(0xffffffffu, first, l), c = __shfl_sync(0xffffffffu, cnt, l);
            const int32_t v = (int32_t)__shfl_sync(0xffffffffu, t, l);
            for (int64_t k = lane; k < c; k += 32) cell_start[f + k] = v;
        }
    }
    if (t < n) {
        uint32_t i = sidx[t];
        // (event index, timestamp): one 64-bit load per candidate.  xs = 5: the index sits above the low 5 bits of the
        // pixel column (order and `<` on the word are those of the index), read by g_search_rows_kernel
        rec[t] = make_int2((int)((i << xs) | (xs ? (skey[t] % (uint32_t)W) & 31u : 0u)), pos[3 * (int64_t)i + 2]);
        if (cell_perm) cell_perm[t] = (int32_t)i;
    }
    const int crowded = __syncthreads_count(t + 3 < n && skey[t] == skey[t + 3]);
    if (threadIdx.x == 0 && crowded) atomicAdd(&flags[2], (uint32_t)crowded);
}

__global__ void __launch_bounds__(256) g_cells_big_kernel(const uint32_t* __restrict__ flags, const int4* __restrict__ gaps,
                                                          int32_t* __restrict__ cell_start) {
    const uint32_t ng = flags[0];
    for (uint32_t e = 0; e < ng; e++) {
        const int4 g = gaps[e];
        for (int64_t k = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; k < g.y; k += (int64_t)gridDim.x * blockDim.x)
            cell_start[(int64_t)g.x + k] = g.z;
    }
}

// The per-cell walk for the event at sorted position p: self loop, then the spiral over (2r+1)^2 pixels, newest-first over
// the newest Q records of each pixel, until K neighbours.  sdx / sdy: the walk of spiral.h:1-15 (layer starts at 1, leg 0),
// tabulated (the per-thread state machine cost more instructions per cell visit than the visit itself).
// xs: rec[].x = index << xs | low bits (g_cells_kernel).
__device__ __forceinline__ void g_walk_event(int64_t p, const uint32_t* __restrict__ skey, const int2* __restrict__ rec,
                                             const int32_t* __restrict__ cell_start, int H, int W, int radius,
                                             float delta_t, int Q, int K, int32_t* __restrict__ nbr,
                                             uint8_t* __restrict__ nbr_cls, uint32_t* __restrict__ count, int xs,
                                             const int8_t* sdx, const int8_t* sdy) {
    const int nsp = (2 * radius + 1) * (2 * radius + 1);
    const int2 me = rec[p];
    const int i = me.x >> xs;
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
        const int sx = sdx[s], sy = sdy[s];
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
                    if ((rec[m].x >> xs) < i) a = m + 1; else b = m;
                }
                q = a - 1;
            }
            for (; q >= lo; q--) {                            // newest first
                int2 r = rec[q];
                r.x >>= xs;
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

// one thread per event, in pixel-sorted order: the per-cell walk for every event (any radius / Q / delta_t)
__global__ void __launch_bounds__(G_THREADS) g_search_kernel(const uint32_t* __restrict__ skey,
                                                             const int2* __restrict__ rec,
                                                             const int32_t* __restrict__ cell_start, int64_t n, int H,
                                                             int W, int radius, float delta_t, int Q, int K,
                                                             int32_t* __restrict__ nbr /*[n][K]*/,
                                                             uint8_t* __restrict__ nbr_cls /*[n][K] or null*/,
                                                             uint32_t* __restrict__ count /*[n]*/, int xs,
                                                             const uint32_t* __restrict__ flags /*null: always*/) {
    if (flags && !g_dense_scene(flags, n)) return;            // sparse scene: g_search_rows_kernel has taken the call
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
    g_walk_event(p, skey, rec, cell_start, H, W, radius, delta_t, Q, K, nbr, nbr_cls, count, xs, g_spiral, g_spiral + nsp);
}

// the per-cell walk for the events g_search_rows_kernel has put on its list
__global__ void __launch_bounds__(G_THREADS) g_search_list_kernel(const uint32_t* __restrict__ skey,
                                                                  const int2* __restrict__ rec,
                                                                  const int32_t* __restrict__ cell_start, int H, int W,
                                                                  int radius, float delta_t, int Q, int K,
                                                                  int32_t* __restrict__ nbr, uint8_t* __restrict__ nbr_cls,
                                                                  uint32_t* __restrict__ count, int xs,
                                                                  const uint32_t* __restrict__ flags,
                                                                  const int32_t* __restrict__ walk_list) {
    const uint32_t total = flags[1];                          // 0 for a dense scene (the row-span kernel has not run)
    if ((uint32_t)blockIdx.x * G_THREADS >= total) return;
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
    const uint32_t tix = (uint32_t)blockIdx.x * G_THREADS + threadIdx.x;
    if (tix >= total) return;
    g_walk_event(walk_list[tix], skey, rec, cell_start, H, W, radius, delta_t, Q, K, nbr, nbr_cls, count, xs, g_spiral,
                 g_spiral + nsp);
}

// The same search over ROW SPANS, for radius <= 7, |delta_t| < 2^24 (the float time gate is then the integer test
// dt <= delta_t) and n < 2^26.  In pixel-sorted order the (2r+1) pixels of one image row are ONE contiguous span of `rec`, so
// a thread reads 2r+1 spans (two cell_start loads per row) instead of (2r+1)^2 buckets of a different length in every lane:
//   pass 1  admissible entries (older than i, inside the time gate) of each row as a bit mask over the 32 entries from
//           the span's (even) start, read two per 128-bit load;
//   count   per admissible entry: spiral position of its pixel (from the 5 column bits packed under the index in the
//           entry's first word and a per-row table in the kernel parameters) -> one byte counter per position in shared
//           memory;
//   prefix  counters -> first output slot of each position (four positions per 32-bit word);
//   place   per admissible entry, newest first inside a pixel: slot = counter++, written straight into the row.
// The result is the list the per-cell walk produces (spiral position, then newest first); nothing is sorted.
// An event one of whose spans holds more than min(32, Q, span_limit) records -- a crowded neighbourhood, where the walk
// stops after a few pixels while the spans would all have to be read, or a pixel whose FIFO may have overflowed, which
// only the walk models -- is put on a list for g_search_list_kernel (the per-cell walk) instead: sparse regions and hot
// pixels of one scene each get the search that suits them.
__device__ __forceinline__ uint32_t gsr_lds8(uint32_t a) { uint32_t v; asm volatile("ld.shared.u8 %0, [%1];" : "=r"(v) : "r"(a) : "memory"); return v; }
__device__ __forceinline__ uint32_t gsr_lds32(uint32_t a) { uint32_t v; asm volatile("ld.shared.u32 %0, [%1];" : "=r"(v) : "r"(a) : "memory"); return v; }
__device__ __forceinline__ void gsr_sts8(uint32_t a, uint32_t v) { asm volatile("st.shared.u8 [%0], %1;" ::"r"(a), "r"(v) : "memory"); }
__device__ __forceinline__ void gsr_sts32(uint32_t a, uint32_t v) { asm volatile("st.shared.u32 [%0], %1;" ::"r"(a), "r"(v) : "memory"); }

struct GsrTable {
    unsigned long long lo[15], hi[15];   // row j = dy + R: spiral position of dx + R = 0..7 (lo) / 8..14 (hi), one byte each
};

template <int R>
__global__ void __launch_bounds__(G_THREADS, R <= 4 ? 10 : 4) g_search_rows_kernel(
    const uint32_t* __restrict__ skey, const int2* __restrict__ rec /* .x = index << 5 | column & 31 */,
    const int32_t* __restrict__ cell_start, int64_t n, int H, int W, int delta_t, int Q, int K, int32_t* __restrict__ nbr,
    uint8_t* __restrict__ nbr_cls, uint32_t* __restrict__ count, int span_limit, const __grid_constant__ GsrTable T,
    uint32_t* __restrict__ flags, int32_t* __restrict__ walk_list) {
    if (g_dense_scene(flags, n)) return;                                   // g_search_kernel walks every event
    constexpr int ROWS = 2 * R + 1, NSP = ROWS * ROWS, NW = (NSP + 3) / 4;
    __shared__ __align__(16) uint32_t s_slot[NW * G_THREADS];              // [NW][thread]: 4 position counters per word
    // shared-space address behind an opaque move: ptxas otherwise rebuilds it (S2R CgaCtaId + 2 more) at every use
    uint32_t slot_a = (uint32_t)__cvta_generic_to_shared(s_slot) + threadIdx.x * 4;   // this thread's counter column
    asm volatile("mov.u32 %0, %0;" : "+r"(slot_a));
#pragma unroll
    for (int w = 0; w < NW; w++) gsr_sts32(slot_a + w * (4 * G_THREADS), 0u);
    int64_t p = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (p >= n) return;
    const int2 me = rec[p];
    const int iw = me.x & ~31, ti = me.y;                                  // entries older than i: word < iw
    const int c = (int)skey[p];
    const int rem = c % (W * H);
    const int y = rem / W, x = rem - y * W;
    const int xl = max(x - R, 0) - x, xh = min(x + R, W - 1) - x;          // clipped dx range
    const int KM = K - 1;
    const int xr = R - x;                                                  // (column + xr) & 31 = dx + R
    const int4* rec2 = reinterpret_cast<const int4*>(rec);                 // entries (2k, 2k + 1)
    int ra[ROWS], rb[ROWS];                                                // span of row j (empty outside the image)
    uint32_t mask[ROWS];
    int longest = 0;
#pragma unroll
    for (int j = 0; j < ROWS; j++) {
        const bool in = (unsigned)(y + j - R) < (unsigned)H;
        const int rowc = in ? c + (j - R) * W : c;
        const int a = cell_start[rowc + xl], b = cell_start[rowc + xh + 1];
        ra[j] = a;
        rb[j] = in ? b : a;
        longest = max(longest, rb[j] - (a & ~1));
    }
    {   // span_limit <= min(32, Q).  Events for the walk are appended to a list (warp-aggregated: neighbours stay together)
        // that g_search_list_kernel, with its own occupancy, works off right behind this kernel
        const bool walk = longest > span_limit;
        const uint32_t wm = __ballot_sync(__activemask(), walk);
        if (walk) {
            const int lane = threadIdx.x & 31, leader = __ffs((int)wm) - 1;
            uint32_t base = 0;
            if (lane == leader) base = atomicAdd(&flags[1], (uint32_t)__popc(wm));
            base = __shfl_sync(wm, base, leader);
            walk_list[base + __popc(wm & ((1u << lane) - 1u))] = (int32_t)p;
            return;
        }
    }
    // ---- pass 1
#pragma unroll
    for (int j = 0; j < ROWS; j++) {
        const int a = ra[j], b = rb[j];
        const int a2 = a & ~1, len = b - a2;                               // <= 32
        ra[j] = a2;
        uint32_t m = 0u;
        int done = 0;
#pragma unroll 2
        for (int q2 = a2 >> 1; done < len; q2++, done += 2) {              // the entry past an odd end is masked off below
            const int4 v = rec2[q2];
            m = __funnelshift_r(m, (v.x < iw && ti - v.y <= delta_t) ? 1u : 0u, 1);
            m = __funnelshift_r(m, (v.z < iw && ti - v.w <= delta_t) ? 1u : 0u, 1);
        }
        m = done ? m >> (32 - done) : 0u;                                  // bit k <-> entry a2 + k
        if (a & 1) m &= ~1u;
        if (len < 32) m &= (1u << max(len, 0)) - 1u;
        mask[j] = m;
    }
    // spiral position of an entry of row j from its packed word, and the address of this thread's counter for it
    auto position = [&](int j, int word) -> uint32_t {
        const int k = (word + xr) & 31;                                    // dx + R
        const unsigned long long t = (ROWS > 8 && k >= 8) ? T.hi[j] : T.lo[j];
        return (uint32_t)(t >> ((k & 7) * 8)) & 255u;
    };
    auto counter = [&](uint32_t s) -> uint32_t { return slot_a + ((s & 0xfcu) << 7) + (s & 3u); };
    // ---- count
#pragma unroll
    for (int j = 0; j < ROWS; j++) {
#pragma unroll 1
        for (uint32_t m = mask[j]; m; m &= m - 1) {
            const uint32_t h = counter(position(j, rec[ra[j] + __ffs((int)m) - 1].x));
            gsr_sts8(h, gsr_lds8(h) + 1);                                  // <= 32
        }
    }
    // ---- prefix: byte k of a word becomes the first slot of its position; everything saturates at 31 >= KM
    uint32_t run = 0u;
#pragma unroll
    for (int w = 0; w < NW; w++) {
        const uint32_t v = __vminu4(gsr_lds32(slot_a + w * (4 * G_THREADS)), 0x1f1f1f1fu);
        const uint32_t incl = v * 0x01010101u;                             // bytes <= 124
        gsr_sts32(slot_a + w * (4 * G_THREADS), (incl << 8) + run * 0x01010101u);   // bytes <= 124 + 31
        run = min(run + (incl >> 24), 31u);
    }
    const int i = me.x >> 5;
    const int m_out = min((int)run, KM);
    int32_t* out = nbr + (int64_t)i * K;
    uint8_t* ocls = nbr_cls ? nbr_cls + (int64_t)i * K : nullptr;
    out[0] = i;                                                            // self edge first (ev_graph.cu:43-46)
    if (ocls) ocls[0] = (uint8_t)(R * ROWS + R);
    count[i] = (uint32_t)(m_out + 1);
    // ---- place, newest first inside every pixel
#pragma unroll
    for (int j = 0; j < ROWS; j++) {
#pragma unroll 1
        for (uint32_t m = mask[j]; m; ) {
            const int bit = 31 - __clz((int)m);
            m &= ~(1u << bit);
            const int word = rec[ra[j] + bit].x;
            const uint32_t h = counter(position(j, word));
            const int slot = (int)gsr_lds8(h);
            if (slot < KM) {
                gsr_sts8(h, (uint32_t)slot + 1);
                out[1 + slot] = word >> 5;
                if (ocls) ocls[1 + slot] = (uint8_t)(((word + xr) & 31) * ROWS + j);   // (dx + R) * (2R + 1) + dy + R
            }
        }
    }
}

// the walk of spiral.h:1-15 (layer starts at 1, leg 0) for one radius: positions packed per row, and step by step
static GsrTable gsr_table(int R) {
    GsrTable T;
    memset(&T, 0, sizeof(T));
    int sx = 0, sy = 0, layer = 1, leg = 0;
    for (int s = 0; s < (2 * R + 1) * (2 * R + 1); s++) {
        const int j = sy + R, k = sx + R;
        (k >= 8 ? T.hi[j] : T.lo[j]) |= (unsigned long long)s << ((k & 7) * 8);
        switch (leg) {
        case 0: ++sx; if (sx == layer) ++leg; break;
        case 1: ++sy; if (sy == layer) ++leg; break;
        case 2: --sx; if (-sx == layer) ++leg; break;
        default: --sy; if (-sy == layer) { leg = 0; ++layer; } break;
        }
    }
    return T;
}

template <int R>
static void g_search_rows_launch(const uint32_t* skey, const int2* rec, const int32_t* cell_start, int64_t n, int H, int W,
                                 int delta_t, int Q, int K, int32_t* nbr, uint8_t* nbr_cls, uint32_t* count,
                                 uint32_t* flags, int32_t* walk_list, cudaStream_t st) {
    static const GsrTable T = gsr_table(R);
    const int span_limit = g_span_limit < Q ? g_span_limit : Q;           // <= 32
    g_search_rows_kernel<R><<<evad_cdiv(n, G_THREADS), G_THREADS, 0, st>>>(skey, rec, cell_start, n, H, W, delta_t, Q, K, nbr,
                                                                         nbr_cls, count, span_limit, T, flags, walk_list);
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
    uint32_t* flags;
    int4* gaps;
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
    w->flags = a.take<uint32_t>(4);
    w->gaps = a.take<int4>(ncells / G_GAP_BIG + 2);
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
    // the row-span kernel (which hands crowded neighbourhoods to the per-cell walk itself) covers radius 1..7, K >= 2,
    // |delta_t| < 2^24, n < 2^26 and sparse scenes (g_dense_scene, found on the device: one of the two paths returns at
    // once); anything else is walked cell by cell.  EVAD_GRAPH_ROWSPAN=0 forces the walk.
    static const bool rowspan_on = [] { const char* e = getenv("EVAD_GRAPH_ROWSPAN"); return !(e && e[0] == '0'); }();
    const int nsp = (2 * radius + 1) * (2 * radius + 1);
    const bool rowspan = rowspan_on && radius >= 1 && radius <= 7 && K >= 2 && delta_t_us > -(1 << 24) &&
                         delta_t_us < (1 << 24) && n < (1ll << 26);
    const int xs = rowspan ? 5 : 0;
    g_keys_kernel<<<evad_cdiv(n, 256), 256, 0, st>>>(batch, pos, n, B, H, W, w.k0, status, w.flags);
    EVAD_CHECK_LAUNCH();
    int key_bits = 1;
    while ((1ll << key_bits) < ncells) key_bits++;
    uint32_t *skey, *sidx;
    EVAD_TRY(evad_sort_pairs_u32(w.k0, w.v0, w.k1, w.v1, 1, n, key_bits, w.sort_tmp, &skey, &sidx, st));
    g_cells_kernel<<<evad_cdiv(n + 1, 256), 256, 0, st>>>(skey, sidx, pos, n, ncells, cell_start, w.rec, cell_perm, w.flags, w.gaps, W, xs);
    EVAD_CHECK_LAUNCH();
    g_cells_big_kernel<<<2 * 148, 256, 0, st>>>(w.flags, w.gaps, cell_start);
    EVAD_CHECK_LAUNCH();
    if (rowspan) {
        // the walk list lives in the key buffer the sort has left free
        int32_t* walk_list = reinterpret_cast<int32_t*>(skey == w.k0 ? w.k1 : w.k0);
        switch (radius) {
#define GSR_CASE(RR) case RR: g_search_rows_launch<RR>(skey, w.rec, cell_start, n, H, W, delta_t_us, Q, K, nbr, nbr_cls, count, w.flags, walk_list, st); break;
        GSR_CASE(1) GSR_CASE(2) GSR_CASE(3) GSR_CASE(4) GSR_CASE(5) GSR_CASE(6) GSR_CASE(7)
#undef GSR_CASE
        }
        EVAD_CHECK_LAUNCH();
        g_search_list_kernel<<<evad_cdiv(n, G_THREADS), G_THREADS, 2 * nsp, st>>>(skey, w.rec, cell_start, H, W, radius, (float)delta_t_us,
                                                                             Q, K, nbr, nbr_cls, count, xs, w.flags, walk_list);
        EVAD_CHECK_LAUNCH();
    }
    g_search_kernel<<<evad_cdiv(n, G_THREADS), G_THREADS, 2 * nsp, st>>>(skey, w.rec, cell_start, n, H, W, radius,
                                                                    (float)delta_t_us, Q, K, nbr, nbr_cls, count, xs,
                                                                    rowspan ? w.flags : nullptr);
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
