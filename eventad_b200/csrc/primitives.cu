// eventad_b200/csrc/primitives.cu -- device-wide exclusive scan and stable LSD radix sort (uint32 pairs).
// Hand-written (no CUB/Thrust): both are used by the graph builder (cell sort), the pooling kernels (cluster
// sort, dense relabel) and the CSR builders.  All work is sized from host-known upper bounds; nothing syncs.
#include "common.cuh"

// ------------------------------------------------------------------------------------------------ scan
#define SCAN_THREADS 512
#define SCAN_ITEMS 4
#define SCAN_TILE (SCAN_THREADS * SCAN_ITEMS)

__device__ __forceinline__ uint32_t warp_inclusive_scan(uint32_t v) {
    const int lane = threadIdx.x & 31;
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) {
        uint32_t t = __shfl_up_sync(0xffffffffu, v, d);
        if (lane >= d) v += t;
    }
    return v;
}

// exclusive scan across the block of one value per thread; returns exclusive prefix, block total in *total
__device__ __forceinline__ uint32_t block_exclusive_scan(uint32_t v, uint32_t* warp_sums /*[32]*/, uint32_t* total) {
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, nwarps = (blockDim.x + 31) >> 5;
    uint32_t inc = warp_inclusive_scan(v);
    if (lane == 31) warp_sums[warp] = inc;
    __syncthreads();
    if (warp == 0) {
        uint32_t w = (lane < nwarps) ? warp_sums[lane] : 0u;
        uint32_t winc = warp_inclusive_scan(w);
        warp_sums[lane] = winc - w;                       // exclusive warp offsets
        if (lane == 31) warp_sums[32] = winc;             // block total
    }
    __syncthreads();
    uint32_t res = warp_sums[warp] + inc - v;
    *total = warp_sums[32];
    __syncthreads();                                      // warp_sums reusable after return
    return res;
}

// per-tile reduction -> tile_sums[tile]
__global__ void __launch_bounds__(SCAN_THREADS) scan_reduce_kernel(const uint32_t* __restrict__ in, long long n,
                                                                   uint32_t* __restrict__ tile_sums) {
    __shared__ uint32_t ws[33];
    long long base = (long long)blockIdx.x * SCAN_TILE + (long long)threadIdx.x * SCAN_ITEMS;
    uint32_t s = 0;
#pragma unroll
    for (int k = 0; k < SCAN_ITEMS; k++)
        if (base + k < n) s += in[base + k];
    uint32_t tot;
    block_exclusive_scan(s, ws, &tot);
    if (threadIdx.x == 0) tile_sums[blockIdx.x] = tot;
}

// single-CTA exclusive scan of n values with a running carry (n small: tile sums, cluster tables, ...)
__global__ void __launch_bounds__(SCAN_THREADS) scan_single_kernel(const uint32_t* __restrict__ in,
                                                                   uint32_t* __restrict__ out, long long n,
                                                                   uint32_t* __restrict__ total) {
    __shared__ uint32_t ws[33];
    uint32_t carry = 0;
    for (long long start = 0; start < n; start += SCAN_TILE) {
        long long base = start + (long long)threadIdx.x * SCAN_ITEMS;
        uint32_t v[SCAN_ITEMS];
        uint32_t s = 0;
#pragma unroll
        for (int k = 0; k < SCAN_ITEMS; k++) {
            v[k] = (base + k < n) ? in[base + k] : 0u;
            s += v[k];
        }
        uint32_t tot;
        uint32_t ex = block_exclusive_scan(s, ws, &tot) + carry;
#pragma unroll
        for (int k = 0; k < SCAN_ITEMS; k++) {
            if (base + k < n) out[base + k] = ex;
            ex += v[k];
        }
        carry += tot;
    }
    if (threadIdx.x == 0 && total != nullptr) *total = carry;
}

// per-tile exclusive scan plus the scanned tile offset
__global__ void __launch_bounds__(SCAN_THREADS) scan_apply_kernel(const uint32_t* __restrict__ in,
                                                                  uint32_t* __restrict__ out, long long n,
                                                                  const uint32_t* __restrict__ tile_offsets) {
    __shared__ uint32_t ws[33];
    long long base = (long long)blockIdx.x * SCAN_TILE + (long long)threadIdx.x * SCAN_ITEMS;
    uint32_t v[SCAN_ITEMS];
    uint32_t s = 0;
#pragma unroll
    for (int k = 0; k < SCAN_ITEMS; k++) {
        v[k] = (base + k < n) ? in[base + k] : 0u;
        s += v[k];
    }
    uint32_t tot;
    uint32_t ex = block_exclusive_scan(s, ws, &tot) + tile_offsets[blockIdx.x];
#pragma unroll
    for (int k = 0; k < SCAN_ITEMS; k++) {
        if (base + k < n) out[base + k] = ex;
        ex += v[k];
    }
}

size_t evad_scan_tmp_elems(long long n) { return (size_t)evad_cdiv(n > 0 ? n : 1, SCAN_TILE) + 64; }

int evad_exclusive_scan_u32(const uint32_t* in, uint32_t* out, long long n, uint32_t* total, uint32_t* tmp,
                            cudaStream_t st) {
    if (n < 0) return EVAD_ERR_ARG;
    if (n <= 4 * SCAN_TILE) {
        scan_single_kernel<<<1, SCAN_THREADS, 0, st>>>(in, out, n, total);
        EVAD_CHECK_LAUNCH();
        return EVAD_OK;
    }
    int tiles = evad_cdiv(n, SCAN_TILE);
    scan_reduce_kernel<<<tiles, SCAN_THREADS, 0, st>>>(in, n, tmp);
    EVAD_CHECK_LAUNCH();
    scan_single_kernel<<<1, SCAN_THREADS, 0, st>>>(tmp, tmp, tiles, total);
    EVAD_CHECK_LAUNCH();
    scan_apply_kernel<<<tiles, SCAN_THREADS, 0, st>>>(in, out, n, tmp);
    EVAD_CHECK_LAUNCH();
    return EVAD_OK;
}

// ------------------------------------------------------------------------------------------------ radix sort
#define RS_THREADS 256
#define RS_WARPS (RS_THREADS / 32)
#define RS_MAX_BINS 1024

// digit histogram of one tile, written digit-major: hist[d * tiles + tile]
template <int ITEMS>
__global__ void __launch_bounds__(RS_THREADS) rs_hist_kernel(const uint32_t* __restrict__ keys, long long n,
                                                             int shift, int bins, uint32_t* __restrict__ hist,
                                                             int tiles) {
    __shared__ uint32_t h[RS_MAX_BINS];
    for (int d = threadIdx.x; d < bins; d += RS_THREADS) h[d] = 0;
    __syncthreads();
    const long long base = (long long)blockIdx.x * (RS_THREADS * ITEMS);
    const uint32_t mask = (uint32_t)bins - 1u;
#pragma unroll
    for (int k = 0; k < ITEMS; k++) {
        long long i = base + (long long)k * RS_THREADS + threadIdx.x;
        if (i < n) atomicAdd(&h[(keys[i] >> shift) & mask], 1u);
    }
    __syncthreads();
    for (int d = threadIdx.x; d < bins; d += RS_THREADS) hist[(size_t)d * tiles + blockIdx.x] = h[d];
}

// stable scatter of one tile.  Order inside a tile = (warp, round, lane) = ascending input index.
template <int ITEMS>
__global__ void __launch_bounds__(RS_THREADS) rs_scatter_kernel(const uint32_t* __restrict__ keys_in,
                                                                const uint32_t* __restrict__ vals_in,
                                                                uint32_t* __restrict__ keys_out,
                                                                uint32_t* __restrict__ vals_out, long long n,
                                                                int shift, int bins,
                                                                const uint32_t* __restrict__ hist_scanned, int tiles) {
    extern __shared__ uint32_t cnt[];                      // [RS_WARPS][bins]
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    for (int d = threadIdx.x; d < RS_WARPS * bins; d += RS_THREADS) cnt[d] = 0;
    __syncthreads();
    const long long wbase = (long long)blockIdx.x * (RS_THREADS * ITEMS) + (long long)warp * (32 * ITEMS);
    const uint32_t mask = (uint32_t)bins - 1u;
    uint32_t key[ITEMS], rank[ITEMS];
    uint32_t* wc = cnt + warp * bins;
    const unsigned lt = evad_lanemask_lt();
#pragma unroll
    for (int r = 0; r < ITEMS; r++) {
        long long i = wbase + r * 32 + lane;
        bool valid = i < n;
        key[r] = valid ? keys_in[i] : 0xffffffffu;
        uint32_t d = valid ? ((key[r] >> shift) & mask) : 0xffffffffu;
        unsigned peers = __match_any_sync(0xffffffffu, d);
        int leader = __ffs(peers) - 1;
        uint32_t old = 0;
        if (valid && lane == leader) {
            old = wc[d];
            wc[d] = old + __popc(peers);
        }
        old = __shfl_sync(0xffffffffu, old, leader);
        rank[r] = old + __popc(peers & lt);
        __syncwarp();
    }
    __syncthreads();
    // exclusive prefix over warps per digit, seeded with the global base of (digit, tile)
    for (int d = threadIdx.x; d < bins; d += RS_THREADS) {
        uint32_t run = hist_scanned[(size_t)d * tiles + blockIdx.x];
#pragma unroll
        for (int w = 0; w < RS_WARPS; w++) {
            uint32_t t = cnt[w * bins + d];
            cnt[w * bins + d] = run;
            run += t;
        }
    }
    __syncthreads();
#pragma unroll
    for (int r = 0; r < ITEMS; r++) {
        long long i = wbase + r * 32 + lane;
        if (i < n) {
            uint32_t d = (key[r] >> shift) & mask;
            uint32_t p = wc[d] + rank[r];
            keys_out[p] = key[r];
            vals_out[p] = vals_in ? vals_in[i] : (uint32_t)i;
        }
    }
}

static inline int rs_items(long long n) { return n >= (1 << 19) ? 8 : 4; }
static inline int rs_tiles(long long n) { return evad_cdiv(n > 0 ? n : 1, RS_THREADS * rs_items(n)); }

size_t evad_sort_tmp_elems(long long n) {
    size_t table = (size_t)RS_MAX_BINS * rs_tiles(n);
    return table + evad_scan_tmp_elems((long long)table) + 64;
}

int evad_sort_pairs_u32(uint32_t* k0, uint32_t* v0, uint32_t* k1, uint32_t* v1, int iota_vals, long long n,
                        int key_bits, uint32_t* tmp, uint32_t** res_k, uint32_t** res_v, cudaStream_t st) {
    if (n < 0 || key_bits < 1 || key_bits > 32) return EVAD_ERR_ARG;
    const int passes = (key_bits + 9) / 10;
    const int bits = (key_bits + passes - 1) / passes;      // <= 10
    const int bins = 1 << bits;
    const int items = rs_items(n);
    const int tiles = rs_tiles(n);
    uint32_t* hist = tmp;
    uint32_t* scan_tmp = tmp + (size_t)RS_MAX_BINS * tiles;
    uint32_t* kb[2] = {k0, k1};
    uint32_t* vb[2] = {v0, v1};
    int cur = 0;                                             // pass p reads buffer cur, writes buffer cur^1
    for (int p = 0; p < passes && n > 0; p++) {
        const uint32_t* cur_k = kb[cur];
        const uint32_t* cur_v = (p == 0 && iota_vals) ? nullptr : vb[cur];
        uint32_t* dst_k = kb[cur ^ 1];
        uint32_t* dst_v = vb[cur ^ 1];
        const int shift = p * bits;
        if (items == 8) {
            rs_hist_kernel<8><<<tiles, RS_THREADS, 0, st>>>(cur_k, n, shift, bins, hist, tiles);
        } else {
            rs_hist_kernel<4><<<tiles, RS_THREADS, 0, st>>>(cur_k, n, shift, bins, hist, tiles);
        }
        EVAD_CHECK_LAUNCH();
        EVAD_TRY(evad_exclusive_scan_u32(hist, hist, (long long)bins * tiles, nullptr, scan_tmp, st));
        size_t smem = (size_t)RS_WARPS * bins * sizeof(uint32_t);
        if (items == 8) {
            rs_scatter_kernel<8><<<tiles, RS_THREADS, smem, st>>>(cur_k, cur_v, dst_k, dst_v, n, shift, bins, hist, tiles);
        } else {
            rs_scatter_kernel<4><<<tiles, RS_THREADS, smem, st>>>(cur_k, cur_v, dst_k, dst_v, n, shift, bins, hist, tiles);
        }
        EVAD_CHECK_LAUNCH();
        cur ^= 1;
    }
    if (n == 0) cur = 1;
    *res_k = kb[cur];
    *res_v = vb[cur];
    return EVAD_OK;
}
