// eventad_b200/csrc/spline_general.cu -- SplineConv for the pooled layers (cin 64..130, cout 64) and Linear.
//
// Reference path replaced (paths relative to /root/reference/src/dagr/model):
//   layers/spline_conv.py:39-47,64-78 (message_lut / PyG basis message, propagate, root weight),
//   layers/components.py:15-22 (Linear), layers/conv.py:23-28,47-56 (BatchNorm, activation, skip add).
// The reference's LUT for these layers would be up to 185x193 matrices of 130x64 floats (1.2 GB); instead the
// sources are aggregated per kernel tap first (step 1) and multiplied by the 25 tap matrices + root matrix in one
// dense fp32 GEMM with a fused BatchNorm/skip/ReLU epilogue (step 2):
//   out_i = sum_k (sum_{e->i} b_{e,k} x_src(e))^T W_k + x_i^T W_root
#include "spline_common.cuh"

// ------------------------------------------------------------------------------------------------ step 1
#define AG_THREADS 128
#define AG_WARPS (AG_THREADS / 32)

__global__ void __launch_bounds__(AG_THREADS) spline_agg_kernel(const int32_t* __restrict__ rowptr,
                                                                const int32_t* __restrict__ src, int64_t n,
                                                                const int32_t* __restrict__ n_dev,
                                                                const float* __restrict__ x, int ldx, int cin,
                                                                EvadPseudo P, EvadSplineMode M, float* __restrict__ A,
                                                                int64_t lda, uint32_t* __restrict__ status) {
    extern __shared__ float rows[];                         // [AG_WARPS][ks*ks*cin]
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int ntap = M.ks * M.ks;
    const int rowlen = ntap * cin;
    float* row = rows + (size_t)warp * rowlen;
    const int64_t nn = n_dev ? min((int64_t)*n_dev, n) : n;
    const int64_t stride = (int64_t)gridDim.x * AG_WARPS;
    for (int64_t i = (int64_t)blockIdx.x * AG_WARPS + warp; i < nn; i += stride) {
        for (int t = lane; t < rowlen; t += 32) row[t] = 0.f;
        __syncwarp();
        const int r0 = rowptr[i], r1 = rowptr[i + 1];
        for (int e = r0; e < r1; e++) {
            const int j = src[e];
            float ax, ay;
            evad_pseudo(P, e, j, (int)i, &ax, &ay);
            int k[4];
            float w[4];
            evad_taps(M, ax, ay, k, w, lane == 0 ? status : status);
            const float* xj = x + (int64_t)j * ldx;
            for (int c = lane; c < cin; c += 32) {
                const float xv = xj[c];
#pragma unroll
                for (int s = 0; s < 4; s++) {
                    float* p = row + k[s] * cin + c;
                    *p = __fadd_rn(*p, __fmul_rn(w[s], xv));
                }
            }
            __syncwarp();
        }
        float* a = A + i * lda;
        for (int t = lane; t < rowlen; t += 32) a[t] = row[t];
        const float* xi = x + i * ldx;
        for (int c = lane; c < cin; c += 32) a[rowlen + c] = xi[c];
        __syncwarp();
    }
}

// Variant for 16-byte aligned rows (ldx % 4 == 0): the source rows of up to 16 in-edges are staged in warp-private
// shared memory with 128-bit loads that are all in flight together (one memory latency per chunk instead of one per
// edge); the taps are then walked in order and each tap row is summed in registers from the staged rows and written
// straight to A (no shared accumulators, no zero fill / copy-out pass).  Per tap the contributions are added in
// ascending edge order with separate fp32 multiply and add, exactly like spline_agg_kernel, so both give the same bits.
#define AG2_WARPS 8
#define AG2_WARPS_H2 4                // H2: a warp also owns a ring of AH2_RING floats through which its rows leave
#define AH2_RING 512                  // floats; >= 256 (flush threshold) + cin + 64 (root slot + K tail)
#define AG2_EDGES 16

// H2 (evad_spline_aggregate_h2): A is not a row-major fp32 matrix but the operand image of evad_gemm64_h2 -- per
// (128-row tile, 64-column K chunk) one 32 KB block [leading fp16 parts 16 KB | fp16 remainders x 2048 16 KB], each a
// 128 x 128-byte tile in the SWIZZLE_128B K-major layout of a UMMA operand (16-byte unit u of row r at (u ^ (r & 7))),
// so that the GEMM fetches a block with ONE bulk copy and hands it to tcgen05.mma without touching it.  `lda` is then
// the number of K chunks per row tile; columns K .. 64 * lda - 1 of valid rows are written as zeros.
#define AH2_BLOCK 32768
template <int CPL, bool H2>
__global__ void __launch_bounds__((H2 ? AG2_WARPS_H2 : AG2_WARPS) * 32) spline_agg2_kernel(const int32_t* __restrict__ rowptr,
                                                                     const int32_t* __restrict__ src, int64_t n,
                                                                     const int32_t* __restrict__ n_dev,
                                                                     const float* __restrict__ x, int ldx, int cin,
                                                                     EvadPseudo P, EvadSplineMode M, float* __restrict__ A,
                                                                     int64_t lda, uint32_t* __restrict__ status, int tsplit) {
    // tsplit > 1 (small inputs, where one node's tap walk IS the kernel's latency): tsplit warps share a node, warp `sub`
    // takes the taps sub, sub + tsplit, ...; every tap is still summed by one warp in ascending edge order (same bits)
    extern __shared__ __align__(16) float ag2_rows[];         // [WARPS][AG2_EDGES][ldx] (+ H2: [WARPS][AH2_RING])
    constexpr int WARPS = H2 ? AG2_WARPS_H2 : AG2_WARPS;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int ntap = M.ks * M.ks;
    const int qpr = ldx >> 2;                                 // 128-bit quads per row
    float* xs = ag2_rows + (size_t)warp * AG2_EDGES * ldx;
    // H2: tap rows go through a small fp32 ring in shared memory (column k at k mod AH2_RING) and leave as 16-byte units
    // (8 columns) of leading parts / remainders as soon as >= 256 columns are complete: all lanes store, and the ring is
    // small enough to keep the occupancy of the fp32 variant.  Rows with more than AG2_EDGES edges add their later
    // chunks to the image by read-modify-write of the same units.
    float* ring = H2 ? ag2_rows + (size_t)WARPS * AG2_EDGES * ldx + (size_t)warp * AH2_RING : nullptr;
    bool h2_ok = true;
    auto ring_flush = [&](int64_t i, int lo, int hi, bool accum) {   // columns [lo, hi), both multiples of 8
        __syncwarp();
        const int r = (int)(i & 127);
        uint8_t* tile = reinterpret_cast<uint8_t*>(A) + (i >> 7) * lda * AH2_BLOCK + r * 128;
        for (int unit = (lo >> 3) + lane; unit < (hi >> 3); unit += 32) {
            const float* rp = ring + ((8 * unit) & (AH2_RING - 1));
            const float4 v0 = *reinterpret_cast<const float4*>(rp), v1 = *reinterpret_cast<const float4*>(rp + 4);
            float v[8] = {v0.x, v0.y, v0.z, v0.w, v1.x, v1.y, v1.z, v1.w};
            uint8_t* dst = tile + (int64_t)(unit >> 3) * AH2_BLOCK + (((unit & 7) ^ (r & 7)) << 4);
            if (accum) {
                const uint4 ph = *reinterpret_cast<const uint4*>(dst), pl = *reinterpret_cast<const uint4*>(dst + AH2_BLOCK / 2);
                const uint32_t hw[4] = {ph.x, ph.y, ph.z, ph.w}, lw[4] = {pl.x, pl.y, pl.z, pl.w};
#pragma unroll
                for (int q = 0; q < 4; q++) {
                    const float2 fh = __half22float2(*reinterpret_cast<const __half2*>(&hw[q]));
                    const float2 fl = __half22float2(*reinterpret_cast<const __half2*>(&lw[q]));
                    v[2 * q] = __fadd_rn(__fmaf_rn(fl.x, 1.f / 2048.f, fh.x), v[2 * q]);
                    v[2 * q + 1] = __fadd_rn(__fmaf_rn(fl.y, 1.f / 2048.f, fh.y), v[2 * q + 1]);
                }
            }
            uint32_t hi2[4], lo2[4];
#pragma unroll
            for (int q = 0; q < 4; q++) {
                h2_ok = h2_ok && fabsf(v[2 * q]) < 32768.f && fabsf(v[2 * q + 1]) < 32768.f;
                const __half2 h = __floats2half2_rn(v[2 * q], v[2 * q + 1]);
                const float2 f = __half22float2(h);
                const __half2 l = __floats2half2_rn(__fmul_rn(__fsub_rn(v[2 * q], f.x), 2048.f),
                                                    __fmul_rn(__fsub_rn(v[2 * q + 1], f.y), 2048.f));
                hi2[q] = *reinterpret_cast<const uint32_t*>(&h);
                lo2[q] = *reinterpret_cast<const uint32_t*>(&l);
            }
            *reinterpret_cast<uint4*>(dst) = make_uint4(hi2[0], hi2[1], hi2[2], hi2[3]);
            *reinterpret_cast<uint4*>(dst + AH2_BLOCK / 2) = make_uint4(lo2[0], lo2[1], lo2[2], lo2[3]);
        }
        __syncwarp();
    };
    const int64_t nn = n_dev ? min((int64_t)*n_dev, n) : n;
    const int64_t gw = (int64_t)blockIdx.x * WARPS + warp;
    const int sub = (int)(gw % tsplit);
    const int64_t stride = (int64_t)gridDim.x * WARPS / tsplit;
    for (int64_t i = gw / tsplit; i < nn; i += stride) {
        const int r0 = rowptr[i], r1 = rowptr[i + 1];
        float* a = H2 ? A : A + i * lda;
        bool first = true;
        for (int e0 = r0; e0 < r1 || first; e0 += AG2_EDGES) {
            const int cnt = max(0, min(AG2_EDGES, r1 - e0));
            int j = 0, k[4] = {-1, -1, -1, -1};
            float w[4] = {0.f, 0.f, 0.f, 0.f};
            if (lane < cnt) {
                j = src[e0 + lane];
                float ax, ay;
                evad_pseudo(P, e0 + lane, j, (int)i, &ax, &ay);
                evad_taps(M, ax, ay, k, w, status);
            }
            __syncwarp();
            for (int base = 0; base < cnt * qpr; base += 32) {   // uniform trip count: every lane takes part in the shuffle
                const int idx = base + lane;
                const bool ok = idx < cnt * qpr;
                const int e = ok ? idx / qpr : 0, q = idx - e * qpr;
                const int jj = __shfl_sync(0xffffffffu, j, e);
                if (ok)
                    reinterpret_cast<float4*>(xs)[idx] = __ldg(reinterpret_cast<const float4*>(x + (int64_t)jj * ldx) + q);
            }
            __syncwarp();
            int flushed = 0;                                  // H2: columns [0, flushed) of this chunk's pass have left the ring
            for (int tap = sub; tap < ntap; tap += tsplit) {
                const float wt = k[0] == tap ? w[0] : (k[1] == tap ? w[1] : (k[2] == tap ? w[2] : w[3]));
                const bool hit = k[0] == tap || k[1] == tap || k[2] == tap || k[3] == tap;
                unsigned mask = __ballot_sync(0xffffffffu, hit);
                float acc[CPL];
#pragma unroll
                for (int u = 0; u < CPL; u++) acc[u] = 0.f;
                const bool any = mask != 0;
                while (mask) {
                    const int e = __ffs(mask) - 1;
                    mask &= mask - 1;
                    const float we = __shfl_sync(0xffffffffu, wt, e);
                    const float* xr = xs + e * ldx;
#pragma unroll
                    for (int u = 0; u < CPL; u++) {
                        const int c = lane + 32 * u;
                        if (c < cin) acc[u] = __fadd_rn(acc[u], __fmul_rn(we, xr[c]));
                    }
                }
                float* at = a + tap * cin;
                if (H2) {
#pragma unroll
                    for (int u = 0; u < CPL; u++) {
                        const int c = lane + 32 * u;
                        if (c < cin) ring[(tap * cin + c) & (AH2_RING - 1)] = acc[u];
                    }
                    const int end8 = ((tap + 1) * cin) & ~7;
                    if (end8 - flushed >= 256) {
                        ring_flush(i, flushed, end8, !first);
                        flushed = end8;
                    }
                } else if (first) {
#pragma unroll
                    for (int u = 0; u < CPL; u++) {
                        const int c = lane + 32 * u;
                        if (c < cin) at[c] = acc[u];
                    }
                } else if (any) {
#pragma unroll
                    for (int u = 0; u < CPL; u++) {
                        const int c = lane + 32 * u;
                        if (c < cin) at[c] = __fadd_rn(at[c], acc[u]);
                    }
                }
            }
            if (H2) {
                // end of the pass over the taps: the first chunk appends the root slot (the node's own row) and the
                // zero K tail; later chunks add nothing there (zeros up to the unit boundary)
                int end = ntap * cin;
                if (first) {
                    const float* xi = x + i * ldx;
#pragma unroll
                    for (int u = 0; u < CPL; u++) {
                        const int c = lane + 32 * u;
                        if (c < cin) ring[(end + c) & (AH2_RING - 1)] = xi[c];
                    }
                    for (int k2 = end + cin + lane; k2 < 64 * (int)lda; k2 += 32) ring[k2 & (AH2_RING - 1)] = 0.f;
                    end = 64 * (int)lda;
                } else {
                    for (int k2 = end + lane; k2 < ((end + 7) & ~7); k2 += 32) ring[k2 & (AH2_RING - 1)] = 0.f;
                    end = (end + 7) & ~7;
                }
                ring_flush(i, flushed, end, !first);
            }
            first = false;
            __syncwarp();
        }
        if (sub == 0 && !H2) {                                // root slot: the node's own row
            const float* xi = x + i * ldx;
#pragma unroll
            for (int u = 0; u < CPL; u++) {
                const int c = lane + 32 * u;
                if (c < cin) a[ntap * cin + c] = xi[c];
            }
        }
    }
    if (H2 && !h2_ok) atomicOr(status, EVAD_STATUS_H16_RANGE);
}

template <int CPL, bool H2 = false>
static int launch_agg2(const int32_t* rowptr, const int32_t* src, int64_t n, const int32_t* n_dev, const float* x, int ldx,
                       int cin, const evad_pseudo_t& P, const evad_spline_mode_t& M, float* A, int64_t lda,
                       uint32_t* status, cudaStream_t st) {
    constexpr int WARPS = H2 ? AG2_WARPS_H2 : AG2_WARPS;
    const size_t smem = (size_t)WARPS * (AG2_EDGES * ldx + (H2 ? AH2_RING : 0)) * sizeof(float);
    if (smem > 200 * 1024) return EVAD_ERR_UNSUPPORTED;
    static size_t attr_bytes = 0;
    if (smem > attr_bytes) {
        if (cudaFuncSetAttribute(spline_agg2_kernel<CPL, H2>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem) !=
            cudaSuccess)
            return EVAD_ERR_LAUNCH;
        attr_bytes = smem;
    }
    const int tsplit = (n <= 2400 && !H2) ? 4 : 1;            // AG2_WARPS % tsplit == 0: a node's warps share a CTA
    const int64_t want = (n * tsplit + WARPS - 1) / WARPS;
    const int64_t cap = 148 * (H2 ? 7 : 4) * tsplit;
    const int grid = (int)(want < cap ? want : cap);
    spline_agg2_kernel<CPL, H2><<<grid, WARPS * 32, smem, st>>>(rowptr, src, n, n_dev, x, ldx, cin, P, M, A, lda, status,
                                                                    tsplit);
    EVAD_CHECK_LAUNCH();
    return EVAD_OK;
}

extern "C" int evad_spline_aggregate(const int32_t* rowptr, const int32_t* src, int64_t n, const int32_t* n_dev,
                                     const float* x, int ldx, int cin, const evad_pseudo_t* pseudo_h,
                                     const evad_spline_mode_t* mode_h, float* A, int64_t lda, uint32_t* status,
                                     void* stream) {
    if (!rowptr || !src || !x || !pseudo_h || !mode_h || !A || !status || cin < 1) return EVAD_ERR_ARG;
    if (mode_h->use_lut && (!mode_h->tab_k || !mode_h->tab_w)) return EVAD_ERR_ARG;
    if (n <= 0) return EVAD_OK;
    const int ntap = mode_h->ks * mode_h->ks;
    if (lda < (int64_t)(ntap + 1) * cin) return EVAD_ERR_ARG;
    if (ldx % 4 == 0 && ((uintptr_t)x & 15) == 0 && cin <= 160 && (size_t)AG2_WARPS * AG2_EDGES * ldx * 4 <= 200 * 1024) {
        const int cpl = (cin + 31) / 32;
        cudaStream_t st = (cudaStream_t)stream;
        switch (cpl) {
        case 1: return launch_agg2<1>(rowptr, src, n, n_dev, x, ldx, cin, *pseudo_h, *mode_h, A, lda, status, st);
        case 2: return launch_agg2<2>(rowptr, src, n, n_dev, x, ldx, cin, *pseudo_h, *mode_h, A, lda, status, st);
        case 3: return launch_agg2<3>(rowptr, src, n, n_dev, x, ldx, cin, *pseudo_h, *mode_h, A, lda, status, st);
        case 4: return launch_agg2<4>(rowptr, src, n, n_dev, x, ldx, cin, *pseudo_h, *mode_h, A, lda, status, st);
        default: return launch_agg2<5>(rowptr, src, n, n_dev, x, ldx, cin, *pseudo_h, *mode_h, A, lda, status, st);
        }
    }
    const size_t smem = (size_t)AG_WARPS * ntap * cin * sizeof(float);
    if (smem > 200 * 1024) return EVAD_ERR_UNSUPPORTED;
    static bool attr_done = false;
    if (!attr_done) {
        if (cudaFuncSetAttribute(spline_agg_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024) != cudaSuccess)
            return EVAD_ERR_LAUNCH;
        attr_done = true;
    }
    int64_t want = (n + AG_WARPS - 1) / AG_WARPS;
    const int64_t cap = 148 * 8;
    int grid = (int)(want < cap ? want : cap);
    spline_agg_kernel<<<grid, AG_THREADS, smem, (cudaStream_t)stream>>>(rowptr, src, n, n_dev, x, ldx, cin, *pseudo_h,
                                                                      *mode_h, A, lda, status);
    EVAD_CHECK_LAUNCH();
    return EVAD_OK;
}

// The same aggregation written as the operand image of evad_gemm64_h2 (see AH2_BLOCK above).  A_h2: evad_gemm64_h2_a_bytes
// (n, K) bytes, K = (ks*ks + 1) * cin.  Values must satisfy |v| < 32768 (EVAD_STATUS_H16_RANGE otherwise).
extern "C" size_t evad_gemm64_h2_a_bytes(int64_t m, int k) {
    if (m <= 0 || k < 1) return 256;
    return (size_t)((m + 127) / 128) * (size_t)((k + 63) / 64) * AH2_BLOCK;
}

extern "C" int evad_spline_aggregate_h2(const int32_t* rowptr, const int32_t* src, int64_t n, const int32_t* n_dev,
                                        const float* x, int ldx, int cin, const evad_pseudo_t* pseudo_h,
                                        const evad_spline_mode_t* mode_h, void* A_h2, uint32_t* status, void* stream) {
    if (!rowptr || !src || !x || !pseudo_h || !mode_h || !A_h2 || !status || cin < 1) return EVAD_ERR_ARG;
    if (mode_h->use_lut && (!mode_h->tab_k || !mode_h->tab_w)) return EVAD_ERR_ARG;
    if (((uintptr_t)A_h2 & 1023) != 0) return EVAD_ERR_ARG;   // blocks are 1024-byte aligned swizzle atoms
    if (n <= 0) return EVAD_OK;
    const int ntap = mode_h->ks * mode_h->ks;
    const int64_t kchunks = ((int64_t)(ntap + 1) * cin + 63) / 64;
    if (!(ldx % 4 == 0 && ((uintptr_t)x & 15) == 0 && cin <= 160 &&
          (size_t)AG2_WARPS_H2 * (AG2_EDGES * ldx + AH2_RING) * 4 <= 200 * 1024 && cin + 320 <= AH2_RING))
        return EVAD_ERR_UNSUPPORTED;
    const int cpl = (cin + 31) / 32;
    cudaStream_t st = (cudaStream_t)stream;
    float* A = reinterpret_cast<float*>(A_h2);
    switch (cpl) {
    case 1: return launch_agg2<1, true>(rowptr, src, n, n_dev, x, ldx, cin, *pseudo_h, *mode_h, A, kchunks, status, st);
    case 2: return launch_agg2<2, true>(rowptr, src, n, n_dev, x, ldx, cin, *pseudo_h, *mode_h, A, kchunks, status, st);
    case 3: return launch_agg2<3, true>(rowptr, src, n, n_dev, x, ldx, cin, *pseudo_h, *mode_h, A, kchunks, status, st);
    case 4: return launch_agg2<4, true>(rowptr, src, n, n_dev, x, ldx, cin, *pseudo_h, *mode_h, A, kchunks, status, st);
    default: return launch_agg2<5, true>(rowptr, src, n, n_dev, x, ldx, cin, *pseudo_h, *mode_h, A, kchunks, status, st);
    }
}

// ------------------------------------------------------------------------------------------------ step 2
// out[m, n0:n0+BN] = epi(A[m,:K] @ W[:K, :]) ; fp32 FFMA, BM x BN tile per CTA, BK = 16, TM x TN micro-tile per thread.
#define GM_BK 16

template <int BM, int BN, int TM, int TN>
__global__ void __launch_bounds__((BM / TM) * (BN / TN)) gemm_epi_kernel(const float* __restrict__ A, int64_t lda,
                                                                         const float* __restrict__ W, int64_t m,
                                                                         const int32_t* __restrict__ m_dev, int K,
                                                                         int N, evad_epilogue_t E,
                                                                         float* __restrict__ out, int ldo) {
    constexpr int THREADS = (BM / TM) * (BN / TN);
    __shared__ float As[GM_BK][BM + 1];
    __shared__ float Bs[GM_BK][BN];
    const int64_t mm = m_dev ? min((int64_t)*m_dev, m) : m;
    const int64_t row0 = (int64_t)blockIdx.x * BM;
    if (row0 >= mm) return;
    const int col0 = blockIdx.y * BN;
    const int tx = threadIdx.x % (BN / TN), ty = threadIdx.x / (BN / TN);
    float acc[TM][TN];
#pragma unroll
    for (int a = 0; a < TM; a++)
#pragma unroll
        for (int b = 0; b < TN; b++) acc[a][b] = 0.f;

    for (int k0 = 0; k0 < K; k0 += GM_BK) {
        // A tile: BM rows x BK cols, threads run along k (contiguous in memory)
        for (int t = threadIdx.x; t < BM * GM_BK; t += THREADS) {
            const int r = t / GM_BK, kk = t % GM_BK;
            const int64_t gr = row0 + r;
            As[kk][r] = (gr < mm && k0 + kk < K) ? A[gr * lda + k0 + kk] : 0.f;
        }
        for (int t = threadIdx.x; t < GM_BK * BN; t += THREADS) {
            const int kk = t / BN, c = t % BN;
            Bs[kk][c] = (k0 + kk < K && col0 + c < N) ? W[(int64_t)(k0 + kk) * N + col0 + c] : 0.f;
        }
        __syncthreads();
#pragma unroll
        for (int kk = 0; kk < GM_BK; kk++) {
            float av[TM], bv[TN];
#pragma unroll
            for (int a = 0; a < TM; a++) av[a] = As[kk][ty * TM + a];
#pragma unroll
            for (int b = 0; b < TN; b++) bv[b] = Bs[kk][tx * TN + b];
#pragma unroll
            for (int a = 0; a < TM; a++)
#pragma unroll
                for (int b = 0; b < TN; b++) acc[a][b] = fmaf(av[a], bv[b], acc[a][b]);
        }
        __syncthreads();
    }
#pragma unroll
    for (int a = 0; a < TM; a++) {
        const int64_t gr = row0 + ty * TM + a;
        if (gr >= mm) continue;
#pragma unroll
        for (int b = 0; b < TN; b++) {
            const int c = col0 + tx * TN + b;
            if (c >= N) continue;
            float v = acc[a][b];
            if (E.scale) v = __fmul_rn(v, E.scale[c]);
            if (E.shift) v = __fadd_rn(v, E.shift[c]);
            if (E.addend) v = __fadd_rn(v, E.addend[gr * E.ld_add + c]);
            if (E.relu) v = fmaxf(v, 0.f);
            out[gr * ldo + c] = v;
        }
    }
}

// N = 64 fast path: BM x 64 tile, BK = 16, 256 threads as 16 (rows) x 16 (column quads), TM = BM/16 rows x 4 columns
// per thread; A is staged transposed so both operands are read with 128-bit shared loads (3 LDS.128 per 32 FFMA at
// BM = 128); the next k-slab is prefetched into registers while the current one is multiplied.
template <int BM>
__global__ void __launch_bounds__(256) gemm64_kernel(const float* __restrict__ A, int64_t lda, const float* __restrict__ W,
                                                     int64_t m, const int32_t* __restrict__ m_dev, int K_total,
                                                     int k_chunk, float* __restrict__ partial,
                                                     evad_epilogue_t E, float* __restrict__ out, int ldo) {
    constexpr int TM = BM / 16;
    constexpr int AV = (BM * GM_BK / 4 + 255) / 256;          // float4 loads of A per thread per slab
    __shared__ __align__(16) float As[GM_BK][BM + 4];
    __shared__ __align__(16) float Bs[GM_BK][64];
    const int64_t mm = m_dev ? min((int64_t)*m_dev, m) : m;
    const int64_t row0 = (int64_t)blockIdx.x * BM;
    if (row0 >= mm) return;
    // split-K: blockIdx.y owns columns [kb0, K) of A / rows of W; partial sums go to `partial` (deterministic reduce)
    const int kb0 = blockIdx.y * k_chunk;
    const int K = min(K_total, kb0 + k_chunk);
    const int tid = threadIdx.x, tx = tid & 15, ty = tid >> 4;
    float acc[TM][4];
#pragma unroll
    for (int a = 0; a < TM; a++) { acc[a][0] = acc[a][1] = acc[a][2] = acc[a][3] = 0.f; }
    float4 ra[AV], rb;
    auto load_slab = [&](int k0) {
#pragma unroll
        for (int v = 0; v < AV; v++) {
            const int idx = tid + v * 256;
            const int r = idx >> 2, kq = idx & 3;
            const int64_t gr = row0 + r;
            const int kk = k0 + kq * 4;
            float4 t = make_float4(0.f, 0.f, 0.f, 0.f);
            if (idx < BM * 4 && gr < mm) {
                const float* p = A + gr * lda + kk;
                if (kk + 3 < K) t = *reinterpret_cast<const float4*>(p);
                else {
                    if (kk < K) t.x = p[0];
                    if (kk + 1 < K) t.y = p[1];
                    if (kk + 2 < K) t.z = p[2];
                }
            }
            ra[v] = t;
        }
        const int kb = k0 + (tid >> 4);
        rb = (kb < K) ? *reinterpret_cast<const float4*>(W + (int64_t)kb * 64 + 4 * tx) : make_float4(0.f, 0.f, 0.f, 0.f);
    };
    auto store_slab = [&]() {
#pragma unroll
        for (int v = 0; v < AV; v++) {
            const int idx = tid + v * 256;
            const int r = idx >> 2, kq = idx & 3;
            if (idx >= BM * 4) continue;
            As[kq * 4 + 0][r] = ra[v].x; As[kq * 4 + 1][r] = ra[v].y;
            As[kq * 4 + 2][r] = ra[v].z; As[kq * 4 + 3][r] = ra[v].w;
        }
        *reinterpret_cast<float4*>(&Bs[tid >> 4][4 * tx]) = rb;
    };
    load_slab(kb0);
    for (int k0 = kb0; k0 < K; k0 += GM_BK) {
        store_slab();
        __syncthreads();
        if (k0 + GM_BK < K) load_slab(k0 + GM_BK);
#pragma unroll
        for (int kk = 0; kk < GM_BK; kk++) {
            const float4 b4 = *reinterpret_cast<const float4*>(&Bs[kk][4 * tx]);
            float av[TM];
#pragma unroll
            for (int a = 0; a < TM; a += 2) {
                const float2 t = *reinterpret_cast<const float2*>(&As[kk][ty * TM + a]);
                av[a] = t.x; av[a + 1] = t.y;
            }
#pragma unroll
            for (int a = 0; a < TM; a++) {
                acc[a][0] = fmaf(av[a], b4.x, acc[a][0]); acc[a][1] = fmaf(av[a], b4.y, acc[a][1]);
                acc[a][2] = fmaf(av[a], b4.z, acc[a][2]); acc[a][3] = fmaf(av[a], b4.w, acc[a][3]);
            }
        }
        __syncthreads();
    }
#pragma unroll
    for (int a = 0; a < TM; a++) {
        const int64_t gr = row0 + ty * TM + a;
        if (gr >= mm) continue;
        float v[4] = {acc[a][0], acc[a][1], acc[a][2], acc[a][3]};
        if (partial) {
            *reinterpret_cast<float4*>(partial + ((int64_t)blockIdx.y * m + gr) * 64 + 4 * tx) = make_float4(v[0], v[1], v[2], v[3]);
            continue;
        }
#pragma unroll
        for (int b = 0; b < 4; b++) {
            const int c = 4 * tx + b;
            if (E.scale) v[b] = __fmul_rn(v[b], E.scale[c]);
            if (E.shift) v[b] = __fadd_rn(v[b], E.shift[c]);
            if (E.addend) v[b] = __fadd_rn(v[b], E.addend[gr * E.ld_add + c]);
            if (E.relu) v[b] = fmaxf(v[b], 0.f);
            out[gr * ldo + c] = v[b];
        }
    }
}

// sums the split-K partials in slice order (deterministic) and applies the epilogue
__global__ void __launch_bounds__(256) gemm64_reduce_kernel(const float* __restrict__ partial, int splits, int64_t m,
                                                            const int32_t* __restrict__ m_dev, evad_epilogue_t E,
                                                            float* __restrict__ out, int ldo) {
    const int64_t t = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    const int64_t mm = m_dev ? min((int64_t)*m_dev, m) : m;
    const int64_t gr = t >> 4;
    const int tx = (int)(t & 15);
    if (gr >= mm) return;
    float4 s4 = *reinterpret_cast<const float4*>(partial + gr * 64 + 4 * tx);
    for (int y = 1; y < splits; y++) {
        const float4 p = *reinterpret_cast<const float4*>(partial + ((int64_t)y * m + gr) * 64 + 4 * tx);
        s4.x += p.x; s4.y += p.y; s4.z += p.z; s4.w += p.w;
    }
    float v[4] = {s4.x, s4.y, s4.z, s4.w};
#pragma unroll
    for (int b = 0; b < 4; b++) {
        const int c = 4 * tx + b;
        if (E.scale) v[b] = __fmul_rn(v[b], E.scale[c]);
        if (E.shift) v[b] = __fadd_rn(v[b], E.shift[c]);
        if (E.addend) v[b] = __fadd_rn(v[b], E.addend[gr * E.ld_add + c]);
        if (E.relu) v[b] = fmaxf(v[b], 0.f);
        out[gr * ldo + c] = v[b];
    }
}

// number of K slices for the small-M path (1 = unsplit); shared by the workspace query and the launcher
static int gemm64_splits(int64_t m, int k, int* k_chunk) {
    *k_chunk = k;
    if (m >= 148 * 64) return 1;
    const int mtiles = evad_cdiv(m, 32);
    int splits = (2 * 148 + mtiles - 1) / mtiles;
    const int max_splits_k = (k + 63) / 64;
    if (splits > max_splits_k) splits = max_splits_k;
    if (splits > 32) splits = 32;
    if (splits < 2) return 1;
    *k_chunk = ((k + splits - 1) / splits + GM_BK - 1) / GM_BK * GM_BK;   // multiple of 16 keeps A loads 16 B aligned
    return (k + *k_chunk - 1) / *k_chunk;
}

extern "C" size_t evad_gemm_ws_bytes(int64_t m, int k, int n) {
    if (n != 64 || m <= 0 || k < 1) return 256;
    int kc;
    const int splits = gemm64_splits(m, k, &kc);
    return (splits > 1 ? (size_t)splits * (size_t)m * 64 * sizeof(float) : 0) + 256;
}

extern "C" int evad_gemm_epilogue(const float* A, int64_t lda, const float* W, int64_t m, const int32_t* m_dev, int k,
                                  int n, const evad_epilogue_t* epi_h, float* out, int ldo, void* ws, size_t ws_bytes,
                                  void* stream) {
    if (!A || !W || !epi_h || !out || k < 1 || n < 1 || lda < k || ldo < n) return EVAD_ERR_ARG;
    if (m <= 0) return EVAD_OK;
    cudaStream_t st = (cudaStream_t)stream;
    if (n == 64 && lda % 4 == 0 && (((uintptr_t)A | (uintptr_t)W) & 15) == 0) {
        if (m >= 148 * 64) {
            gemm64_kernel<128><<<evad_cdiv(m, 128), 256, 0, st>>>(A, lda, W, m, m_dev, k, k, nullptr, *epi_h, out, ldo);
            EVAD_CHECK_LAUNCH();
            return EVAD_OK;
        }
        // small M (pooled graphs of a few windows): split K so that about two waves of CTAs exist
        const int mtiles = evad_cdiv(m, 32);
        int k_chunk;
        const int splits = gemm64_splits(m, k, &k_chunk);
        const size_t need = (size_t)splits * (size_t)m * 64 * sizeof(float);
        if (splits < 2 || !ws || ws_bytes < need) {
            gemm64_kernel<32><<<mtiles, 256, 0, st>>>(A, lda, W, m, m_dev, k, k, nullptr, *epi_h, out, ldo);
            EVAD_CHECK_LAUNCH();
            return EVAD_OK;
        }
        gemm64_kernel<32><<<dim3(mtiles, splits), 256, 0, st>>>(A, lda, W, m, m_dev, k, k_chunk, (float*)ws, *epi_h, out, ldo);
        EVAD_CHECK_LAUNCH();
        gemm64_reduce_kernel<<<evad_cdiv(m * 16, 256), 256, 0, st>>>((const float*)ws, splits, m, m_dev, *epi_h, out, ldo);
        EVAD_CHECK_LAUNCH();
        return EVAD_OK;
    }
    if (m >= 148 * 64) {
        dim3 grid(evad_cdiv(m, 64), evad_cdiv(n, 64));
        gemm_epi_kernel<64, 64, 4, 4><<<grid, 256, 0, st>>>(A, lda, W, m, m_dev, k, n, *epi_h, out, ldo);
    } else {
        dim3 grid(evad_cdiv(m, 16), evad_cdiv(n, 64));
        gemm_epi_kernel<16, 64, 1, 4><<<grid, 256, 0, st>>>(A, lda, W, m, m_dev, k, n, *epi_h, out, ldo);
    }
    EVAD_CHECK_LAUNCH();
    return EVAD_OK;
}
