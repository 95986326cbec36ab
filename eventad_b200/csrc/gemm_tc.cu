// eventad_b200/csrc/gemm_tc.cu -- tensor-core GEMM with fp32-grade accuracy (3xTF32) for the SplineConv weight
// contraction:  out[m, 0:64] = epi(A[m, 0:K] @ W[K, 64]).
//
// Reference path replaced (paths relative to /root/reference/src/dagr/model): the per-kernel-slot weight contraction of
// layers/spline_conv.py:39-47,64-78 (einsum / spline_weighting + root `lin`) and layers/components.py:15-22 (Linear),
// with layers/conv.py:23-28,47-56 (BatchNorm, skip add, activation) in the epilogue.
//
// Round-1 ncu evidence (profiles/r01_summary.md): this contraction is FMA-issue bound on the CUDA cores, not HBM bound,
// which is the case north_star reserves tensor cores for.  fp32 parity (rtol 1e-4 / atol 1e-5) rules out plain TF32
// (10-bit mantissa), so every product is split  a = a_hi + a_lo, w = w_hi + w_lo  (each part exactly a TF32 number) and
// accumulated as  a_lo*w_hi + a_hi*w_lo + a_hi*w_hi  in fp32: 3 tensor-core MMAs per tile, ~2^-21 relative error.
// Both operands are split in registers while the fragments are loaded (hi = upper 19 bits, lo = exact remainder, of
// which the tensor core again reads the upper 19 bits), so W streams from L2 once as plain fp32.
// evad_split_tf32 (round-to-nearest parts) serves kernels that keep a weight resident in shared memory (spline_l1.cu).
#include "common.cuh"

#define TG_BK 32
#define TG_LDA 36          // A tile row stride in floats: bank = 4*g + t  -> conflict-free fragment loads
#define TG_LDB 72          // W tile row stride in floats: bank = 8*t + g  -> conflict-free fragment loads
#define TG_STAGES 3
#define TG_THREADS 128

__device__ __forceinline__ uint32_t tg_tf32(float x) {
    uint32_t r;
    asm("cvt.rna.tf32.f32 %0, %1;" : "=r"(r) : "f"(x));
    return r;
}

__device__ __forceinline__ void tg_mma(float (&d)[4], const uint32_t (&a)[4], const uint32_t (&b)[2]) {
    asm volatile(
        "mma.sync.aligned.m16n8k8.row.col.f32.tf32.tf32.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
        : "+f"(d[0]), "+f"(d[1]), "+f"(d[2]), "+f"(d[3])
        : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(b[0]), "r"(b[1]));
}

__device__ __forceinline__ void tg_cp16(void* smem, const void* gmem, int src_bytes) {
    const uint32_t s = (uint32_t)__cvta_generic_to_shared(smem);
    asm volatile("cp.async.cg.shared.global [%0], [%1], 16, %2;" ::"r"(s), "l"(gmem), "r"(src_bytes));
}
__device__ __forceinline__ void tg_commit() { asm volatile("cp.async.commit_group;"); }
template <int N>
__device__ __forceinline__ void tg_wait() { asm volatile("cp.async.wait_group %0;" ::"n"(N)); }

__global__ void __launch_bounds__(256) split_tf32_kernel(const float* __restrict__ w, int64_t n, float* __restrict__ hi,
                                                         float* __restrict__ lo) {
    const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const float v = w[i];
    const float h = __uint_as_float(tg_tf32(v));
    hi[i] = h;
    lo[i] = __uint_as_float(tg_tf32(__fsub_rn(v, h)));
}

extern "C" int evad_split_tf32(const float* w, int64_t n, float* hi, float* lo, void* stream) {
    if (!w || !hi || !lo || n < 0) return EVAD_ERR_ARG;
    if (n == 0) return EVAD_OK;
    split_tf32_kernel<<<evad_cdiv(n, 256), 256, 0, (cudaStream_t)stream>>>(w, n, hi, lo);
    EVAD_CHECK_LAUNCH();
    return EVAD_OK;
}

// BM x 64 tile per CTA; 4 warps as WM x WN; warp tile (BM/WM) x (64/WN) = MT m16-tiles x NT n8-tiles.
// A and W slabs of 32 k-columns are staged by cp.async in a 3-deep ring (zero-filled beyond m and K).
// blockIdx.y owns the k-range [y*k_chunk, (y+1)*k_chunk): with `partial` the raw sums go to partial[y][m][64]
// (summed in slice order by tg_reduce_kernel -- deterministic), otherwise the epilogue is applied here.
template <int BM, int WM, int WN>
__global__ void __launch_bounds__(TG_THREADS) gemm64_tf32x3_kernel(
    const float* __restrict__ A, int64_t lda, const float* __restrict__ W, int64_t m,
    const int32_t* __restrict__ m_dev, int K_total, int k_chunk, float* __restrict__ partial, evad_epilogue_t E,
    float* __restrict__ out, int ldo) {
    constexpr int MT = BM / WM / 16, NT = 64 / WN / 8;
    extern __shared__ __align__(16) float tg_smem[];
    float* As = tg_smem;                                     // [STAGES][BM][LDA]
    float* Bs = As + TG_STAGES * BM * TG_LDA;                // [STAGES][BK][LDB], fp32 W (split in registers)
    const int64_t mm = m_dev ? min((int64_t)*m_dev, m) : m;
    const int64_t row0 = (int64_t)blockIdx.x * BM;
    if (row0 >= mm) return;
    const int kb0 = blockIdx.y * k_chunk;
    const int K = min(K_total, kb0 + k_chunk);
    const int nk = (K - kb0 + TG_BK - 1) / TG_BK;
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int wm = warp / WN, wn = warp % WN, g = lane >> 2, t = lane & 3;

    auto load_stage = [&](int slot, int k0) {
        for (int idx = tid; idx < BM * 8; idx += TG_THREADS) {
            const int r = idx >> 3, c = idx & 7;
            const int64_t gr = row0 + r;
            const int kk = k0 + 4 * c;
            int bytes = (gr < mm) ? min(16, (K - kk) * 4) : 0;
            bytes = max(bytes, 0);
            const float* src = bytes > 0 ? A + gr * lda + kk : A;
            tg_cp16(As + (slot * BM + r) * TG_LDA + 4 * c, src, bytes);
        }
        for (int idx = tid; idx < TG_BK * 16; idx += TG_THREADS) {
            const int r = idx >> 4, c = idx & 15;
            const int kk = k0 + r;
            const int bytes = kk < K ? 16 : 0;
            const int64_t off = bytes ? (int64_t)kk * 64 + 4 * c : 0;
            tg_cp16(Bs + (slot * TG_BK + r) * TG_LDB + 4 * c, W + off, bytes);
        }
    };

    float acc[MT][NT][4];
#pragma unroll
    for (int a = 0; a < MT; a++)
#pragma unroll
        for (int b = 0; b < NT; b++) { acc[a][b][0] = acc[a][b][1] = acc[a][b][2] = acc[a][b][3] = 0.f; }

#pragma unroll
    for (int s = 0; s < TG_STAGES - 1; s++) {
        if (s < nk) load_stage(s, kb0 + s * TG_BK);
        tg_commit();
    }
    for (int kt = 0; kt < nk; kt++) {
        tg_wait<TG_STAGES - 2>();
        __syncthreads();
        {
            const int nxt = kt + TG_STAGES - 1;
            if (nxt < nk) load_stage(nxt % TG_STAGES, kb0 + nxt * TG_BK);
            tg_commit();
        }
        const int slot = kt % TG_STAGES;
        // the tensor cores truncate when they align the fp32 accumulator: sum each 32-column slab from zero and add it
        // to the running total with a round-to-nearest FADD, so the truncation bias cannot build up over K
        float part[MT][NT][4];
#pragma unroll
        for (int a = 0; a < MT; a++)
#pragma unroll
            for (int b = 0; b < NT; b++) { part[a][b][0] = part[a][b][1] = part[a][b][2] = part[a][b][3] = 0.f; }
        const float* a_s = As + (slot * BM + wm * (BM / WM)) * TG_LDA;
        const float* b_s = Bs + slot * TG_BK * TG_LDB + wn * (64 / WN);
#pragma unroll
        for (int ks = 0; ks < TG_BK / 8; ks++) {
            const int kk = ks * 8;
            uint32_t ah[MT][4], al[MT][4];
#pragma unroll
            for (int mt = 0; mt < MT; mt++) {
                const float* p = a_s + (mt * 16 + g) * TG_LDA + kk + t;
                const float v[4] = {p[0], p[8 * TG_LDA], p[4], p[8 * TG_LDA + 4]};
#pragma unroll
                for (int q = 0; q < 4; q++) {
                    // hi = upper 19 bits (what the tensor core reads of a TF32 operand), lo = exact remainder
                    ah[mt][q] = __float_as_uint(v[q]) & 0xffffe000u;
                    al[mt][q] = __float_as_uint(__fsub_rn(v[q], __uint_as_float(ah[mt][q])));
                }
            }
#pragma unroll
            for (int nt = 0; nt < NT; nt++) {
                const int o0 = (kk + t) * TG_LDB + nt * 8 + g, o1 = o0 + 4 * TG_LDB;
                const float w0 = b_s[o0], w1 = b_s[o1];
                const uint32_t bh[2] = {__float_as_uint(w0) & 0xffffe000u, __float_as_uint(w1) & 0xffffe000u};
                const uint32_t bl[2] = {__float_as_uint(__fsub_rn(w0, __uint_as_float(bh[0]))),
                                        __float_as_uint(__fsub_rn(w1, __uint_as_float(bh[1])))};
#pragma unroll
                for (int mt = 0; mt < MT; mt++) {
                    tg_mma(part[mt][nt], al[mt], bh);
                    tg_mma(part[mt][nt], ah[mt], bl);
                    tg_mma(part[mt][nt], ah[mt], bh);
                }
            }
        }
#pragma unroll
        for (int a = 0; a < MT; a++)
#pragma unroll
            for (int b = 0; b < NT; b++)
#pragma unroll
                for (int q = 0; q < 4; q++) acc[a][b][q] += part[a][b][q];
    }
    tg_wait<0>();

#pragma unroll
    for (int mt = 0; mt < MT; mt++)
#pragma unroll
        for (int h = 0; h < 2; h++) {
            const int64_t gr = row0 + wm * (BM / WM) + mt * 16 + g + 8 * h;
            if (gr >= mm) continue;
#pragma unroll
            for (int nt = 0; nt < NT; nt++) {
                const int c = wn * (64 / WN) + nt * 8 + 2 * t;
                float v[2] = {acc[mt][nt][2 * h], acc[mt][nt][2 * h + 1]};
                if (partial) {
                    *reinterpret_cast<float2*>(partial + ((int64_t)blockIdx.y * m + gr) * 64 + c) = make_float2(v[0], v[1]);
                    continue;
                }
#pragma unroll
                for (int b = 0; b < 2; b++) {
                    if (E.scale) v[b] = __fmul_rn(v[b], E.scale[c + b]);
                    if (E.shift) v[b] = __fadd_rn(v[b], E.shift[c + b]);
                    if (E.addend) v[b] = __fadd_rn(v[b], E.addend[gr * E.ld_add + c + b]);
                    if (E.relu) v[b] = fmaxf(v[b], 0.f);
                }
                *reinterpret_cast<float2*>(out + gr * ldo + c) = make_float2(v[0], v[1]);
            }
        }
}

// sums the split-K partials in slice order (deterministic) and applies the epilogue
__global__ void __launch_bounds__(256) tg_reduce_kernel(const float* __restrict__ partial, int splits, int64_t m,
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
    }
    *reinterpret_cast<float4*>(out + gr * ldo + 4 * tx) = make_float4(v[0], v[1], v[2], v[3]);
}

// shared with gemm_tc5.cu (declared in common.cuh)
int evad_gemm64_reduce(const float* partial, int splits, int64_t m, const int32_t* m_dev, const evad_epilogue_t* E,
                       float* out, int ldo, cudaStream_t st) {
    tg_reduce_kernel<<<evad_cdiv(m * 16, 256), 256, 0, st>>>(partial, splits, m, m_dev, *E, out, ldo);
    EVAD_CHECK_LAUNCH();
    return EVAD_OK;
}

// tile height and number of K slices: enough CTAs to fill the resident slots of the 148 SMs
static void tg_plan(int64_t m, int k, int* bm, int* splits, int* k_chunk) {
    *bm = (m >= 64 * 148) ? 64 : 32;
    const int mtiles = evad_cdiv(m, *bm);
    int s = 1;
    if (mtiles < 4 * 148) {                                  // 4 CTAs are resident per SM (55 KB of shared memory each)
        s = (4 * 148 + mtiles - 1) / mtiles;
        const int max_s = (k + 4 * TG_BK - 1) / (4 * TG_BK);      // at least 4 slabs per slice
        if (s > max_s) s = max_s;
        if (s > 32) s = 32;
        if (s < 1) s = 1;
    }
    int kc = ((k + s - 1) / s + TG_BK - 1) / TG_BK * TG_BK;
    *k_chunk = kc;
    *splits = (k + kc - 1) / kc;
}

extern "C" size_t evad_gemm64_tc_ws_bytes(int64_t m, int k) {
    if (m <= 0 || k < 1) return 256;
    int bm, splits, kc;
    tg_plan(m, k, &bm, &splits, &kc);
    return (splits > 1 ? (size_t)splits * (size_t)m * 64 * sizeof(float) : 0) + 256;
}

template <int BM, int WM, int WN>
static int tg_launch(const float* A, int64_t lda, const float* W, int64_t m, const int32_t* m_dev,
                     int k, int splits, int k_chunk, float* partial, const evad_epilogue_t& E, float* out, int ldo,
                     cudaStream_t st) {
    const size_t smem = (size_t)(TG_STAGES * BM * TG_LDA + TG_STAGES * TG_BK * TG_LDB) * sizeof(float);
    static bool attr_done = false;
    if (!attr_done) {
        if (cudaFuncSetAttribute(gemm64_tf32x3_kernel<BM, WM, WN>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                 (int)smem) != cudaSuccess)
            return EVAD_ERR_LAUNCH;
        attr_done = true;
    }
    dim3 grid(evad_cdiv(m, BM), splits);
    gemm64_tf32x3_kernel<BM, WM, WN><<<grid, TG_THREADS, smem, st>>>(A, lda, W, m, m_dev, k, k_chunk, partial, E,
                                                                     out, ldo);
    EVAD_CHECK_LAUNCH();
    return EVAD_OK;
}

extern "C" int evad_gemm64_tc(const float* A, int64_t lda, const float* W, int64_t m, const int32_t* m_dev, int k, const evad_epilogue_t* epi_h, float* out, int ldo, void* ws,
                              size_t ws_bytes, void* stream) {
    if (!A || !W || !epi_h || !out || k < 1 || lda < k || ldo < 64) return EVAD_ERR_ARG;
    if (lda % 4 != 0 || ldo % 2 != 0 || (((uintptr_t)A | (uintptr_t)W) & 15) || ((uintptr_t)out & 7))
        return EVAD_ERR_UNSUPPORTED;
    if (m <= 0) return EVAD_OK;
    cudaStream_t st = (cudaStream_t)stream;
    int bm, splits, kc;
    tg_plan(m, k, &bm, &splits, &kc);
    float* partial = nullptr;
    if (splits > 1) {
        const size_t need = (size_t)splits * (size_t)m * 64 * sizeof(float);
        if (ws && ws_bytes >= need) partial = (float*)ws;
        else { splits = 1; kc = (k + TG_BK - 1) / TG_BK * TG_BK; }
    }
    if (bm == 64) EVAD_TRY((tg_launch<64, 2, 2>(A, lda, W, m, m_dev, k, splits, kc, partial, *epi_h, out, ldo, st)));
    else EVAD_TRY((tg_launch<32, 1, 4>(A, lda, W, m, m_dev, k, splits, kc, partial, *epi_h, out, ldo, st)));
    if (partial) {
        if (((uintptr_t)out & 15) || ldo % 4 != 0) return EVAD_ERR_UNSUPPORTED;
        tg_reduce_kernel<<<evad_cdiv(m * 16, 256), 256, 0, st>>>(partial, splits, m, m_dev, *epi_h, out, ldo);
        EVAD_CHECK_LAUNCH();
    }
    return EVAD_OK;
}
