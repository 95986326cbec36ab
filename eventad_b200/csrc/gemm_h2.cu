// eventad_b200/csrc/gemm_h2.cu -- the SplineConv weight contraction of the pooled layers as a bulk-copy fed, warp-specialised
// tcgen05 pipeline on fp16 operand pairs:   out[m, 0:64] = epi(A[m, 0:K] @ W[K, 64]).
//
// Reference path replaced: as csrc/gemm_tc5.cu (layers/spline_conv.py:39-47,64-78 einsum / spline_weighting + `lin`,
// layers/conv.py:23-28,47-56 in the epilogue).
//
// Why (profiles/r01h_summary.md): gemm64_tcgen05_kernel streams the fp32 operand A[n, 26*Cin] at
// 27-34 % of the HBM peak -- every 32-column slab goes cp.async -> split into TF32 hi/lo by all threads in shared memory
// -> __syncthreads -> 12 MMAs -> drain, in lock step.  Here nothing touches the operand between HBM and the tensor core:
//   * evad_spline_aggregate_h2 writes A as fp16 pairs (v = a1 + a2 * 2^-11, the same 4 bytes per element) ALREADY in the
//     shared-memory image of a UMMA operand: per (128-row tile, 64-column K chunk) one 32 KB block [a1 tile | a2 tile],
//     each 128 rows x 128 bytes, SWIZZLE_128B K-major.  The weights are packed the same way once on the host
//     (16 KB per chunk: [w1 | w2], 64 rows each).
//   * one producer thread fetches a stage (A block + W block, 48 KB) with two cp.async.bulk copies that complete on an
//     mbarrier (the TMA engine, no tensor map needed because the source is already tiled); 4 stages in flight.
//   * one MMA thread issues per stage 4 x { a1 x [w1 | w2] (N = 128), a2 x w1 (N = 64, into the second half) }
//     tcgen05.mma kind::f16 into one of two 128-column TMEM buffers and commits to the stage's "empty" barrier and the
//     buffer's "full" barrier.
//   * eight drain warps add the buffer to fp32 registers (the tensor core truncates when it accumulates, so a buffer
//     only collects one 64-column chunk from zero) and release it; epilogue out = acc1 + acc2 * 2^-11, BN / skip / ReLU,
//     or raw split-K partials for gemm_tc.cu's deterministic reduce.
// No __syncthreads in the main loop; the three roles meet only on mbarriers.
#include "tc5.cuh"

#define GH_THREADS 320                 // warp 0: producer, warp 1: MMA issuer, warps 2..9: drain
#define GH_STAGES 4
#define GH_A_BYTES 32768               // [a1 | a2] block of one (row tile, K chunk)
#define GH_W_BYTES 16384               // [w1 | w2] block of one K chunk
#define GH_STAGE_BYTES (GH_A_BYTES + GH_W_BYTES)
#define GH_TMEM_COLS 256               // two buffers x (64 columns a1 w1 | 64 columns a1 w2 + a2 w1)
#define GH_BK 64

// instruction descriptor: D = F32 (1 << 4), A = B = F16 (0), both K-major, N >> 3 at bit 17, M = 128 (M >> 4 at bit 24)
#define GH_IDESC(N) ((1u << 4) | (((uint32_t)(N) >> 3) << 17) | ((128u >> 4) << 24))

__device__ __forceinline__ void gh_umma(uint32_t tmem_d, uint64_t adesc, uint64_t bdesc, uint32_t idesc, uint32_t accumulate) {
    asm volatile(
        "{\n\t"
        ".reg .pred p;\n\t"
        "setp.ne.b32 p, %4, 0;\n\t"
        "tcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n\t"
        "}\n" ::"r"(tmem_d), "l"(adesc), "l"(bdesc), "r"(idesc), "r"(accumulate)
        : "memory");
}
__device__ __forceinline__ void gh_expect_tx(uint32_t bar, uint32_t bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(bytes) : "memory");
}
__device__ __forceinline__ void gh_bulk_load(uint32_t dst, const void* src, uint32_t bytes, uint32_t bar) {
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(dst),
                 "l"(src), "r"(bytes), "r"(bar)
                 : "memory");
}

__global__ void __launch_bounds__(GH_THREADS, 1) gemm64_h2_kernel(
    const uint8_t* __restrict__ A, const uint8_t* __restrict__ W, int64_t m, const int32_t* __restrict__ m_dev,
    int kchunks_total, int kchunks_per_cta, float* __restrict__ partial, evad_epilogue_t E, float* __restrict__ out, int ldo) {
    extern __shared__ uint8_t gh_smem[];
    __shared__ __align__(8) uint64_t s_full[GH_STAGES], s_empty[GH_STAGES], s_accfull[2], s_accempty[2];
    __shared__ uint32_t s_tmem;
    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const int64_t mm = m_dev ? min((int64_t)*m_dev, m) : m;
    const int64_t row0 = (int64_t)blockIdx.x * 128;
    if (row0 >= mm) return;                                  // whole CTA leaves before any barrier / TMEM allocation
    const int kc0 = blockIdx.y * kchunks_per_cta;
    const int nk = min(kchunks_total, kc0 + kchunks_per_cta) - kc0;
    const uint32_t st_base = (t5_smem(gh_smem) + 1023u) & ~1023u;       // SWIZZLE_128B tiles: 1024-byte aligned

    if (tid == 0) {
        for (int s = 0; s < GH_STAGES; s++) {
            t5_mbar_init(t5_smem(&s_full[s]), 1);
            t5_mbar_init(t5_smem(&s_empty[s]), 1);
        }
        for (int b = 0; b < 2; b++) {
            t5_mbar_init(t5_smem(&s_accfull[b]), 1);
            t5_mbar_init(t5_smem(&s_accempty[b]), 8);        // one arrival per drain warp
        }
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    if (warp == 0) {
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(t5_smem(&s_tmem)),
                     "n"(GH_TMEM_COLS)
                     : "memory");
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
    }
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
    const uint32_t tmem = s_tmem;

    float acc1[32], acc2[32];                                // drain warps: row 32 * (warp & 3) + lane, 32 of the 64 columns
#pragma unroll
    for (int j = 0; j < 32; j++) acc1[j] = acc2[j] = 0.f;

    if (warp == 0) {
        // ---------------------------------------------------------------- producer: two bulk copies per stage
        if (lane == 0) {
            const uint8_t* asrc = A + ((int64_t)blockIdx.x * kchunks_total + kc0) * GH_A_BYTES;
            const uint8_t* wsrc = W + (int64_t)kc0 * GH_W_BYTES;
            for (int it = 0; it < nk; it++) {
                const int s = it % GH_STAGES;
                t5_mbar_wait(t5_smem(&s_empty[s]), (uint32_t)(((it / GH_STAGES) & 1) ^ 1));   // first round passes
                const uint32_t bar = t5_smem(&s_full[s]), dst = st_base + s * GH_STAGE_BYTES;
                gh_expect_tx(bar, GH_STAGE_BYTES);
                gh_bulk_load(dst, asrc + (int64_t)it * GH_A_BYTES, GH_A_BYTES, bar);
                gh_bulk_load(dst + GH_A_BYTES, wsrc + (int64_t)it * GH_W_BYTES, GH_W_BYTES, bar);
            }
        }
    } else if (warp == 1) {
        // ---------------------------------------------------------------- MMA issuer
        if (lane == 0) {
            for (int it = 0; it < nk; it++) {
                const int s = it % GH_STAGES, b = it & 1;
                t5_mbar_wait(t5_smem(&s_accempty[b]), (uint32_t)(((it >> 1) & 1) ^ 1));       // buffer drained (first use passes)
                t5_mbar_wait(t5_smem(&s_full[s]), (uint32_t)((it / GH_STAGES) & 1));          // stage landed
                asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
                const uint32_t sa = st_base + s * GH_STAGE_BYTES;
                const uint64_t a1 = t5_desc(sa), a2 = t5_desc(sa + GH_A_BYTES / 2);
                const uint64_t w12 = t5_desc(sa + GH_A_BYTES);                                 // 128 rows: w1 then w2
                const uint32_t d = tmem + (uint32_t)(b * 128);
#pragma unroll
                for (int k = 0; k < GH_BK / 16; k++) {       // 16 halves = 32 bytes = 2 descriptor units per K step
                    const uint64_t o = (uint64_t)(2 * k);
                    gh_umma(d, a1 + o, w12 + o, GH_IDESC(128), k != 0);       // [a1 w1 | a1 w2]
                    gh_umma(d + 64u, a2 + o, w12 + o, GH_IDESC(64), 1u);      // + a2 w1 into the second half
                }
                t5_mma_commit(t5_smem(&s_empty[s]));         // the stage may be refilled when these MMAs are done
                t5_mma_commit(t5_smem(&s_accfull[b]));       // ... and the buffer drained
            }
        }
    } else {
        // ---------------------------------------------------------------- drain: TMEM -> fp32 registers, round to nearest
        const int quarter = warp & 3, chalf = (warp - 2) >> 2;
        for (int it = 0; it < nk; it++) {
            const int b = it & 1;
            t5_mbar_wait(t5_smem(&s_accfull[b]), (uint32_t)((it >> 1) & 1));
            asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
            const uint32_t taddr = tmem + (uint32_t)(b * 128 + chalf * 32) + ((uint32_t)(quarter * 32) << 16);
            uint32_t r1[32], r2[32];
            t5_ld32(taddr, r1);
            t5_ld32(taddr + 64u, r2);
            asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
            __syncwarp();
            if (lane == 0) t5_mbar_arrive(t5_smem(&s_accempty[b]));
#pragma unroll
            for (int j = 0; j < 32; j++) {
                acc1[j] += __uint_as_float(r1[j]);
                acc2[j] += __uint_as_float(r2[j]);
            }
        }
        // ---- epilogue: thread owns output row 32 * quarter + lane, columns 32 * chalf .. +31
        const int64_t gr = row0 + quarter * 32 + lane;
        if (gr < mm) {
#pragma unroll
            for (int j = 0; j < 32; j += 4) {
                const int c = chalf * 32 + j;
                float v[4];
#pragma unroll
                for (int q = 0; q < 4; q++) v[q] = __fmaf_rn(acc2[j + q], 1.f / 2048.f, acc1[j + q]);
                if (partial) {
                    *reinterpret_cast<float4*>(partial + ((int64_t)blockIdx.y * m + gr) * 64 + c) =
                        make_float4(v[0], v[1], v[2], v[3]);
                    continue;
                }
#pragma unroll
                for (int q = 0; q < 4; q++) {
                    if (E.scale) v[q] = __fmul_rn(v[q], E.scale[c + q]);
                    if (E.shift) v[q] = __fadd_rn(v[q], E.shift[c + q]);
                    if (E.addend) v[q] = __fadd_rn(v[q], E.addend[gr * E.ld_add + c + q]);
                    if (E.relu) v[q] = fmaxf(v[q], 0.f);
                }
                *reinterpret_cast<float4*>(out + gr * ldo + c) = make_float4(v[0], v[1], v[2], v[3]);
            }
        }
    }
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    if (warp == 0)
        asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem), "n"(GH_TMEM_COLS) : "memory");
}

// K chunks per CTA so that every SM gets a CTA (one CTA per SM: 193 KB of shared memory)
static void gh_plan(int64_t m, int kchunks, int* splits, int* per_cta) {
    const int mtiles = evad_cdiv(m, 128);
    int s = 1;
    if (mtiles < 148) {
        s = 148 / mtiles;
        const int max_s = (kchunks + 3) / 4;                 // at least four chunks per CTA
        if (s > max_s) s = max_s;
        if (s > 32) s = 32;
        if (s < 1) s = 1;
    }
    const int pc = (kchunks + s - 1) / s;
    *per_cta = pc;
    *splits = (kchunks + pc - 1) / pc;
}

extern "C" size_t evad_gemm64_h2_w_bytes(int k) { return (size_t)((k + 63) / 64) * GH_W_BYTES; }

extern "C" size_t evad_gemm64_h2_ws_bytes(int64_t m, int k) {
    if (m <= 0 || k < 1) return 256;
    int splits, pc;
    gh_plan(m, (k + 63) / 64, &splits, &pc);
    return (splits > 1 ? (size_t)splits * (size_t)m * 64 * sizeof(float) : 0) + 256;
}

extern "C" int evad_gemm64_h2(const void* A_h2, const void* W_h2, int64_t m, const int32_t* m_dev, int k,
                              const evad_epilogue_t* epi_h, float* out, int ldo, void* ws, size_t ws_bytes, void* stream) {
    if (!A_h2 || !W_h2 || !epi_h || !out || k < 1 || ldo < 64) return EVAD_ERR_ARG;
    if (ldo % 4 != 0 || (((uintptr_t)A_h2 | (uintptr_t)W_h2) & 1023) || ((uintptr_t)out & 15)) return EVAD_ERR_UNSUPPORTED;
    if (m <= 0) return EVAD_OK;
    cudaStream_t st = (cudaStream_t)stream;
    const int kchunks = (k + 63) / 64;
    int splits, pc;
    gh_plan(m, kchunks, &splits, &pc);
    float* partial = nullptr;
    if (splits > 1) {
        const size_t need = (size_t)splits * (size_t)m * 64 * sizeof(float);
        if (ws && ws_bytes >= need) partial = (float*)ws;
        else { splits = 1; pc = kchunks; }
    }
    const size_t smem = (size_t)GH_STAGES * GH_STAGE_BYTES + 1024;
    static bool attr_done = false;
    if (!attr_done) {
        if (cudaFuncSetAttribute(gemm64_h2_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem) != cudaSuccess)
            return EVAD_ERR_LAUNCH;
        attr_done = true;
    }
    dim3 grid(evad_cdiv(m, 128), splits);
    gemm64_h2_kernel<<<grid, GH_THREADS, smem, st>>>((const uint8_t*)A_h2, (const uint8_t*)W_h2, m, m_dev, kchunks, pc, partial,
                                                     *epi_h, out, ldo);
    EVAD_CHECK_LAUNCH();
    if (partial) EVAD_TRY(evad_gemm64_reduce(partial, splits, m, m_dev, epi_h, out, ldo, st));
    return EVAD_OK;
}
