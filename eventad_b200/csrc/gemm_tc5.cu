// eventad_b200/csrc/gemm_tc5.cu -- the SplineConv weight contraction on Blackwell's 5th-generation tensor cores:
//   out[m, 0:64] = epi(A[m, 0:K] @ W[K, 64])   with tcgen05.mma (kind::tf32), accumulator in tensor memory (TMEM).
//
// Reference path replaced: as csrc/gemm_tc.cu (layers/spline_conv.py:39-47,64-78 einsum / spline_weighting + `lin`,
// layers/components.py:15-22, layers/conv.py:23-28,47-56 in the epilogue).
//
// Why: profiles/r01c_summary.md -- on the legacy mma.sync path the 3xTF32 contraction is tensor-pipe bound
// (HMMA.1688.TF32 ~410 MAC/clk/SM); tcgen05 kind::tf32 has ~4.6x that.  fp32 accuracy is kept the same way: both
// operands are split into TF32 parts x = hi + lo (hi = upper 19 bits, lo = exact remainder) and every 128x64x8 step is
// issued three times, (A_lo, W_hi), (A_hi, W_lo), (A_hi, W_hi), into the same TMEM accumulator.
//
// One CTA = 128 threads = one 128 x 64 output tile; K is walked in slabs of 32 floats (= one 128-byte swizzle row):
//   cp.async (16-byte chunks, written in the SWIZZLE_128B K-major layout the UMMA descriptors expect) ->
//   split pass (each thread rewrites its chunks as hi in place and lo into a second tile) -> fence.proxy.async ->
//   one thread issues 12 tcgen05.mma and commits them to the slab's mbarrier -> (two slabs in flight) ->
//   tcgen05.ld of the accumulator rows (one row per thread), BatchNorm / skip add / ReLU epilogue, or raw split-K
//   partials that gemm_tc.cu's deterministic reduce kernel sums.
#include "tc5.cuh"

#define T5_THREADS 256
#define T5_RAW_BYTES (T5_A_BYTES + T5_B_BYTES)               // raw fp32 slab: 24 KB
#define T5_RAW_STAGES 4

// A [m, lda] row-major; Wt [64, ldw] = W transposed (row n holds the K weights of output n, zero padded to ldw).
// Shared memory: a T5_RAW_STAGES-deep ring of raw fp32 slabs filled by cp.async (enough bytes in flight to cover the
// HBM / L2 latency) and two operand buffers {A_hi, A_lo, W_hi, W_lo} in the UMMA layout, so that the MMAs of slab kt run
// while the threads split slab kt+1 and drain the accumulator of slab kt-1.
__global__ void __launch_bounds__(T5_THREADS, 1) gemm64_tcgen05_kernel(
    const float* __restrict__ A, int64_t lda, const float* __restrict__ Wt, int64_t ldw, int64_t m,
    const int32_t* __restrict__ m_dev, int K_total, int k_chunk, float* __restrict__ partial, evad_epilogue_t E,
    float* __restrict__ out, int ldo) {
    extern __shared__ uint8_t t5_raw[];
    __shared__ __align__(8) uint64_t s_bar[2];
    __shared__ uint32_t s_tmem;
    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const int64_t mm = m_dev ? min((int64_t)*m_dev, m) : m;
    const int64_t row0 = (int64_t)blockIdx.x * T5_BM;
    if (row0 >= mm) return;                                  // whole CTA leaves before any barrier / TMEM allocation
    const int kb0 = blockIdx.y * k_chunk;
    const int K = min(K_total, kb0 + k_chunk);
    const int nk = (K - kb0 + T5_BK - 1) / T5_BK;
    // operand buffers first, 1024-byte aligned (the swizzle pattern is a function of address bits 4..9), then the raw ring
    const uint32_t ob_base = (t5_smem(t5_raw) + 1023u) & ~1023u;
    uint8_t* ob_g = t5_raw + (ob_base - t5_smem(t5_raw));
    const uint32_t raw_base = ob_base + 2 * T5_OB_BYTES;
    uint8_t* raw_g = ob_g + 2 * T5_OB_BYTES;

    if (tid == 0) {
        t5_mbar_init(t5_smem(&s_bar[0]), 1);
        t5_mbar_init(t5_smem(&s_bar[1]), 1);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    if (warp == 0) {                                          // one warp allocates the accumulator columns
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(t5_smem(&s_tmem)),
                     "n"(T5_TMEM_COLS)
                     : "memory");
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
    }
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
    const uint32_t tmem = s_tmem;

    // raw slab: A rows then W rows, 128 bytes each, plain order; thread t owns chunks t, t + 256, ... in load AND split.
    // Everything that does not depend on the slab is resolved once: source row pointers, validity, swizzled targets.
    constexpr int CH = (T5_BM + T5_BN) * 8 / T5_THREADS;     // 16-byte chunks per thread and slab (6)
    const float* rowp[CH];                                   // global address of the chunk at k = 0 (nullptr: zero fill)
    uint32_t dst_off[CH], lo_off[CH];                        // byte offsets of the hi chunk in an operand buffer / hi -> lo
    int col4[CH];                                            // first k column of the chunk inside a slab
#pragma unroll
    for (int j = 0; j < CH; j++) {
        const int idx = tid + j * T5_THREADS;
        const int r = idx >> 3, c = idx & 7;
        col4[j] = 4 * c;
        if (r < T5_BM) {
            const int64_t gr = row0 + r;
            rowp[j] = gr < mm ? A + gr * lda + 4 * c : nullptr;
            dst_off[j] = r * 128 + ((c ^ (r & 7)) << 4);
            lo_off[j] = T5_A_BYTES;
        } else {
            const int rb = r - T5_BM;
            rowp[j] = Wt + (int64_t)rb * ldw + 4 * c;
            dst_off[j] = 2 * T5_A_BYTES + rb * 128 + ((c ^ (rb & 7)) << 4);
            lo_off[j] = T5_B_BYTES;
        }
    }
    auto load_stage = [&](int slot, int k0) {
        const uint32_t sr = raw_base + slot * T5_RAW_BYTES + tid * 16;
        if (k0 + T5_BK <= K) {
#pragma unroll
            for (int j = 0; j < CH; j++)
                t5_cp16(sr + j * (T5_THREADS * 16), rowp[j] ? rowp[j] + k0 : A, rowp[j] ? 16 : 0);
        } else {                                              // K tail: clip the chunk, zero fill the rest
#pragma unroll
            for (int j = 0; j < CH; j++) {
                int bytes = rowp[j] ? min(16, (K - k0 - col4[j]) * 4) : 0;
                bytes = max(bytes, 0);
                t5_cp16(sr + j * (T5_THREADS * 16), bytes > 0 ? rowp[j] + k0 : A, bytes);
            }
        }
    };
    // split: raw chunk -> hi and lo chunks at the swizzled position r*128 + ((c ^ (r & 7)) << 4) of the operand tiles
    auto split_stage = [&](int slot, int ob) {
        const uint8_t* gr_ = raw_g + slot * T5_RAW_BYTES + tid * 16;
        uint8_t* go = ob_g + ob * T5_OB_BYTES;
        float4 v[CH];
#pragma unroll
        for (int j = 0; j < CH; j++) v[j] = *reinterpret_cast<const float4*>(gr_ + j * (T5_THREADS * 16));
#pragma unroll
        for (int j = 0; j < CH; j++) {
            float4 h, l;
            h.x = __uint_as_float(__float_as_uint(v[j].x) & 0xffffe000u); l.x = __fsub_rn(v[j].x, h.x);
            h.y = __uint_as_float(__float_as_uint(v[j].y) & 0xffffe000u); l.y = __fsub_rn(v[j].y, h.y);
            h.z = __uint_as_float(__float_as_uint(v[j].z) & 0xffffe000u); l.z = __fsub_rn(v[j].z, h.z);
            h.w = __uint_as_float(__float_as_uint(v[j].w) & 0xffffe000u); l.w = __fsub_rn(v[j].w, h.w);
            *reinterpret_cast<float4*>(go + dst_off[j]) = h;
            *reinterpret_cast<float4*>(go + dst_off[j] + lo_off[j]) = l;
        }
    };

    // The tensor core truncates when it aligns the fp32 accumulator, a bias that grows with the number of accumulating
    // MMAs and with the magnitude of the accumulator.  So a TMEM buffer only collects one slab (12 MMAs) from zero; it
    // is drained into registers with round-to-nearest adds while the next slab runs in the other buffer.
    // Warp w reads TMEM lanes 32*(w & 3) .. +31 (its hardware lane quarter), columns 32*(w >> 2) .. +31.
    const int quarter = warp & 3, chalf = warp >> 2;
    float racc[32];
#pragma unroll
    for (int j = 0; j < 32; j++) racc[j] = 0.f;
    auto drain = [&](int slab) {
        t5_mbar_wait(t5_smem(&s_bar[slab & 1]), (uint32_t)((slab >> 1) & 1));
        asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
        const uint32_t taddr = tmem + ((slab & 1) ? 64u : 0u) + chalf * 32 + ((uint32_t)(quarter * 32) << 16);
        uint32_t r[32];
        asm volatile(
            "tcgen05.ld.sync.aligned.32x32b.x32.b32 "
            "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, "
            "%16, %17, %18, %19, %20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31}, [%32];"
            : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]),
              "=r"(r[8]), "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15]),
              "=r"(r[16]), "=r"(r[17]), "=r"(r[18]), "=r"(r[19]), "=r"(r[20]), "=r"(r[21]), "=r"(r[22]), "=r"(r[23]),
              "=r"(r[24]), "=r"(r[25]), "=r"(r[26]), "=r"(r[27]), "=r"(r[28]), "=r"(r[29]), "=r"(r[30]), "=r"(r[31])
            : "r"(taddr)
            : "memory");
        asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
#pragma unroll
        for (int j = 0; j < 32; j++) racc[j] += __uint_as_float(r[j]);
        asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    };

#pragma unroll
    for (int s = 0; s < T5_RAW_STAGES - 1; s++) {
        if (s < nk) load_stage(s, kb0 + s * T5_BK);
        t5_commit_group();
    }
    for (int kt = 0; kt < nk; kt++) {
        {   // slab kt + STAGES-1 goes into the ring slot this thread finished splitting one iteration ago
            const int nxt = kt + T5_RAW_STAGES - 1;
            if (nxt < nk) load_stage(nxt % T5_RAW_STAGES, kb0 + nxt * T5_BK);
            t5_commit_group();
        }
        t5_wait_group<T5_RAW_STAGES - 1>();                   // this thread's chunks of slab kt have landed
        const int ob = kt & 1;
        // operand buffer ob was read by the MMAs of slab kt-2; every thread waited for them in drain(kt-2)
        split_stage(kt % T5_RAW_STAGES, ob);
        asm volatile("fence.proxy.async.shared::cta;" ::: "memory");   // generic-proxy writes -> visible to the tensor core
        asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
        __syncthreads();
        if (tid == 0) {
            asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
            const uint32_t sa = ob_base + ob * T5_OB_BYTES;
            const uint32_t tacc = tmem + (ob ? 64u : 0u);
            const uint64_t a_hi = t5_desc(sa), a_lo = t5_desc(sa + T5_A_BYTES);
            const uint64_t b_hi = t5_desc(sa + 2 * T5_A_BYTES), b_lo = t5_desc(sa + 2 * T5_A_BYTES + T5_B_BYTES);
            // the two cross terms first (small, their truncation does not matter), then the four hi*hi steps
#pragma unroll
            for (int k = 0; k < T5_BK / 8; k++) {            // 8 floats = 32 bytes = 2 descriptor units per K step
                const uint64_t o = (uint64_t)(2 * k);
                t5_mma(tacc, a_lo + o, b_hi + o, k != 0);
                t5_mma(tacc, a_hi + o, b_lo + o, 1u);
            }
#pragma unroll
            for (int k = 0; k < T5_BK / 8; k++) {
                const uint64_t o = (uint64_t)(2 * k);
                t5_mma(tacc, a_hi + o, b_hi + o, 1u);
            }
            t5_mma_commit(t5_smem(&s_bar[ob]));               // arrives when these (and all earlier) MMAs are done
        }
        if (kt > 0) drain(kt - 1);                            // overlaps with the MMAs just issued
    }
    drain(nk - 1);

    // ---- epilogue: thread owns output row 32*quarter + lane, columns 32*chalf .. +31
    const int64_t gr = row0 + quarter * 32 + lane;
    if (gr < mm) {
#pragma unroll
        for (int j = 0; j < 32; j += 4) {
            const int c = chalf * 32 + j;
            float v[4] = {racc[j], racc[j + 1], racc[j + 2], racc[j + 3]};
            if (partial) {
                *reinterpret_cast<float4*>(partial + ((int64_t)blockIdx.y * m + gr) * 64 + c) =
                    make_float4(v[0], v[1], v[2], v[3]);
                continue;
            }
#pragma unroll
            for (int b = 0; b < 4; b++) {
                if (E.scale) v[b] = __fmul_rn(v[b], E.scale[c + b]);
                if (E.shift) v[b] = __fadd_rn(v[b], E.shift[c + b]);
                if (E.addend) v[b] = __fadd_rn(v[b], E.addend[gr * E.ld_add + c + b]);
                if (E.relu) v[b] = fmaxf(v[b], 0.f);
            }
            *reinterpret_cast<float4*>(out + gr * ldo + c) = make_float4(v[0], v[1], v[2], v[3]);
        }
    }
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    if (warp == 0)
        asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem), "n"(T5_TMEM_COLS) : "memory");
}

// K slices so that every SM gets a CTA (one CTA per SM: 193 KB of shared memory)
static void t5_plan(int64_t m, int k, int* splits, int* k_chunk) {
    const int mtiles = evad_cdiv(m, T5_BM);
    int s = 1;
    if (mtiles < 148) {
        s = 148 / mtiles;
        const int max_s = (k + 4 * T5_BK - 1) / (4 * T5_BK);
        if (s > max_s) s = max_s;
        if (s > 32) s = 32;
        if (s < 1) s = 1;
    }
    const int kc = ((k + s - 1) / s + T5_BK - 1) / T5_BK * T5_BK;
    *k_chunk = kc;
    *splits = (k + kc - 1) / kc;
}

extern "C" size_t evad_gemm64_tc5_ws_bytes(int64_t m, int k) {
    if (m <= 0 || k < 1) return 256;
    int splits, kc;
    t5_plan(m, k, &splits, &kc);
    return (splits > 1 ? (size_t)splits * (size_t)m * 64 * sizeof(float) : 0) + 256;
}

extern "C" int evad_gemm64_tc5(const float* A, int64_t lda, const float* Wt, int64_t ldw, int64_t m, const int32_t* m_dev,
                               int k, const evad_epilogue_t* epi_h, float* out, int ldo, void* ws, size_t ws_bytes,
                               void* stream) {
    if (!A || !Wt || !epi_h || !out || k < 1 || lda < k || ldw < k || ldo < 64) return EVAD_ERR_ARG;
    if (lda % 4 != 0 || ldw % 4 != 0 || ldo % 4 != 0 || (((uintptr_t)A | (uintptr_t)Wt | (uintptr_t)out) & 15))
        return EVAD_ERR_UNSUPPORTED;
    if (m <= 0) return EVAD_OK;
    cudaStream_t st = (cudaStream_t)stream;
    int splits, kc;
    t5_plan(m, k, &splits, &kc);
    float* partial = nullptr;
    if (splits > 1) {
        const size_t need = (size_t)splits * (size_t)m * 64 * sizeof(float);
        if (ws && ws_bytes >= need) partial = (float*)ws;
        else { splits = 1; kc = (k + T5_BK - 1) / T5_BK * T5_BK; }
    }
    const size_t smem = (size_t)2 * T5_OB_BYTES + (size_t)T5_RAW_STAGES * T5_RAW_BYTES + 1024;
    static bool attr_done = false;
    if (!attr_done) {
        if (cudaFuncSetAttribute(gemm64_tcgen05_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem) != cudaSuccess)
            return EVAD_ERR_LAUNCH;
        attr_done = true;
    }
    dim3 grid(evad_cdiv(m, T5_BM), splits);
    gemm64_tcgen05_kernel<<<grid, T5_THREADS, smem, st>>>(A, lda, Wt, ldw, m, m_dev, k, kc, partial, *epi_h, out, ldo);
    EVAD_CHECK_LAUNCH();
    if (partial) EVAD_TRY(evad_gemm64_reduce(partial, splits, m, m_dev, epi_h, out, ldo, st));
    return EVAD_OK;
}
