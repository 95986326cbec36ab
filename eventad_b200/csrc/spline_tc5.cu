// eventad_b200/csrc/spline_tc5.cu -- SplineConv of the pooled layers (Cout = 64) as ONE kernel: tap aggregation feeding the
// 5th-generation tensor cores directly, without the aggregated operand A[n, 26*Cin] ever going to HBM.
//
// Reference path replaced per call (paths relative to /root/reference/src/dagr/model):
//   layers/spline_conv.py:39-47,64-78  message (LUT or torch_spline_conv basis) + sum aggregation + root weight `lin`
//   layers/conv.py:23-28,47-56         eval BatchNorm, skip add, activation (epilogue)
//
// Why: with csrc/spline_general.cu + csrc/gemm_tc5.cu the operand A is written and read back once per convolution --
// 1.2 GB per launch at 64 windows of 100k events, more HBM traffic than the rest of the layer together, and the GEMM
// kernel spends its issue slots splitting A into TF32 parts (profiles/r01d_summary.md).  Here
//   out_i = sum_k W_k^T (sum_{e in N(i), tap(e) = k} b_e x_src(e)) + W_root^T x_i
// is walked tap by tap: for tap k the CTA's 256 threads build the [128 nodes x Cin] slice of A in registers from a
// per-node tap-sorted entry list (one pass over the graph, shared by both convolutions of a Layer), write it as TF32
// hi / lo parts in the SWIZZLE_128B K-major layout, cp.async brings the matching pre-split weight rows, and one thread
// issues the tcgen05.mma (kind::tf32, 3xTF32: lo*hi, hi*lo, hi*hi) into a TMEM accumulator that is drained into fp32
// registers every 32 K columns, exactly as in gemm_tc5.cu.
#include "spline_common.cuh"
#include "tc5.cuh"

#define F5_THREADS 512                                        // 16 warps on the one resident CTA: 4 per scheduler
#define F5_ITEMS (T5_BM * 8 / F5_THREADS)                     // (row, 16-byte chunk) items per thread and K block: 2
#define F5_RSTEP (F5_THREADS / 8)                             // row distance between a thread's items: 64
#define F5_DCOLS (64 / (F5_THREADS / 128))                    // accumulator columns a thread drains: 16
#define F5_WSLOTS 4                                           // weight ring: {W_hi, W_lo} blocks, 16 KB each, fetched 2 stages ahead
#define F5_A2_BYTES (2 * T5_A_BYTES)                          // one A operand buffer: A_hi, A_lo = 32 KB

// ------------------------------------------------------------------------------------------------ tap lists
// Per node: the (source, basis value) entries of its in-edges, grouped by kernel tap (ascending), CSR order inside a tap;
// tptr[i*(ntap+1) + k] .. tptr[i*(ntap+1) + k + 1] is the entry range of tap k.  Entries with basis value 0 are dropped.
// Node i owns the slots [4*rowptr[i], 4*rowptr[i+1]) of `ent`.  One thread per node, two passes over its row.
__global__ void __launch_bounds__(128) tap_list_kernel(const int32_t* __restrict__ rowptr, const int32_t* __restrict__ src,
                                                       int64_t n, const int32_t* __restrict__ n_dev, EvadPseudo P,
                                                       EvadSplineMode M, int32_t* __restrict__ tptr, int2* __restrict__ ent,
                                                       uint32_t* __restrict__ status) {
    const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    const int64_t nn = n_dev ? min((int64_t)*n_dev, n) : n;
    if (i >= nn) return;
    const int ntap = M.ks * M.ks;
    const int r0 = rowptr[i], r1 = rowptr[i + 1];
    int cnt[32];
#pragma unroll
    for (int k = 0; k < 32; k++) cnt[k] = 0;
    for (int e = r0; e < r1; e++) {
        float ax, ay, w[4];
        int k[4];
        evad_pseudo(P, e, src[e], (int)i, &ax, &ay);
        evad_taps(M, ax, ay, k, w, status);
#pragma unroll
        for (int s = 0; s < 4; s++)
            if (w[s] != 0.f) cnt[k[s]]++;
    }
    int32_t* tp = tptr + i * (ntap + 1);
    int run = 4 * r0;
    for (int k = 0; k < ntap; k++) {
        tp[k] = run;
        const int c = cnt[k];
        cnt[k] = run;                                         // becomes the write cursor of tap k
        run += c;
    }
    tp[ntap] = run;
    for (int e = r0; e < r1; e++) {
        float ax, ay, w[4];
        int k[4];
        const int j = src[e];
        evad_pseudo(P, e, j, (int)i, &ax, &ay);
        evad_taps(M, ax, ay, k, w, status);
#pragma unroll
        for (int s = 0; s < 4; s++)
            if (w[s] != 0.f) ent[cnt[k[s]]++] = make_int2(j, __float_as_int(w[s]));
    }
}

extern "C" int evad_spline_tap_lists(const int32_t* rowptr, const int32_t* src, int64_t n, const int32_t* n_dev,
                                     const evad_pseudo_t* pseudo_h, const evad_spline_mode_t* mode_h, int32_t* tptr,
                                     void* entries, uint32_t* status, void* stream) {
    if (!rowptr || !src || !pseudo_h || !mode_h || !tptr || !entries || !status) return EVAD_ERR_ARG;
    if (mode_h->use_lut && (!mode_h->tab_k || !mode_h->tab_w)) return EVAD_ERR_ARG;
    if (mode_h->ks * mode_h->ks > 32) return EVAD_ERR_UNSUPPORTED;
    if (n <= 0) return EVAD_OK;
    tap_list_kernel<<<evad_cdiv(n, 128), 128, 0, (cudaStream_t)stream>>>(rowptr, src, n, n_dev, *pseudo_h, *mode_h, tptr,
                                                                          (int2*)entries, status);
    EVAD_CHECK_LAUNCH();
    return EVAD_OK;
}

// ------------------------------------------------------------------------------------------------ fused convolution
// x [n, ldx] rows zero padded to ldx (multiple of 4); K layout: tap-major, `cp` columns per tap (multiple of 8, >= ldx
// rounded up), taps 0 .. ntap-1 then the root "tap" ntap; Wt_hi / Wt_lo [64, ldw]: TF32 parts of the transposed stacked
// weight in that K layout, zero in the padding columns.  blockIdx.y selects a range of taps (split for small graphs:
// raw partial sums, reduced by evad_gemm64_reduce).
__global__ void __launch_bounds__(F5_THREADS, 1) spline_fused_tc5_kernel(
    const int32_t* __restrict__ tptr, const int2* __restrict__ ent, const float* __restrict__ x, int ldx, int64_t n,
    const int32_t* __restrict__ n_dev, int ntap, int cp, const float* __restrict__ Wt_hi, const float* __restrict__ Wt_lo,
    int64_t ldw, int taps_per_split, float* __restrict__ partial, evad_epilogue_t E, float* __restrict__ out, int ldo) {
    extern __shared__ uint8_t f5_raw[];
    __shared__ __align__(8) uint64_t s_bar[2];
    __shared__ uint32_t s_tmem;
    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const int64_t mm = n_dev ? min((int64_t)*n_dev, n) : n;
    const int64_t row0 = (int64_t)blockIdx.x * T5_BM;
    if (row0 >= mm) return;                                  // whole CTA leaves before any barrier / TMEM allocation
    const int tap_lo = blockIdx.y * taps_per_split, tap_hi = min(ntap + 1, tap_lo + taps_per_split);
    const int nb = (cp + T5_BK - 1) / T5_BK;                  // K blocks (swizzle rows) per tap
    const int nstage = (tap_hi - tap_lo) * nb;
    const uint32_t ob_base = (t5_smem(f5_raw) + 1023u) & ~1023u;
    uint8_t* ob_g = f5_raw + (ob_base - t5_smem(f5_raw));

    if (tid == 0) {
        t5_mbar_init(t5_smem(&s_bar[0]), 1);
        t5_mbar_init(t5_smem(&s_bar[1]), 1);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    if (warp == 0) {
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(t5_smem(&s_tmem)),
                     "n"(T5_TMEM_COLS)
                     : "memory");
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
    }
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
    const uint32_t tmem = s_tmem;

    // A producer: thread -> (row r = (tid >> 3) + F5_RSTEP j, 16-byte chunk c = tid & 7 of the 128-byte K block)
    const int c8 = tid & 7, rbase = tid >> 3;
    // W loader: 64 rows x 8 chunks x {hi, lo} = 1024 chunks: chunk idx = tid + F5_THREADS j -> (part, row, chunk)
    const uint32_t wring = ob_base + 2 * F5_A2_BYTES;
    auto load_w = [&](int slot, int tap, int b) {
        const int64_t kcol = (int64_t)tap * cp + b * T5_BK;
        const int kw = min(T5_BK, cp - b * T5_BK);
        const uint32_t wb = wring + slot * (2 * T5_B_BYTES);
#pragma unroll
        for (int j = 0; j < 1024 / F5_THREADS; j++) {
            const int idx = tid + j * F5_THREADS;
            const int part = idx >> 9, rr = (idx >> 3) & 63, cc = idx & 7;
            if (4 * cc < kw) {
                const float* g = (part ? Wt_lo : Wt_hi) + (int64_t)rr * ldw + kcol + 4 * cc;
                t5_cp16(wb + part * T5_B_BYTES + rr * 128 + ((cc ^ (rr & 7)) << 4), g, 16);
            }
        }
    };
    // ---- A producer, software pipelined: the gathers of stage st+1 are issued before the threads wait for stage st's
    //      barrier / drain, and the entry-list heads of the following tap are fetched a whole tap ahead, so that no L2
    //      round trip sits between two MMAs.
    // heads = per row: start and length of the (node, tap) entry range and its first two (source, basis value) entries
    // (longer ranges are rare and finished with a plain loop)
    int cp0[F5_ITEMS], ccnt[F5_ITEMS], np0[F5_ITEMS], ncnt[F5_ITEMS];                     // c*: tap of the gathers in flight, n*: the tap after it
    int2 ce0[F5_ITEMS], ce1[F5_ITEMS], ne0[F5_ITEMS], ne1[F5_ITEMS];
    float4 g0[F5_ITEMS], g1[F5_ITEMS];                                      // gathered source-row quads of the stage to finish next
    auto heads_ptr = [&](int tap, int (&p0)[F5_ITEMS], int (&cnt)[F5_ITEMS], int2 (&e0)[F5_ITEMS]) {
#pragma unroll
        for (int j = 0; j < F5_ITEMS; j++) {
            const int64_t node = row0 + rbase + F5_RSTEP * j;
            p0[j] = 0;
            cnt[j] = 0;
            if (node < mm && tap <= ntap) {
                if (tap == ntap) {                            // root weight: the node's own row, basis value 1
                    cnt[j] = 1;
                    e0[j] = make_int2((int)node, __float_as_int(1.f));
                } else {
                    p0[j] = __ldg(tptr + node * (ntap + 1) + tap);
                    cnt[j] = __ldg(tptr + node * (ntap + 1) + tap + 1) - p0[j];
                }
            }
        }
    };
    auto heads_ent = [&](int tap, const int (&p0)[F5_ITEMS], const int (&cnt)[F5_ITEMS], int2 (&e0)[F5_ITEMS], int2 (&e1)[F5_ITEMS]) {
        if (tap >= ntap) return;
#pragma unroll
        for (int j = 0; j < F5_ITEMS; j++) {
            if (cnt[j] > 0) e0[j] = __ldg(ent + p0[j]);
            if (cnt[j] > 1) e1[j] = __ldg(ent + p0[j] + 1);
        }
    };
    auto issue_gathers = [&](int b) {                         // for the c* tap
        const int col = b * T5_BK + 4 * c8;
        const bool live = col < ldx && 4 * c8 < min(T5_BK, cp - b * T5_BK);
#pragma unroll
        for (int j = 0; j < F5_ITEMS; j++) {
            g0[j] = g1[j] = make_float4(0.f, 0.f, 0.f, 0.f);
            if (live && ccnt[j] > 0) g0[j] = __ldg(reinterpret_cast<const float4*>(x + (int64_t)ce0[j].x * ldx + col));
            if (live && ccnt[j] > 1) g1[j] = __ldg(reinterpret_cast<const float4*>(x + (int64_t)ce1[j].x * ldx + col));
        }
    };
    auto finish_a = [&](int ob, int b) {                      // basis-weighted sum, TF32 split, swizzled store
        uint8_t* go = ob_g + ob * F5_A2_BYTES;
        const int col = b * T5_BK + 4 * c8;
        if (4 * c8 >= min(T5_BK, cp - b * T5_BK)) return;     // chunk beyond the tap's K extent: never read by the MMAs
        const bool live = col < ldx;
#pragma unroll
        for (int j = 0; j < F5_ITEMS; j++) {
            float4 acc = make_float4(0.f, 0.f, 0.f, 0.f);
            if (ccnt[j] > 0) {
                const float w = __int_as_float(ce0[j].y);
                acc.x = __fmul_rn(w, g0[j].x); acc.y = __fmul_rn(w, g0[j].y);
                acc.z = __fmul_rn(w, g0[j].z); acc.w = __fmul_rn(w, g0[j].w);
            }
            if (ccnt[j] > 1) {
                const float w = __int_as_float(ce1[j].y);
                acc.x = fmaf(w, g1[j].x, acc.x); acc.y = fmaf(w, g1[j].y, acc.y);
                acc.z = fmaf(w, g1[j].z, acc.z); acc.w = fmaf(w, g1[j].w, acc.w);
            }
            if (live)
                for (int e = cp0[j] + 2; e < cp0[j] + ccnt[j]; e++) {
                    const int2 en = __ldg(ent + e);
                    const float4 xv = __ldg(reinterpret_cast<const float4*>(x + (int64_t)en.x * ldx + col));
                    const float w = __int_as_float(en.y);
                    acc.x = fmaf(w, xv.x, acc.x); acc.y = fmaf(w, xv.y, acc.y);
                    acc.z = fmaf(w, xv.z, acc.z); acc.w = fmaf(w, xv.w, acc.w);
                }
            const int r = rbase + F5_RSTEP * j;
            float4 h, l;
            t5_split4(acc, &h, &l);
            const uint32_t off = r * 128 + ((c8 ^ (r & 7)) << 4);
            *reinterpret_cast<float4*>(go + off) = h;
            *reinterpret_cast<float4*>(go + off + T5_A_BYTES) = l;
        }
    };

    // warp w drains TMEM lanes 32*(w & 3) .. +31 (its hardware lane quarter), columns F5_DCOLS*(w >> 2) .. +F5_DCOLS-1
    const int quarter = warp & 3, cpart = warp >> 2;
    float racc[F5_DCOLS];
#pragma unroll
    for (int j = 0; j < F5_DCOLS; j++) racc[j] = 0.f;
    auto drain = [&](int st) {
        t5_mbar_wait(t5_smem(&s_bar[st & 1]), (uint32_t)((st >> 1) & 1));
        asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
        uint32_t r[F5_DCOLS];
        t5_ld16(tmem + ((st & 1) ? 64u : 0u) + cpart * F5_DCOLS + ((uint32_t)(quarter * 32) << 16), r);
#pragma unroll
        for (int j = 0; j < F5_DCOLS; j++) racc[j] += __uint_as_float(r[j]);
        asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    };

    // weights of stages 0 and 1; stage st+2 is requested in iteration st (ring slot last read by the MMAs of stage st-2)
    load_w(0, tap_lo, 0);
    t5_commit_group();
    if (nstage > 1) load_w(1, nb > 1 ? tap_lo : tap_lo + 1, nb > 1 ? 1 : 0);
    t5_commit_group();
    // prologue: heads of the first tap (exposed once), its first gathers, and the pointer fetch of the second tap
    heads_ptr(tap_lo, cp0, ccnt, ce0);
    heads_ent(tap_lo, cp0, ccnt, ce0, ce1);
    issue_gathers(0);
    heads_ptr(tap_lo + 1 < tap_hi ? tap_lo + 1 : ntap + 1, np0, ncnt, ne0);
    if (nb == 1) heads_ent(tap_lo + 1 < tap_hi ? tap_lo + 1 : ntap + 1, np0, ncnt, ne0, ne1);
    int tap = tap_lo, b = 0;                                  // stage st = (tap, b)
    for (int st = 0; st < nstage; st++) {
        const int ob = st & 1;
        // operand buffer ob was read by the MMAs of stage st-2; every thread waited for them in drain(st-2)
        finish_a(ob, b);
        t5_wait_group<1>();                                   // this thread's weight chunks of stage st have landed
        asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
        asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
        __syncthreads();
        if (tid == 0) {
            asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
            const uint32_t sa = ob_base + ob * F5_A2_BYTES, sw = wring + (st % F5_WSLOTS) * (2 * T5_B_BYTES);
            const uint32_t tacc = tmem + (ob ? 64u : 0u);
            const uint64_t a_hi = t5_desc(sa), a_lo = t5_desc(sa + T5_A_BYTES);
            const uint64_t b_hi = t5_desc(sw), b_lo = t5_desc(sw + T5_B_BYTES);
            const int ksteps = min(T5_BK, cp - b * T5_BK) >> 3;
            for (int k = 0; k < ksteps; k++) {               // cross terms first, then hi*hi (see gemm_tc5.cu)
                const uint64_t o = (uint64_t)(2 * k);
                t5_mma(tacc, a_lo + o, b_hi + o, k != 0);
                t5_mma(tacc, a_hi + o, b_lo + o, 1u);
            }
            for (int k = 0; k < ksteps; k++) {
                const uint64_t o = (uint64_t)(2 * k);
                t5_mma(tacc, a_hi + o, b_hi + o, 1u);
            }
            t5_mma_commit(t5_smem(&s_bar[ob]));
        }
        // ---- next stage: gathers (and, a tap ahead, entry-list heads) go out before this thread waits for anything
        if (++b == nb) {
            b = 0;
            tap++;
#pragma unroll
            for (int j = 0; j < F5_ITEMS; j++) { cp0[j] = np0[j]; ccnt[j] = ncnt[j]; ce0[j] = ne0[j]; ce1[j] = ne1[j]; }
        }
        if (st + 1 < nstage) {
            issue_gathers(b);
            if (b == 0) heads_ptr(tap + 1 < tap_hi ? tap + 1 : ntap + 1, np0, ncnt, ne0);
            if (b == (nb > 1 ? 1 : 0)) heads_ent(tap + 1 < tap_hi ? tap + 1 : ntap + 1, np0, ncnt, ne0, ne1);
        }
        if (st > 0) drain(st - 1);                            // overlaps with the MMAs just issued; frees buffer ob ^ 1
        if (st + 2 < nstage) {                                // (tap, b) is stage st+1 here; its successor is st+2
            const int b2 = b + 1 == nb ? 0 : b + 1, tap2 = b + 1 == nb ? tap + 1 : tap;
            load_w((st + 2) % F5_WSLOTS, tap2, b2);
        }
        t5_commit_group();
    }
    drain(nstage - 1);

    // ---- epilogue: thread owns output row 32*quarter + lane, columns F5_DCOLS*cpart .. +F5_DCOLS-1
    const int64_t gr = row0 + quarter * 32 + lane;
    if (gr < mm) {
#pragma unroll
        for (int j = 0; j < F5_DCOLS; j += 4) {
            const int c = cpart * F5_DCOLS + j;
            float v[4] = {racc[j], racc[j + 1], racc[j + 2], racc[j + 3]};
            if (partial) {
                *reinterpret_cast<float4*>(partial + ((int64_t)blockIdx.y * n + gr) * 64 + c) =
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
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    if (warp == 0)
        asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem), "n"(T5_TMEM_COLS) : "memory");
}

// taps per CTA so that small graphs still fill the SMs
static void f5_plan(int64_t m, int ntap1, int* splits, int* taps_per_split) {
    const int mtiles = evad_cdiv(m, T5_BM);
    int s = 1;
    if (mtiles < 148) {
        s = 148 / mtiles;
        if (s > ntap1 / 2) s = ntap1 / 2;
        if (s < 1) s = 1;
    }
    const int tps = (ntap1 + s - 1) / s;
    *taps_per_split = tps;
    *splits = (ntap1 + tps - 1) / tps;
}

extern "C" size_t evad_spline_conv64_fused_ws_bytes(int64_t n, int ntap) {
    if (n <= 0 || ntap < 1) return 256;
    int splits, tps;
    f5_plan(n, ntap + 1, &splits, &tps);
    return (splits > 1 ? (size_t)splits * (size_t)n * 64 * sizeof(float) : 0) + 256;
}

extern "C" int evad_spline_conv64_fused(const int32_t* tptr, const void* entries, const float* x, int ldx, int64_t n,
                                        const int32_t* n_dev, int ntap, int cp, const float* Wt_hi, const float* Wt_lo,
                                        int64_t ldw, const evad_epilogue_t* epi_h, float* out, int ldo, void* ws,
                                        size_t ws_bytes, void* stream) {
    if (!tptr || !entries || !x || !Wt_hi || !Wt_lo || !epi_h || !out || ntap < 1 || ntap > 32) return EVAD_ERR_ARG;
    if (ldx % 4 != 0 || cp % 8 != 0 || cp < ldx || ldw < (int64_t)(ntap + 1) * cp || ldw % 4 != 0 || ldo < 64 || ldo % 4 != 0)
        return EVAD_ERR_ARG;
    if ((((uintptr_t)x | (uintptr_t)Wt_hi | (uintptr_t)Wt_lo | (uintptr_t)out | (uintptr_t)entries) & 15) != 0) return EVAD_ERR_ARG;
    if (n <= 0) return EVAD_OK;
    cudaStream_t st = (cudaStream_t)stream;
    int splits, tps;
    f5_plan(n, ntap + 1, &splits, &tps);
    float* partial = nullptr;
    if (splits > 1) {
        const size_t need = (size_t)splits * (size_t)n * 64 * sizeof(float);
        if (ws && ws_bytes >= need) partial = (float*)ws;
        else { splits = 1; tps = ntap + 1; }
    }
    const size_t smem = (size_t)2 * F5_A2_BYTES + (size_t)F5_WSLOTS * 2 * T5_B_BYTES + 1024;
    static bool attr_done = false;
    if (!attr_done) {
        if (cudaFuncSetAttribute(spline_fused_tc5_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem) != cudaSuccess)
            return EVAD_ERR_LAUNCH;
        attr_done = true;
    }
    dim3 grid(evad_cdiv(n, T5_BM), splits);
    spline_fused_tc5_kernel<<<grid, F5_THREADS, smem, st>>>(tptr, (const int2*)entries, x, ldx, n, n_dev, ntap, cp, Wt_hi,
                                                            Wt_lo, ldw, tps, partial, *epi_h, out, ldo);
    EVAD_CHECK_LAUNCH();
    if (partial) EVAD_TRY(evad_gemm64_reduce(partial, splits, n, n_dev, epi_h, out, ldo, st));
    return EVAD_OK;
}
