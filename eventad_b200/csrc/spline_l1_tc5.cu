// eventad_b200/csrc/spline_l1_tc5.cu -- full-resolution SplineConv (layer 1: 19->16 and 16->16 on the event graph) with the
// weight contraction on Blackwell's tcgen05 tensor cores (UMMA, accumulator in tensor memory).
//
// Reference path replaced per call (paths relative to /root/reference/src/dagr/model): as csrc/spline_l1.cu --
//   layers/spline_conv.py:39-47 (message_lut), PyG propagate sum aggregation, spline_conv.py:72-76 (root weight),
//   layers/conv.py:23-28,47-56 (eval BatchNorm, activation, skip Linear + BatchNorm).
//
// Same re-association as spline_l1_mma_kernel:  out_i = W_cat^T P_i,  P_i[slot] = sum_e tapw[class_e][slot] * x_src(e)
// (15 tap slots + the node's own row as slot 15).  What changes against the mma.sync kernels:
//   * the [tile x 16*Cin] @ [16*Cin x 16] contraction leaves mma.sync: a 128-node P tile is written ONCE to shared
//     memory in the K-major SWIZZLE_128B operand layout and contracted by tcgen05.mma (M 128, N 16, K 16 per
//     instruction) issued by one thread; the weights sit in shared memory behind UMMA descriptors for the life of the
//     persistent CTA.  No weight-fragment or A-fragment LDS, no partial-sum tiles.
//   * fp32 accuracy without an 8-byte operand: P is split into two fp16 parts, P = p1 + p2 * 2^-11 (22 significant bits;
//     the weights the same, W * s_w = w1 + w2 * 2^-11, split and block-scaled once on the host), so the operand stays at
//     4 bytes per element = 160 KB per tile and three kind::f16 products (p1 w1 | p1 w2 + p2 w1) give 2^-22 relative
//     accuracy; the leading product is accumulated in 64-column K chunks in separate TMEM accumulators that are added
//     in fp32 registers (the tensor core truncates when it accumulates).
//   * epilogue from tensor memory (tcgen05.ld, one row per thread): BatchNorm, skip add, ReLU, 64-byte row store.
// The per-node tap aggregation is spline_l1_h16_kernel's (csrc/spline_l1_h16.cu): features arrive as h2 words, 9 / 6
// mma.sync m16n8k16 per node, two accumulators (hi x hi | hi x lo + lo x hi) that are merged while P is split.
// Measured (profiles/r02_summary.md): equal to spline_l1_h16_kernel at Cin 16, 30 % slower at Cin 20: opt-in
// (EVAD_L1_TC5=1).
#include <cuda_fp16.h>

#include "spline_common.cuh"
#include "tc5.cuh"

#define L1T_ROWS 128                  // UMMA tile = nodes per contraction
#define L1T_SUB 64                    // nodes per producer round: 16 warps x 4 nodes
#define L1T_WARPS 16                  // producer warps (aggregation, operand conversion, epilogue)
#define L1T_THREADS (32 * L1T_WARPS)
#define L1T_SLOTS 16
#define L1T_TW_LD 24                  // words per class row of the tap table: 8 x (hi pair, lo pair) of (slot g, slot g+8) + pad
#define L1T_CHUNK 16384               // bytes of one K chunk of the P tile: 128 rows x 64 halves
#define L1T_WCHUNK 8192               // bytes of one K chunk of the weight operand: 64 rows [w2 | w1 | s1 | s2] x 64 halves
#define L1T_ACC_COLS 256              // per K chunk ch < 3: 32 columns [p1 w2 | p1 w1]; ch 3, 4: 64 [p1 w2 | p1 w1 | p1 s1 | p1 s2];
                                      // then 32 columns [p2 w1 | p2 s1] accumulated over all chunks
#define L1T_TMEM_COLS 512             // two accumulator buffers: the MMAs of tile T run while tile T-1 is drained
#define L1T_COL_P2 224                // first column of the p2 region

// instruction descriptor: D = F32 (1 << 4), A = B = F16 (0), both K-major, N >> 3 at bit 17, M = 128 (M >> 4 at bit 24)
#define L1T_IDESC(N) ((1u << 4) | (((uint32_t)(N) >> 3) << 17) | ((128u >> 4) << 24))

template <int CP>
struct L1TCfg {
    static constexpr int NCH = CP > 16 ? 5 : 4;              // K chunks of 64 halves: 256 (+ 64 for channels 16..19)
    static constexpr int NT = (CP + 7) / 8;                  // channel tiles of the aggregation MMA
    static constexpr int P_BYTES = 2 * NCH * L1T_CHUNK;      // p1 chunks, then p2 chunks
    static constexpr int W_BYTES = NCH * L1T_WCHUNK;         // per K chunk 64 operand rows: w2, w1, s1, s2 (16 each)
    static size_t smem_bytes(int ncls) {
        return 1024 + (size_t)P_BYTES + W_BYTES + (size_t)(ncls + 1) * L1T_TW_LD * 4 + 4 * L1T_ROWS * 4 + 64 * 4 + 64;
    }
};

struct L1THead {                      // what a lane keeps of one node's row: id, length, its 4 sources and their classes
    int i, deg;
    uint32_t s[4];                    // sources of row positions t, t+4, t+8, t+12, clamped to valid rows
    uint32_t c;                       // their class bytes; positions beyond the row length hold the all-zero class
};

__device__ __forceinline__ void l1t_mma_sync(float (&d)[4], const uint32_t (&a)[4], uint32_t b0, uint32_t b1) {
    asm("mma.sync.aligned.m16n8k16.row.col.f32.f16.f16.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
        : "+f"(d[0]), "+f"(d[1]), "+f"(d[2]), "+f"(d[3])
        : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(b0), "r"(b1));
}

__device__ __forceinline__ void l1t_umma(uint32_t tmem_d, uint64_t adesc, uint64_t bdesc, uint32_t idesc, uint32_t accumulate) {
    asm volatile(
        "{\n\t"
        ".reg .pred p;\n\t"
        "setp.ne.b32 p, %4, 0;\n\t"
        "tcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n\t"
        "}\n" ::"r"(tmem_d), "l"(adesc), "l"(bdesc), "r"(idesc), "r"(accumulate)
        : "memory");
}
__device__ __forceinline__ void l1t_ldtm4(uint32_t taddr, uint32_t (&r)[4]) {
    asm volatile("tcgen05.ld.sync.aligned.32x32b.x4.b32 {%0, %1, %2, %3}, [%4];"
                 : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]) : "r"(taddr) : "memory");
}

// two fp32 values v + e * 2^-11 (the two accumulators of the aggregation) -> packed fp16 leading parts and packed fp16
// remainders * 2^11
__device__ __forceinline__ void l1t_split2(float v0, float e0, float v1, float e1, uint32_t* h1, uint32_t* h2) {
    const __half2 p1 = __floats2half2_rn(v0, v1);
    const float2 f1 = __half22float2(p1);
    const __half2 p2 = __floats2half2_rn(__fmaf_rn(__fsub_rn(v0, f1.x), 2048.f, e0), __fmaf_rn(__fsub_rn(v1, f1.y), 2048.f, e1));
    *h1 = *reinterpret_cast<const uint32_t*>(&p1);
    *h2 = *reinterpret_cast<const uint32_t*>(&p2);
}
#define L1T_LO(a, b) __byte_perm((a), (b), 0x5410)            // (low half of a, low half of b)
#define L1T_HI(a, b) __byte_perm((a), (b), 0x7632)            // (high half of a, high half of b)

// Operand layout (K index of the contraction -> (slot, channel)), chosen so that a lane's D fragments of the aggregation
// are one 128-bit store per part:
//   K <  256: unit u = K >> 3 (16 bytes), u = 4 g + t; halves 0..3 = slot g, channels 4t..4t+3; halves 4..7 = slot g + 8
//   K >= 256: q = K - 256 = 8 g + 4 t + 2 h + c  (t < 2): slot g + 8 h, channel 16 + 2 t + c
// The host packs the weights in the same order (eventad_b200/ops.py: l1_pack16).
//
// Weight operand in shared memory, per K chunk 64 rows of 128 bytes: rows 0..15 = w2, 16..31 = w1, 32..47 = s1 (skip
// Linear, leading part), 48..63 = s2.  One MMA p1 x rows [0, 32) (or [0, 64) in the chunks that hold the root slot when
// there is a skip Linear) yields [p1 w2 | p1 w1 | p1 s1 | p1 s2] in adjacent TMEM columns, one MMA p2 x rows [16, 32) (or
// [16, 48)) yields [p2 w1 | p2 s1]: the P tile is read twice per K step instead of three (five) times.
//
// 16 warps, all of them producers (aggregation -> operand conversion -> P tile) and drainers (epilogue, one tile behind).
// The tcgen05.mma of tile T are issued by lane 0 of warp T mod 16 after a named barrier on which the other warps only
// ARRIVE: no warp waits for the slowest one except that tile's issuer, and the issuing cost rotates.
template <int CP>
__global__ void __launch_bounds__(L1T_THREADS, 1) spline_l1_tc5_kernel(
    const int32_t* __restrict__ count, const int32_t* __restrict__ src, const uint8_t* __restrict__ cls,
    const int32_t* __restrict__ perm, int64_t n, const uint32_t* __restrict__ x, int ldx, const float* __restrict__ tapw,
    int ncls, const uint16_t* __restrict__ W16, float inv_sw, const uint16_t* __restrict__ S16, float inv_ss,
    evad_epilogue_t E, float* __restrict__ out, int ldo, int out_h2, evad_epilogue_t ES, float* __restrict__ skip_out,
    int ld_skip, uint32_t* __restrict__ status) {
    using C = L1TCfg<CP>;
    constexpr int NT = C::NT, NCH = C::NCH;
    extern __shared__ uint8_t l1t_smem[];
    __shared__ __align__(8) uint64_t s_bar[2];
    __shared__ uint32_t s_tmem;
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int g = lane >> 2, t = lane & 3;
    const bool has_skip = S16 != nullptr;

    const uint32_t p_base = (t5_smem(l1t_smem) + 1023u) & ~1023u;      // SWIZZLE_128B operands: 1024-byte aligned
    uint8_t* p_g = l1t_smem + (p_base - t5_smem(l1t_smem));
    const uint32_t w_base = p_base + C::P_BYTES;
    uint8_t* w_g = p_g + C::P_BYTES;
    uint32_t* tw = reinterpret_cast<uint32_t*>(w_g + C::W_BYTES);
    int* ids = reinterpret_cast<int*>(tw + (ncls + 1) * L1T_TW_LD);    // [4 tiles in flight][128] node ids of the tile rows
    float* epc = reinterpret_cast<float*>(ids + 4 * L1T_ROWS);         // epilogue scale / shift [4][16]

    // ---------------------------------------------------------------- one-time staging
    if (tid == 0) {
        t5_mbar_init(t5_smem(&s_bar[0]), 1);
        t5_mbar_init(t5_smem(&s_bar[1]), 1);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    if (warp == 0) {
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(t5_smem(&s_tmem)),
                     "n"(L1T_TMEM_COLS)
                     : "memory");
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
    }
    if (tid < 64) {                                          // epilogue constants: scale (1 when absent), shift (0 when absent)
        const int c = tid & 15, w = tid >> 4;
        const float* pq = w == 0 ? E.scale : w == 1 ? E.shift : w == 2 ? ES.scale : ES.shift;
        epc[tid] = pq ? pq[c] : (w & 1) ? 0.f : 1.f;
    }
    // weights: global W16 [part][16 rows][K halves] / S16 [part][16 rows][128 halves] -> per K chunk 64 operand rows of
    // 128 bytes (w2, w1, s1, s2), 16-byte units swizzled by (row & 7)
    for (int idx = tid; idx < NCH * 64 * 8; idx += L1T_THREADS) {
        const int u = idx & 7, orow = (idx >> 3) & 63, ch = idx >> 9;
        const int blk = orow >> 4, r = orow & 15;              // 0: w2, 1: w1, 2: s1, 3: s2
        uint4 v = make_uint4(0, 0, 0, 0);
        if (blk < 2) {
            const int part = blk == 0 ? 1 : 0;
            v = __ldg(reinterpret_cast<const uint4*>(W16 + ((size_t)(part * 16 + r) * (NCH * 64) + ch * 64 + u * 8)));
        } else if (has_skip && ch >= 3) {
            const int part = blk - 2;
            v = __ldg(reinterpret_cast<const uint4*>(S16 + ((size_t)(part * 16 + r) * 128 + (ch - 3) * 64 + u * 8)));
        }
        *reinterpret_cast<uint4*>(w_g + ch * L1T_WCHUNK + orow * 128 + ((u ^ (orow & 7)) << 4)) = v;
    }
    // tap table as h2 words: row c, lane group g: word 2g = (hi[c][g], hi[c][g+8]), word 2g+1 = (lo[c][g], lo[c][g+8]);
    // row ncls = zeros
    for (int idx = tid; idx < (ncls + 1) * 8; idx += L1T_THREADS) {
        const int r = idx >> 3, k = idx & 7;
        uint32_t a = 0u, b = 0u;
        if (r < ncls) {
            a = evad_h2_split(__ldg(tapw + r * L1T_SLOTS + k));
            b = evad_h2_split(__ldg(tapw + r * L1T_SLOTS + k + 8));
        }
        *reinterpret_cast<uint2*>(tw + r * L1T_TW_LD + 2 * k) = make_uint2(L1T_LO(a, b), L1T_HI(a, b));
    }
    for (int idx = tid; idx < C::P_BYTES / 16; idx += L1T_THREADS)     // rows of a partial last tile read as zeros
        reinterpret_cast<uint4*>(p_g)[idx] = make_uint4(0, 0, 0, 0);
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
    const uint32_t tmem = s_tmem;

    const int64_t nsub = (n + L1T_SUB - 1) / L1T_SUB;        // 64-node sub-tiles; UMMA tile T = sub-tiles 2T, 2T + 1
    const int64_t ntiles = (nsub + 1) / 2;
    // q-th sub-tile of this CTA (q = 0, 1, 2, ...) -> global sub-tile index, or -1 past the end
    auto subtile = [&](int64_t q) -> int64_t {
        const int64_t T = (int64_t)blockIdx.x + (q >> 1) * gridDim.x;
        const int64_t s = 2 * T + (q & 1);
        return (T < ntiles && s < nsub) ? s : -1;
    };

    {
        // ================================================================ all 16 warps produce; warp (tile number mod 16) also issues
        // the tile's tcgen05.mma once every warp has arrived on the "tile stored" named barrier -- the other warps only
        // arrive and go on (epilogue of the previous tile, next tile's aggregation): nobody waits for the slowest warp
        // except the issuer of that one tile, and the issuing cost rotates over the warps.
        const uint32_t zero4 = (uint32_t)ncls * 0x01010101u; // four copies of the all-zero class
        const uint32_t nmax = (uint32_t)(n - 1);
        // lane bases of the row gathers; an address is base + row * (4 ldx) as one 32 x 32 + 64 bit multiply-add
        const char* xg = reinterpret_cast<const char*>(x + 2 * g);
        const char* xg16 = reinterpret_cast<const char*>(x + 16 + g);
        const char* xt = reinterpret_cast<const char*>(x + 4 * t);
        const char* xt16 = reinterpret_cast<const char*>(x + 16 + 2 * t);
        const uint32_t ldxb = 4u * (uint32_t)ldx;
        const uint32_t* twg = tw + 2 * g;
        L1THead hd[4];
        uint4 craw[4];
        int inext[4];
        auto load_ids = [&](int64_t sub) {
#pragma unroll
            for (int k = 0; k < 4; k++) {
                const int64_t ip = sub * L1T_SUB + warp * 4 + k;
                inext[k] = -1;
                if (sub >= 0 && ip < n) inext[k] = perm ? perm[ip] : (int)ip;
            }
        };
        auto load_head = [&](int k) {                        // row head of node inext[k]: length, class bytes, 4 sources
            const int i = inext[k];
            hd[k].i = i;
            hd[k].deg = 0;
            craw[k] = make_uint4(0, 0, 0, 0);
#pragma unroll
            for (int j = 0; j < 4; j++) hd[k].s[j] = 0u;
            if (i >= 0) {
                hd[k].deg = __ldg(count + i);
                craw[k] = __ldg(reinterpret_cast<const uint4*>(cls) + i);
                const uint32_t* srow = reinterpret_cast<const uint32_t*>(
                    reinterpret_cast<const char*>(src + t) + (uint64_t)(uint32_t)i * 64u);
#pragma unroll
                for (int j = 0; j < 4; j++) hd[k].s[j] = __ldg(srow + 4 * j);
            }
        };
        auto pack_class = [&](int k) {
            const uint32_t c = __byte_perm(__byte_perm(craw[k].x, craw[k].y, t | ((4 + t) << 4)),
                                           __byte_perm(craw[k].z, craw[k].w, t | ((4 + t) << 4)), 0x5410);
            // positions t + 4j < deg  <=>  j < nv; slots beyond the row hold garbage: all-zero class, any valid source row
            const int nv = min(4, (hd[k].deg - t + 3) >> 2);
            const uint32_t live = __funnelshift_rc(0xffffffffu, 0u, 32 - 8 * nv);
            hd[k].c = (c & live) | (zero4 & ~live);
#pragma unroll
            for (int j = 0; j < 4; j++) hd[k].s[j] = min(hd[k].s[j], nmax);
        };
        uint2 xa[2][4];                                      // [buffer][row position j]: h2 words of channels 2g, 2g+1
        uint32_t xb[2][4];                                   // channel 16 + g (Cin 20)
        uint4 xr[2];                                         // the node's own row (lanes g == 7): channels 4t..4t+3
        uint2 xr2[2];                                        //   and 16+2t, 17+2t
        auto gather = [&](int k, int buf) {
#pragma unroll
            for (int j = 0; j < 4; j++) {
                const uint64_t off = (uint64_t)hd[k].s[j] * ldxb;
                xa[buf][j] = __ldg(reinterpret_cast<const uint2*>(xg + off));
                if (CP > 16) xb[buf][j] = g < 4 ? __ldg(reinterpret_cast<const uint32_t*>(xg16 + off)) : 0u;
            }
            xr[buf] = make_uint4(0u, 0u, 0u, 0u);
            xr2[buf] = make_uint2(0u, 0u);
            if (g == 7 && hd[k].i >= 0) {
                const uint64_t off = (uint64_t)(uint32_t)hd[k].i * ldxb;
                xr[buf] = __ldg(reinterpret_cast<const uint4*>(xt + off));
                if (CP > 16 && t < 2) xr2[buf] = __ldg(reinterpret_cast<const uint2*>(xt16 + off));
            }
        };
        // Epilogue of UMMA tile number `tn` of this CTA (accumulator buffer tn & 1, ids / rsc slot tn & 3): every producer
        // warp drains its TMEM lane quarter (warp & 3), four of the 16 output columns (warp >> 2): thread = (row, quad).
        auto epilogue = [&](int tn) {
            t5_mbar_wait(t5_smem(&s_bar[tn & 1]), (uint32_t)((tn >> 1) & 1));
            asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
            const int quarter = warp & 3, cq = warp >> 2;
            const int r = quarter * 32 + lane;
            const int io = ids[(tn & 3) * L1T_ROWS + r];
            const uint32_t ta = tmem + (uint32_t)((tn & 1) * L1T_ACC_COLS + 4 * cq) + ((uint32_t)(quarter * 32) << 16);
            float4 ad = make_float4(0.f, 0.f, 0.f, 0.f);
            if (io >= 0 && E.addend) ad = __ldg(reinterpret_cast<const float4*>(E.addend + (int64_t)io * E.ld_add) + cq);
            // the leading products of the K chunks are added here, in fp32 round-to-nearest (the tensor core truncates)
            float acc[4] = {0.f, 0.f, 0.f, 0.f}, cor[4], sacc[4] = {0.f, 0.f, 0.f, 0.f}, scor[4];
            {
                uint32_t cb[4], sb[4] = {0u, 0u, 0u, 0u};
                l1t_ldtm4(ta + L1T_COL_P2, cb);               // p2 w1
                if (has_skip) l1t_ldtm4(ta + L1T_COL_P2 + 16u, sb);   // p2 s1
                asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
#pragma unroll
                for (int b = 0; b < 4; b++) { cor[b] = __uint_as_float(cb[b]); scor[b] = __uint_as_float(sb[b]); }
            }
#pragma unroll
            for (int ch = 0; ch < NCH; ch++) {
                const uint32_t col = ch < 3 ? 32u * ch : 96u + 64u * (ch - 3);
                uint32_t a2[4], a1[4], s0[4] = {0u, 0u, 0u, 0u}, s1[4] = {0u, 0u, 0u, 0u};
                l1t_ldtm4(ta + col, a2);                      // p1 w2
                l1t_ldtm4(ta + col + 16u, a1);                // p1 w1
                if (has_skip && ch >= 3) {
                    l1t_ldtm4(ta + col + 32u, s0);            // p1 s1
                    l1t_ldtm4(ta + col + 48u, s1);            // p1 s2
                }
                asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
#pragma unroll
                for (int b = 0; b < 4; b++) {
                    acc[b] += __uint_as_float(a1[b]);
                    cor[b] += __uint_as_float(a2[b]);
                    sacc[b] += __uint_as_float(s0[b]);
                    scor[b] += __uint_as_float(s1[b]);
                }
            }
            asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
            if (io < 0) return;
            const float um = inv_sw;
            const float adv[4] = {ad.x, ad.y, ad.z, ad.w};
            float rr[4];
#pragma unroll
            for (int b = 0; b < 4; b++) {
                const float v = __fmul_rn(fmaf(cor[b], 4.8828125e-4f, acc[b]), um);
                const int j = 4 * cq + b;
                rr[b] = __fadd_rn(__fmul_rn(v, epc[j]), epc[16 + j]);
                if (E.addend) rr[b] = __fadd_rn(rr[b], adv[b]);
                if (E.relu) rr[b] = fmaxf(rr[b], 0.f);
            }
            if (out_h2) {                                    // the next convolution of the layer gathers h2 words
                if (!(fabsf(rr[0]) < EVAD_H2_RANGE && fabsf(rr[1]) < EVAD_H2_RANGE && fabsf(rr[2]) < EVAD_H2_RANGE &&
                      fabsf(rr[3]) < EVAD_H2_RANGE) && status)
                    atomicOr(status, EVAD_STATUS_H16_RANGE);
                *reinterpret_cast<uint4*>(out + (int64_t)io * ldo + 4 * cq) =
                    make_uint4(evad_h2_split(rr[0]), evad_h2_split(rr[1]), evad_h2_split(rr[2]), evad_h2_split(rr[3]));
            } else {
                *reinterpret_cast<float4*>(out + (int64_t)io * ldo + 4 * cq) = make_float4(rr[0], rr[1], rr[2], rr[3]);
            }
            if (has_skip) {
                const float us = inv_ss;
#pragma unroll
                for (int b = 0; b < 4; b++) {
                    const int j = 4 * cq + b;
                    const float sk = __fmul_rn(fmaf(scor[b], 4.8828125e-4f, sacc[b]), us);
                    rr[b] = __fadd_rn(__fmul_rn(sk, epc[32 + j]), epc[48 + j]);
                    if (ES.addend) rr[b] = __fadd_rn(rr[b], ES.addend[(int64_t)io * ES.ld_add + j]);
                    if (ES.relu) rr[b] = fmaxf(rr[b], 0.f);
                }
                *reinterpret_cast<float4*>(skip_out + (int64_t)io * ld_skip + 4 * cq) = make_float4(rr[0], rr[1], rr[2], rr[3]);
            }
        };

        load_ids(subtile(0));
#pragma unroll
        for (int k = 0; k < 4; k++) load_head(k);
        pack_class(0);
        gather(0, 0);
        load_ids(subtile(1));

        int tn = 0;                                          // UMMA tiles completed by this CTA
        for (int64_t q = 0; subtile(q) >= 0; q++) {
            const int sub = (int)(q & 1);
            const bool more = subtile(q + 1) >= 0;
            const bool close = sub == 1 || !more;            // this sub-tile completes a UMMA tile
            if (lane < 4) {                                  // node ids (perm order) and row validity for the epilogue
                int idv = hd[0].i;
                if (lane == 1) idv = hd[1].i;
                if (lane == 2) idv = hd[2].i;
                if (lane == 3) idv = hd[3].i;
                ids[(tn & 3) * L1T_ROWS + sub * L1T_SUB + warp * 4 + lane] = idv;
                if (sub == 0 && !more) ids[(tn & 3) * L1T_ROWS + L1T_SUB + warp * 4 + lane] = -1;   // odd tail: no second half
            }
            // ------------------------------------------------------------ (1) tap aggregation: one node per warp at a time
#pragma unroll
            for (int k = 0; k < 4; k++) {
                const int buf = k & 1;
                if (k + 1 < 4) {
                    pack_class(k + 1);
                    gather(k + 1, buf ^ 1);
                }
                // MMA k slots (2t, 2t+1, 2t+8, 2t+9) of a lane = its row positions t, t+4, t+8, t+12 (j = 0..3)
                uint32_t ah[4], al[4];
#pragma unroll
                for (int p = 0; p < 2; p++) {
                    const uint32_t c0 = (hd[k].c >> (16 * p)) & 0xffu, c1 = (hd[k].c >> (16 * p + 8)) & 0xffu;
                    const uint2 r0 = *reinterpret_cast<const uint2*>(twg + c0 * L1T_TW_LD);
                    const uint2 r1 = *reinterpret_cast<const uint2*>(twg + c1 * L1T_TW_LD);
                    ah[2 * p] = L1T_LO(r0.x, r1.x);          // slot g,     positions j = 2p, 2p+1
                    ah[2 * p + 1] = L1T_HI(r0.x, r1.x);      // slot g + 8
                    al[2 * p] = L1T_LO(r0.y, r1.y);
                    al[2 * p + 1] = L1T_HI(r0.y, r1.y);
                }
                float d1[NT][4], d2[NT][4];                  // hi x hi | hi x lo + lo x hi (weight 2^-11)
#pragma unroll
                for (int nt = 0; nt < NT; nt++)
#pragma unroll
                    for (int b = 0; b < 4; b++) d1[nt][b] = 0.f, d2[nt][b] = 0.f;
                uint32_t bh[NT][2], bl[NT][2];
#pragma unroll
                for (int nt = 0; nt < NT; nt++)
#pragma unroll
                    for (int p = 0; p < 2; p++) {
                        const uint32_t w0 = nt == 0 ? xa[buf][2 * p].x : nt == 1 ? xa[buf][2 * p].y : xb[buf][2 * p];
                        const uint32_t w1 = nt == 0 ? xa[buf][2 * p + 1].x : nt == 1 ? xa[buf][2 * p + 1].y : xb[buf][2 * p + 1];
                        bh[nt][p] = L1T_LO(w0, w1);
                        bl[nt][p] = L1T_HI(w0, w1);
                    }
#pragma unroll
                for (int nt = 0; nt < NT; nt++) l1t_mma_sync(d1[nt], ah, bh[nt][0], bh[nt][1]);
#pragma unroll
                for (int nt = 0; nt < NT; nt++) l1t_mma_sync(d2[nt], ah, bl[nt][0], bl[nt][1]);
#pragma unroll
                for (int nt = 0; nt < NT; nt++) l1t_mma_sync(d2[nt], al, bh[nt][0], bh[nt][1]);
                // the row head of this position in the NEXT sub-tile (its node id arrived a sub-tile ago): in flight
                // until that sub-tile's aggregation, three to four nodes from now
                load_head(k);
                // lane (g, t): slot g and slot g + 8 of channels 4t..4t+3 (and 16+2t, 17+2t for t < 2); slot 15 = own row,
                // already h2 words.  fp16 pairs without a block scale: |P| < 16 * 2048 fits, small values keep 2^-36 absolute
                const int r = sub * L1T_SUB + warp * 4 + k;                             // row of the tile
                uint4 h1, h2;
                l1t_split2(d1[0][0], d2[0][0], d1[1][0], d2[1][0], &h1.x, &h2.x);
                l1t_split2(d1[0][1], d2[0][1], d1[1][1], d2[1][1], &h1.y, &h2.y);
                l1t_split2(d1[0][2], d2[0][2], d1[1][2], d2[1][2], &h1.z, &h2.z);
                l1t_split2(d1[0][3], d2[0][3], d1[1][3], d2[1][3], &h1.w, &h2.w);
                if (g == 7) {
                    h1.z = L1T_LO(xr[buf].x, xr[buf].y); h1.w = L1T_LO(xr[buf].z, xr[buf].w);
                    h2.z = L1T_HI(xr[buf].x, xr[buf].y); h2.w = L1T_HI(xr[buf].z, xr[buf].w);
                }
                // the P tile is free once the MMAs of the previous UMMA tile have completed (first store of a tile only)
                if (k == 0 && sub == 0 && tn > 0)
                    t5_mbar_wait(t5_smem(&s_bar[(tn - 1) & 1]), (uint32_t)(((tn - 1) >> 1) & 1));
                {
                    const int ch = g >> 1, cu = (g & 1) * 4 + t;
                    uint8_t* dst = p_g + ch * L1T_CHUNK + r * 128 + ((cu ^ (r & 7)) << 4);
                    *reinterpret_cast<uint4*>(dst) = h1;
                    *reinterpret_cast<uint4*>(dst + NCH * L1T_CHUNK) = h2;
                }
                if (CP > 16 && t < 2) {
                    uint2 e1, e2;
                    l1t_split2(d1[NT - 1][0], d2[NT - 1][0], d1[NT - 1][1], d2[NT - 1][1], &e1.x, &e2.x);
                    l1t_split2(d1[NT - 1][2], d2[NT - 1][2], d1[NT - 1][3], d2[NT - 1][3], &e1.y, &e2.y);
                    if (g == 7) {
                        e1.y = L1T_LO(xr2[buf].x, xr2[buf].y);
                        e2.y = L1T_HI(xr2[buf].x, xr2[buf].y);
                    }
                    uint8_t* dst = p_g + 4 * L1T_CHUNK + r * 128 + ((g ^ (r & 7)) << 4) + t * 8;
                    *reinterpret_cast<uint2*>(dst) = e1;
                    *reinterpret_cast<uint2*>(dst + NCH * L1T_CHUNK) = e2;
                }
            }
            // node ids of the sub-tile after the next (the next one's heads are in flight already); first node's rows
            load_ids(subtile(q + 2));
            if (more) {
                pack_class(0);
                gather(0, 0);
            }
            if (!close) continue;
            // ------------------------------------------------------------ (2) hand the tile to the MMA warp, do not wait
            asm volatile("fence.proxy.async.shared::cta;" ::: "memory");   // generic-proxy writes of P -> visible to the tensor core
            asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
            if ((tn & (L1T_WARPS - 1)) != warp) {
                asm volatile("bar.arrive 1, %0;" ::"n"(L1T_THREADS) : "memory");
            } else {
                asm volatile("bar.sync 1, %0;" ::"n"(L1T_THREADS) : "memory");   // every warp has stored (and fenced) its rows
                asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
                if (lane == 0) {
                    const uint32_t tacc = tmem + (uint32_t)((tn & 1) * L1T_ACC_COLS);
#pragma unroll
                    for (int ch = 0; ch < NCH; ch++) {
                        const bool wide = has_skip && ch >= 3;
                        const uint64_t p1 = t5_desc(p_base + ch * L1T_CHUNK), p2 = t5_desc(p_base + (NCH + ch) * L1T_CHUNK);
                        const uint64_t wa = t5_desc(w_base + ch * L1T_WCHUNK), wb = t5_desc(w_base + ch * L1T_WCHUNK + 16 * 128);
                        const uint32_t dcol = tacc + (ch < 3 ? 32u * ch : 96u + 64u * (ch - 3));
#pragma unroll
                        for (int kk = 0; kk < 4; kk++) {     // 16 halves = 32 bytes = 2 descriptor units per K step
                            const uint64_t o = (uint64_t)(2 * kk);
                            if (wide) {
                                l1t_umma(dcol, p1 + o, wa + o, L1T_IDESC(64), kk != 0);
                                l1t_umma(tacc + L1T_COL_P2, p2 + o, wb + o, L1T_IDESC(32), (ch | kk) != 0);
                            } else if (has_skip) {           // the p2 region keeps its 32-column shape: s1 rows are zero here
                                l1t_umma(dcol, p1 + o, wa + o, L1T_IDESC(32), kk != 0);
                                l1t_umma(tacc + L1T_COL_P2, p2 + o, wb + o, L1T_IDESC(32), (ch | kk) != 0);
                            } else {
                                l1t_umma(dcol, p1 + o, wa + o, L1T_IDESC(32), kk != 0);
                                l1t_umma(tacc + L1T_COL_P2, p2 + o, wb + o, L1T_IDESC(16), (ch | kk) != 0);
                            }
                        }
                    }
                    t5_mma_commit(t5_smem(&s_bar[tn & 1]));   // arrives when every MMA above has completed
                }
                __syncwarp();
            }
            // ------------------------------------------------------------ (3) epilogue of the PREVIOUS tile: its MMAs
            // finished about a tile ago; this tile's MMAs run meanwhile
            if (tn > 0) epilogue(tn - 1);
            tn++;
        }
        if (tn > 0) epilogue(tn - 1);
    }
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    if (warp == 0)
        asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem), "n"(L1T_TMEM_COLS) : "memory");
}

template <int CP>
static int launch_l1_tc5(const int32_t* count, const int32_t* src, const uint8_t* cls, const int32_t* perm, int64_t n,
                         const uint32_t* x, int ldx, const float* tapw, int ncls, const uint16_t* W16, float inv_sw,
                         const uint16_t* S16, float inv_ss, const evad_epilogue_t& E, float* out, int ldo, int out_h2,
                         const evad_epilogue_t& ES, float* skip_out, int ld_skip, uint32_t* status, cudaStream_t st) {
    const size_t smem = L1TCfg<CP>::smem_bytes(ncls);
    if (smem > 227 * 1024) return EVAD_ERR_UNSUPPORTED;
    static size_t attr_bytes = 0;
    if (smem > attr_bytes) {
        if (cudaFuncSetAttribute(spline_l1_tc5_kernel<CP>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem) != cudaSuccess)
            return EVAD_ERR_LAUNCH;
        attr_bytes = smem;
    }
    const int64_t ntiles = (n + L1T_ROWS - 1) / L1T_ROWS;
    const int grid = (int)(ntiles < 148 ? ntiles : 148);     // persistent: one CTA per SM
    spline_l1_tc5_kernel<CP><<<grid, L1T_THREADS, smem, st>>>(count, src, cls, perm, n, x, ldx, tapw, ncls, W16, inv_sw, S16,
                                                               inv_ss, E, out, ldo, out_h2, ES, skip_out, ld_skip, status);
    EVAD_CHECK_LAUNCH();
    return EVAD_OK;
}

extern "C" int evad_spline_conv_l1_tc5(const int32_t* count, const int32_t* nbr, const uint8_t* nbr_cls, const int32_t* perm,
                                       int64_t n, const uint32_t* x, int ldx, int cin_pad, const float* tapw, int ncls,
                                       const uint16_t* w16, float inv_sw, const uint16_t* s16, float inv_ss,
                                       const evad_epilogue_t* epi_h, float* out, int ldo, int out_h2,
                                       const evad_epilogue_t* skip_epi_h, float* skip_out, int ld_skip, uint32_t* status,
                                       void* stream) {
    if (!count || !nbr || !nbr_cls || !x || !tapw || !w16 || !epi_h || !out) return EVAD_ERR_ARG;
    if (cin_pad != 16 && cin_pad != 20) return EVAD_ERR_UNSUPPORTED;
    if (n * 16 >= (1ll << 31) || n * (int64_t)ldx >= (1ll << 32)) return EVAD_ERR_ARG;   // 32-bit row offsets in the kernel
    if (epi_h->addend && ((((uintptr_t)epi_h->addend) & 15) != 0 || epi_h->ld_add % 4 != 0)) return EVAD_ERR_ARG;
    if (ldx < cin_pad || ldx % 4 != 0 || ldo % 4 != 0 || ncls < 1 || ncls > 255) return EVAD_ERR_ARG;
    if ((((uintptr_t)x | (uintptr_t)tapw | (uintptr_t)w16 | (uintptr_t)out | (uintptr_t)nbr | (uintptr_t)nbr_cls) & 15) != 0)
        return EVAD_ERR_ARG;
    const bool has_skip = s16 && skip_epi_h && skip_out;
    if (has_skip && ((((uintptr_t)s16 | (uintptr_t)skip_out) & 15) != 0 || ld_skip % 4 != 0)) return EVAD_ERR_ARG;
    if (n <= 0) return EVAD_OK;
    const evad_epilogue_t es = has_skip ? *skip_epi_h : *epi_h;
    cudaStream_t st = (cudaStream_t)stream;
    if (cin_pad == 20)
        return launch_l1_tc5<20>(count, nbr, nbr_cls, perm, n, x, ldx, tapw, ncls, w16, inv_sw, has_skip ? s16 : nullptr,
                                 inv_ss, *epi_h, out, ldo, out_h2, es, skip_out, ld_skip, status, st);
    return launch_l1_tc5<16>(count, nbr, nbr_cls, perm, n, x, ldx, tapw, ncls, w16, inv_sw, has_skip ? s16 : nullptr, inv_ss,
                             *epi_h, out, ldo, out_h2, es, skip_out, ld_skip, status, st);
}
