// eventad_b200/csrc/spline_l1_h16.cu -- full-resolution SplineConv (layer 1) with BOTH tensor-core products on fp16 operand
// pairs (mma.sync.m16n8k16) instead of 3xTF32 (m16n8k8): half the HMMA instructions per node and no in-register operand
// splitting.
//
// Reference path replaced per call (paths relative to /root/reference/src/dagr/model): as csrc/spline_l1.cu --
//   layers/spline_conv.py:39-47 (message_lut), PyG propagate sum aggregation, spline_conv.py:72-76 (root weight),
//   layers/conv.py:23-28,47-56 (eval BatchNorm, activation, skip Linear + BatchNorm).
//
// Why (profiles/r02_summary.md, "layer-1 sensitivity"): spline_l1_mma_kernel issues 18 + 15 legacy HMMA.1688.TF32 per node
// and ~60 LOP/FADD per node to split operands into TF32 hi/lo; halving either product's MMA work moved the kernel by
// 12-15 %, the row gathers by 5 %.  The legacy tensor path issues m16n8k16 f16 at the rate of m16n8k8 tf32, so:
//   * every fp32 value v travels as an "h2 word": low half hi = rn16(v), high half lo = rn16((v - hi) * 2048), i.e.
//     v = hi + lo * 2^-11 to 2^-22 relative accuracy (|v| < 2048 is required and checked where the words are produced:
//     EVAD_STATUS_H16_RANGE).  The node features x arrive in HBM in this form (evad_split_h2, or the epilogue of this
//     kernel for the layer's second convolution): same 4 bytes per element, and a B fragment is one PRMT per register.
//   * a product a * b is hi_a hi_b (accumulator 1) + hi_a lo_b + lo_a hi_b (accumulator 2, weight 2^-11): three
//     m16n8k16 per (16 x 8 x 16) tile where 3xTF32 needs six m16n8k8.
//   * aggregation  P_i[slot][c] = sum_e tapw[class_e][slot] x_{src_e}[c]: 9 (Cin 20) / 6 (Cin 16) HMMA per node (was 18 / 12);
//     contraction  out = P @ W_cat over K = 16 * Cin: 7.5 / 6 HMMA per node (was 15 / 12).
//   * the P tile in shared memory is two row-major fp16 planes (leading parts, scaled remainders): an A fragment of the
//     contraction is one ldmatrix.x4 per part (registers arrive in fragment order, no moves), the aggregation's D
//     fragments leave as one 64-bit store per part and slot.
// Everything else (32-node tiles in pixel order, 8 warps, two CTAs per SM, prefetch of ids / row heads / first gather,
// four k quarters + partial sums, fused BN / skip / ReLU epilogue) is spline_l1_mma_kernel's.
#include <cuda_fp16.h>

#include "spline_common.cuh"

#define L1H_NODES 32
#define L1H_THREADS 256
#define L1H_SLOTS 16
#define L1H_TW_LD 24                  // words per class row of the tap table: 8 x (hi pair, lo pair) + pad
#define L1H_LDC 20                    // C tile row stride (floats)
#define L1H_AUX 128                   // words: node ids of the tile [2 parities][32] + epilogue scale / shift [4][16]
#define L1H_RANGE EVAD_H2_RANGE        // |v| bound of an h2 word (P sums 16 of them: < 32768 < 65504)

template <int CP>
struct L1HCfg {
    static constexpr int KTOT = L1H_SLOTS * CP;              // 320 / 256 contraction columns
    static constexpr int NKB = KTOT / 16;                    // k16 blocks: 20 / 16
    static constexpr int LDW = KTOT / 2 + 4;                 // words per row of one fp16 plane: 16 bytes (mod 128), ldmatrix rows
                                                             // of an 8 x 8 block fall into distinct bank groups
    static constexpr int SKB = CP > 16 ? 2 : 1;              // k16 blocks of the skip Linear
    static constexpr int NT = (CP + 7) / 8;                  // channel tiles of the aggregation MMA
    static constexpr int P_WORDS = 2 * L1H_NODES * LDW;      // plane of leading parts, plane of remainders
    static constexpr int W_WORDS = NKB * 2 * 32 * 4;
    static constexpr int S_WORDS = SKB * 2 * 32 * 4;
    static constexpr int C_FLOATS = 5 * L1H_NODES * L1H_LDC; // four k-quarter partials + skip
    static size_t smem_bytes(int ncls) {
        return (size_t)(P_WORDS + W_WORDS + S_WORDS + C_FLOATS + (ncls + 1) * L1H_TW_LD + L1H_AUX) * 4;
    }
    // P column k -> row of W_cat ([slot][channel], channel stride CP).  Columns 0..255: slot k / 16, channel k % 16;
    // columns 256.. (Cin 20): 4-column groups (g, t), g < 8, t < 2 = (slot g | slot g + 8) x channels 16 + 2t, 17 + 2t
    __host__ __device__ static constexpr int wrow(int k) {
        return k < 256 ? (k >> 4) * CP + (k & 15)
                       : (((k - 256) >> 3) + 8 * ((k >> 1) & 1)) * CP + 16 + 2 * (((k - 256) >> 2) & 1) + (k & 1);
    }
};

struct L1HHead {                      // what a lane keeps of one node's row: id, length, its 4 sources and their classes
    int i, deg;
    uint32_t s[4];                    // sources of row positions t, t+4, t+8, t+12, clamped to valid rows
    uint32_t c;                       // their class bytes; positions beyond the row length hold the all-zero class
};

__device__ __forceinline__ uint32_t l1h_split(float v) { return evad_h2_split(v); }
__device__ __forceinline__ float l1h_join(uint32_t w) {      // h2 word -> fp32
    return __fmaf_rn(__half2float(__ushort_as_half((unsigned short)(w >> 16))), 1.f / 2048.f,
                     __half2float(__ushort_as_half((unsigned short)(w & 0xffffu))));
}
// two fp32 values a = v + e * 2^-11 (v, e: the two accumulators) -> packed leading halves and packed scaled remainders
__device__ __forceinline__ void l1h_split2(float v0, float e0, float v1, float e1, uint32_t* p1, uint32_t* p2) {
    const __half2 h = __floats2half2_rn(v0, v1);
    const float2 f = __half22float2(h);
    const __half2 l = __floats2half2_rn(__fmaf_rn(__fsub_rn(v0, f.x), 2048.f, e0), __fmaf_rn(__fsub_rn(v1, f.y), 2048.f, e1));
    *p1 = *reinterpret_cast<const uint32_t*>(&h);
    *p2 = *reinterpret_cast<const uint32_t*>(&l);
}
__device__ __forceinline__ void l1h_mma(float (&d)[4], const uint32_t (&a)[4], uint32_t b0, uint32_t b1) {
    asm("mma.sync.aligned.m16n8k16.row.col.f32.f16.f16.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
        : "+f"(d[0]), "+f"(d[1]), "+f"(d[2]), "+f"(d[3])
        : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(b0), "r"(b1));
}
__device__ __forceinline__ void l1h_ldmatrix(uint32_t (&r)[4], uint32_t addr) {
    asm volatile("ldmatrix.sync.aligned.m8n8.x4.shared.b16 {%0,%1,%2,%3}, [%4];"
                 : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]) : "r"(addr) : "memory");
}
#define L1H_LO(a, b) __byte_perm((a), (b), 0x5410)            // (low half of a, low half of b)
#define L1H_HI(a, b) __byte_perm((a), (b), 0x7632)            // (high half of a, high half of b)

// ------------------------------------------------------------------------------------------------ fp32 -> h2 words
__global__ void __launch_bounds__(256) split_h2_kernel(const float4* __restrict__ in, uint4* __restrict__ out, int64_t n4,
                                                       uint32_t* __restrict__ status) {
    const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n4) return;
    const float4 v = __ldg(in + i);
    if (!(fabsf(v.x) < L1H_RANGE && fabsf(v.y) < L1H_RANGE && fabsf(v.z) < L1H_RANGE && fabsf(v.w) < L1H_RANGE) && status)
        atomicOr(status, EVAD_STATUS_H16_RANGE);
    out[i] = make_uint4(l1h_split(v.x), l1h_split(v.y), l1h_split(v.z), l1h_split(v.w));
}

extern "C" int evad_split_h2(const float* in, uint32_t* out, int64_t n_words, uint32_t* status, void* stream) {
    if (!in || !out || n_words < 0 || n_words % 4 != 0 || ((((uintptr_t)in | (uintptr_t)out) & 15) != 0)) return EVAD_ERR_ARG;
    if (n_words == 0) return EVAD_OK;
    split_h2_kernel<<<evad_cdiv(n_words / 4, 256), 256, 0, (cudaStream_t)stream>>>(
        reinterpret_cast<const float4*>(in), reinterpret_cast<uint4*>(out), n_words / 4, status);
    EVAD_CHECK_LAUNCH();
    return EVAD_OK;
}

// ------------------------------------------------------------------------------------------------ the convolution
template <int CP>
__global__ void __launch_bounds__(L1H_THREADS, 2) spline_l1_h16_kernel(
    const int32_t* __restrict__ count, const int32_t* __restrict__ src, const uint8_t* __restrict__ cls,
    const int32_t* __restrict__ perm, int64_t n, int64_t n_rows, const uint32_t* __restrict__ x, int ldx,
    const float* __restrict__ tapw, int ncls, const float* __restrict__ Whi, const float* __restrict__ Wlo, const float* __restrict__ Shi,
    const float* __restrict__ Slo, evad_epilogue_t E, float* __restrict__ out, int ldo, int out_h2, evad_epilogue_t ES,
    float* __restrict__ skip_out, int ld_skip, uint32_t* __restrict__ status) {
    using C = L1HCfg<CP>;
    constexpr int LDW = C::LDW, NT = C::NT;
    extern __shared__ __align__(16) uint32_t l1h_smem[];
    uint32_t* P1 = l1h_smem;                                 // [32][LDW] words = [32][2 LDW] halves
    uint32_t* P2 = P1 + L1H_NODES * LDW;
    uint4* Wf = reinterpret_cast<uint4*>(P1 + C::P_WORDS);
    uint4* Sf = Wf + C::W_WORDS / 4;
    float* Cs = reinterpret_cast<float*>(Sf + C::S_WORDS / 4);
    uint32_t* tw = reinterpret_cast<uint32_t*>(Cs + C::C_FLOATS);
    int* ids = reinterpret_cast<int*>(tw + (ncls + 1) * L1H_TW_LD);
    float* epc = reinterpret_cast<float*>(ids + 64);
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int g = lane >> 2, t = lane & 3;
    const bool has_skip = Shi != nullptr;

    if (tid < 64) {                                          // epilogue constants: scale (1 when absent), shift (0 when absent)
        const int c = tid & 15, w = tid >> 4;
        const float* pq = w == 0 ? E.scale : w == 1 ? E.shift : w == 2 ? ES.scale : ES.shift;
        epc[tid] = pq ? pq[c] : (w & 1) ? 0.f : 1.f;
    }
    // one-time staging.  Weight fragments: entry (kb, nt, lane) = W_cat rows of P columns 16 kb + {2t, 2t+1, 2t+8, 2t+9},
    // column 8 nt + g, as (w1 k0 k1, w1 k2 k3, w2 k0 k1, w2 k2 k3); W = Whi + Wlo exactly (the TF32 split of the pack)
    for (int idx = tid; idx < C::NKB * 64; idx += L1H_THREADS) {
        const int kb = idx >> 6, nt = (idx >> 5) & 1, l = idx & 31, k0 = 16 * kb + 2 * (l & 3), col = 8 * nt + (l >> 2);
        uint32_t w[4];
#pragma unroll
        for (int e = 0; e < 4; e++) {
            const int r = C::wrow(k0 + (e & 1) + 8 * (e >> 1));
            w[e] = l1h_split(__fadd_rn(__ldg(Whi + r * 16 + col), __ldg(Wlo + r * 16 + col)));
        }
        Wf[idx] = make_uint4(L1H_LO(w[0], w[1]), L1H_LO(w[2], w[3]), L1H_HI(w[0], w[1]), L1H_HI(w[2], w[3]));
    }
    if (has_skip)
        for (int idx = tid; idx < C::SKB * 64; idx += L1H_THREADS) {
            // skip block 0 reads P columns 240..255 (root slot, channels 0..15); block 1 (Cin 20) reads columns 304..319,
            // of which 314, 315, 318, 319 are the root slot's channels 16..19
            const int sb = idx >> 6, nt = (idx >> 5) & 1, l = idx & 31, col = 8 * nt + (l >> 2);
            uint32_t w[4] = {0u, 0u, 0u, 0u};
#pragma unroll
            for (int e = 0; e < 4; e++) {
                const int kk = 2 * (l & 3) + (e & 1) + 8 * (e >> 1);          // column inside the block
                int ch = -1;
                if (sb == 0) ch = kk;
                else if (kk >= 8 && (kk & 2)) ch = 16 + 2 * ((kk >> 2) & 1) + (kk & 1);
                if (ch >= 0) w[e] = l1h_split(__fadd_rn(__ldg(Shi + ch * 16 + col), __ldg(Slo + ch * 16 + col)));
            }
            Sf[idx] = make_uint4(L1H_LO(w[0], w[1]), L1H_LO(w[2], w[3]), L1H_HI(w[0], w[1]), L1H_HI(w[2], w[3]));
        }
    // tap table: row c, lane group g: word 2g = (hi[c][g], hi[c][g+8]), word 2g+1 = (lo[c][g], lo[c][g+8]); row ncls = zeros
    for (int idx = tid; idx < (ncls + 1) * 8; idx += L1H_THREADS) {
        const int r = idx >> 3, k = idx & 7;
        uint32_t a = 0u, b = 0u;
        if (r < ncls) {
            a = l1h_split(__ldg(tapw + r * L1H_SLOTS + k));
            b = l1h_split(__ldg(tapw + r * L1H_SLOTS + k + 8));
        }
        *reinterpret_cast<uint2*>(tw + r * L1H_TW_LD + 2 * k) = make_uint2(L1H_LO(a, b), L1H_HI(a, b));
    }
    for (int idx = tid; idx < C::P_WORDS; idx += L1H_THREADS) P1[idx] = 0u;
    __syncthreads();

    const int mt = warp & 1, kq = warp >> 1;                 // contraction role: m16 tile, k quarter
    const int64_t ntiles = (n + L1H_NODES - 1) / L1H_NODES;
    const uint32_t zero4 = (uint32_t)ncls * 0x01010101u;     // four copies of the all-zero class
    const uint32_t nmax = (uint32_t)(n_rows - 1);
    // lane bases of the row gathers; an address is base + row * (4 ldx) as ONE 32 x 32 + 64 bit multiply-add
    const char* xg = reinterpret_cast<const char*>(x + 2 * g);
    const char* xg16 = reinterpret_cast<const char*>(x + 16 + g);
    const char* xt = reinterpret_cast<const char*>(x + 4 * t);
    const char* xt16 = reinterpret_cast<const char*>(x + 16 + 2 * t);
    const uint32_t ldxb = 4u * (uint32_t)ldx;

    const uint32_t* twg = tw + 2 * g;
    L1HHead hd[4];
    uint4 craw[4];
    int inext[4];
    auto load_ids = [&](int64_t tile) {
#pragma unroll
        for (int k = 0; k < 4; k++) {
            const int64_t ip = tile * L1H_NODES + warp * 4 + k;
            inext[k] = -1;
            if (tile < ntiles && ip < n) inext[k] = perm ? perm[ip] : (int)ip;
        }
    };
    auto load_heads = [&]() {
#pragma unroll
        for (int k = 0; k < 4; k++) {
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
        }
    };
    auto pack_classes = [&]() {
#pragma unroll
        for (int k = 0; k < 4; k++) {
            const uint32_t c = __byte_perm(__byte_perm(craw[k].x, craw[k].y, t | ((4 + t) << 4)),
                                           __byte_perm(craw[k].z, craw[k].w, t | ((4 + t) << 4)), 0x5410);
            // positions t + 4j < deg  <=>  j < nv; slots beyond the row hold garbage: all-zero class, any valid source row
            const int nv = min(4, (hd[k].deg - t + 3) >> 2);
            const uint32_t live = __funnelshift_rc(0xffffffffu, 0u, 32 - 8 * nv);
            hd[k].c = (c & live) | (zero4 & ~live);
#pragma unroll
            for (int j = 0; j < 4; j++) hd[k].s[j] = min(hd[k].s[j], nmax);
        }
    };
    uint2 xa[2][4];                                          // [buffer][row position j]: h2 words of channels 2g, 2g+1
    uint32_t xb[2][4];                                       // channel 16 + g (Cin 20)
    uint4 xr[2];                                             // the node's own row (lanes g == 7): channels 4t..4t+3
    uint2 xr2[2];                                            //   and 16+2t, 17+2t
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
    load_ids(blockIdx.x);
    load_heads();
    pack_classes();
    gather(0, 0);
    load_ids((int64_t)blockIdx.x + gridDim.x);

    int par = 0;
    for (int64_t tile = blockIdx.x; tile < ntiles; tile += gridDim.x, par ^= 32) {
        if (lane < 4) {                                      // this tile's node ids for the epilogue threads
            int idv = hd[0].i;
            if (lane == 1) idv = hd[1].i;
            if (lane == 2) idv = hd[2].i;
            if (lane == 3) idv = hd[3].i;
            ids[par + warp * 4 + lane] = idv;
        }
        // ---------------------------------------------------------------- (1) tap aggregation: one node per warp at a time
        // MMA k slots (2t, 2t+1, 2t+8, 2t+9) of a lane = its row positions t, t+4, t+8, t+12 (j = 0..3)
#pragma unroll
        for (int k = 0; k < 4; k++) {
            const int buf = k & 1;
            if (k + 1 < 4) gather(k + 1, buf ^ 1);
            uint32_t ah[4], al[4];
#pragma unroll
            for (int p = 0; p < 2; p++) {
                const uint32_t c0 = (hd[k].c >> (16 * p)) & 0xffu, c1 = (hd[k].c >> (16 * p + 8)) & 0xffu;
                const uint2 r0 = *reinterpret_cast<const uint2*>(twg + c0 * L1H_TW_LD);
                const uint2 r1 = *reinterpret_cast<const uint2*>(twg + c1 * L1H_TW_LD);
                ah[2 * p] = L1H_LO(r0.x, r1.x);              // slot g,     positions j = 2p, 2p+1
                ah[2 * p + 1] = L1H_HI(r0.x, r1.x);          // slot g + 8
                al[2 * p] = L1H_LO(r0.y, r1.y);
                al[2 * p + 1] = L1H_HI(r0.y, r1.y);
            }
            float d1[NT][4], d2[NT][4];
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
                    bh[nt][p] = L1H_LO(w0, w1);
                    bl[nt][p] = L1H_HI(w0, w1);
                }
            // product by product over the channel tiles: consecutive MMAs are independent
#pragma unroll
            for (int nt = 0; nt < NT; nt++) l1h_mma(d1[nt], ah, bh[nt][0], bh[nt][1]);
#pragma unroll
            for (int nt = 0; nt < NT; nt++) l1h_mma(d2[nt], ah, bl[nt][0], bl[nt][1]);
#pragma unroll
            for (int nt = 0; nt < NT; nt++) l1h_mma(d2[nt], al, bh[nt][0], bh[nt][1]);
            // D fragments -> the two fp16 planes of the P row: 4 columns (channels 4t..4t+3) of slot g and of slot g + 8
            const int prow = (warp * 4 + k) * LDW;
            uint2 q0, r0, q1, r1;                            // q: leading parts, r: remainders
            l1h_split2(d1[0][0], d2[0][0], d1[1][0], d2[1][0], &q0.x, &r0.x);     // channels 4t, 4t+1
            l1h_split2(d1[0][1], d2[0][1], d1[1][1], d2[1][1], &q0.y, &r0.y);     // channels 4t+2, 4t+3
            l1h_split2(d1[0][2], d2[0][2], d1[1][2], d2[1][2], &q1.x, &r1.x);
            l1h_split2(d1[0][3], d2[0][3], d1[1][3], d2[1][3], &q1.y, &r1.y);
            if (g == 7) {                                    // slot 15 = the node's own row (root weight), already h2 words
                q1 = make_uint2(L1H_LO(xr[buf].x, xr[buf].y), L1H_LO(xr[buf].z, xr[buf].w));
                r1 = make_uint2(L1H_HI(xr[buf].x, xr[buf].y), L1H_HI(xr[buf].z, xr[buf].w));
            }
            *reinterpret_cast<uint2*>(P1 + prow + 8 * g + 2 * t) = q0;
            *reinterpret_cast<uint2*>(P2 + prow + 8 * g + 2 * t) = r0;
            *reinterpret_cast<uint2*>(P1 + prow + 8 * (g + 8) + 2 * t) = q1;
            *reinterpret_cast<uint2*>(P2 + prow + 8 * (g + 8) + 2 * t) = r1;
            if (CP > 16 && t < 2) {                          // columns 256 + 4 (2g + t) ..: (slot g | slot g + 8) x channels 16+2t, 17+2t
                uint2 q2, r2;
                l1h_split2(d1[NT - 1][0], d2[NT - 1][0], d1[NT - 1][1], d2[NT - 1][1], &q2.x, &r2.x);
                l1h_split2(d1[NT - 1][2], d2[NT - 1][2], d1[NT - 1][3], d2[NT - 1][3], &q2.y, &r2.y);
                if (g == 7) {
                    q2.y = L1H_LO(xr2[buf].x, xr2[buf].y);
                    r2.y = L1H_HI(xr2[buf].x, xr2[buf].y);
                }
                *reinterpret_cast<uint2*>(P1 + prow + 128 + 2 * (2 * g + t)) = q2;
                *reinterpret_cast<uint2*>(P2 + prow + 128 + 2 * (2 * g + t)) = r2;
            }
        }
        // next tile's row heads (their node ids arrived a tile ago): in flight over the first half of the contraction;
        // the node ids of the tile after that: in flight over a whole tile
        const bool more = tile + gridDim.x < ntiles;
        load_heads();
        load_ids(tile + 2 * (int64_t)gridDim.x);
        __syncthreads();
        // epilogue threads fetch their addend row now: it lands during the contraction
        const int e_nl = tid >> 2, e_oq = tid & 3;
        int e_io = -1;
        float4 e_add = make_float4(0.f, 0.f, 0.f, 0.f);
        if (tid < 4 * L1H_NODES) {
            e_io = ids[par + e_nl];
            if (e_io >= 0 && E.addend) e_add = __ldg(reinterpret_cast<const float4*>(E.addend + (int64_t)e_io * E.ld_add) + e_oq);
        }
        // ---------------------------------------------------------------- (2) contraction P tile @ W_cat, k quarter per warp
        {
            float a1[2][4] = {{0.f, 0.f, 0.f, 0.f}, {0.f, 0.f, 0.f, 0.f}}, a2[2][4] = {{0.f, 0.f, 0.f, 0.f}, {0.f, 0.f, 0.f, 0.f}};
            constexpr int KQ = C::NKB / 4;
            // ldmatrix.x4: lane l gives the row address of 8x8 block l / 8 = (rows 0-7 | 8-15) x (columns 0-7 | 8-15)
            const uint32_t lm1 = (uint32_t)__cvta_generic_to_shared(P1 + (mt * 16 + (lane & 7) + 8 * ((lane >> 3) & 1)) * LDW +
                                                                    4 * (lane >> 4));
            const uint32_t lm2 = lm1 + 4u * L1H_NODES * LDW;
#pragma unroll
            for (int s = 0; s < KQ; s++) {
                if (s == KQ / 2 && more) {                   // first node of the next tile: rows in flight over the rest
                    pack_classes();
                    gather(0, 0);
                }
                const int kb = kq * KQ + s;
                uint32_t p1[4], p2[4];
                l1h_ldmatrix(p1, lm1 + 32u * kb);
                l1h_ldmatrix(p2, lm2 + 32u * kb);
                const uint4 w0 = Wf[(kb * 2 + 0) * 32 + lane], w1 = Wf[(kb * 2 + 1) * 32 + lane];
                l1h_mma(a1[0], p1, w0.x, w0.y);
                l1h_mma(a1[1], p1, w1.x, w1.y);
                l1h_mma(a2[0], p1, w0.z, w0.w);
                l1h_mma(a2[1], p1, w1.z, w1.w);
                l1h_mma(a2[0], p2, w0.x, w0.y);
                l1h_mma(a2[1], p2, w1.x, w1.y);
            }
#pragma unroll
            for (int nt = 0; nt < 2; nt++)
#pragma unroll
                for (int h = 0; h < 2; h++) {
                    const int row = mt * 16 + g + 8 * h, col = nt * 8 + 2 * t;
                    *reinterpret_cast<float2*>(Cs + (kq * L1H_NODES + row) * L1H_LDC + col) =
                        make_float2(__fmaf_rn(a2[nt][2 * h], 1.f / 2048.f, a1[nt][2 * h]),
                                    __fmaf_rn(a2[nt][2 * h + 1], 1.f / 2048.f, a1[nt][2 * h + 1]));
                }
            if (has_skip && kq == 3) {                       // skip Linear on the root slot
                float s1[2][4] = {{0.f, 0.f, 0.f, 0.f}, {0.f, 0.f, 0.f, 0.f}}, s2[2][4] = {{0.f, 0.f, 0.f, 0.f}, {0.f, 0.f, 0.f, 0.f}};
#pragma unroll
                for (int sb = 0; sb < C::SKB; sb++) {
                    const int kb = sb == 0 ? 15 : 19;
                    uint32_t p1[4], p2[4];
                    l1h_ldmatrix(p1, lm1 + 32u * kb);
                    l1h_ldmatrix(p2, lm2 + 32u * kb);
                    const uint4 w0 = Sf[(sb * 2 + 0) * 32 + lane], w1 = Sf[(sb * 2 + 1) * 32 + lane];
                    l1h_mma(s1[0], p1, w0.x, w0.y);
                    l1h_mma(s1[1], p1, w1.x, w1.y);
                    l1h_mma(s2[0], p1, w0.z, w0.w);
                    l1h_mma(s2[1], p1, w1.z, w1.w);
                    l1h_mma(s2[0], p2, w0.x, w0.y);
                    l1h_mma(s2[1], p2, w1.x, w1.y);
                }
#pragma unroll
                for (int nt = 0; nt < 2; nt++)
#pragma unroll
                    for (int h = 0; h < 2; h++) {
                        const int row = mt * 16 + g + 8 * h, col = nt * 8 + 2 * t;
                        *reinterpret_cast<float2*>(Cs + (4 * L1H_NODES + row) * L1H_LDC + col) =
                            make_float2(__fmaf_rn(s2[nt][2 * h], 1.f / 2048.f, s1[nt][2 * h]),
                                        __fmaf_rn(s2[nt][2 * h + 1], 1.f / 2048.f, s1[nt][2 * h + 1]));
                    }
            }
        }
        __syncthreads();
        // ---------------------------------------------------------------- (3) epilogue: BN / skip add / ReLU
        if (e_io >= 0) {
            float4 v = *reinterpret_cast<const float4*>(Cs + e_nl * L1H_LDC + 4 * e_oq);
#pragma unroll
            for (int p = 1; p < 4; p++) {
                const float4 c = *reinterpret_cast<const float4*>(Cs + (p * L1H_NODES + e_nl) * L1H_LDC + 4 * e_oq);
                v.x += c.x; v.y += c.y; v.z += c.z; v.w += c.w;
            }
            const float4 sc = *reinterpret_cast<const float4*>(epc + 4 * e_oq);
            const float4 sh = *reinterpret_cast<const float4*>(epc + 16 + 4 * e_oq);
            float r[4] = {__fadd_rn(__fmul_rn(v.x, sc.x), sh.x), __fadd_rn(__fmul_rn(v.y, sc.y), sh.y),
                          __fadd_rn(__fmul_rn(v.z, sc.z), sh.z), __fadd_rn(__fmul_rn(v.w, sc.w), sh.w)};
            const float ad[4] = {e_add.x, e_add.y, e_add.z, e_add.w};
#pragma unroll
            for (int b = 0; b < 4; b++) {
                if (E.addend) r[b] = __fadd_rn(r[b], ad[b]);
                if (E.relu) r[b] = fmaxf(r[b], 0.f);
            }
            if (out_h2) {                                    // the next convolution of the layer gathers h2 words
                if (!(fabsf(r[0]) < L1H_RANGE && fabsf(r[1]) < L1H_RANGE && fabsf(r[2]) < L1H_RANGE && fabsf(r[3]) < L1H_RANGE) &&
                    status)
                    atomicOr(status, EVAD_STATUS_H16_RANGE);
                *reinterpret_cast<uint4*>(out + (int64_t)e_io * ldo + 4 * e_oq) =
                    make_uint4(l1h_split(r[0]), l1h_split(r[1]), l1h_split(r[2]), l1h_split(r[3]));
            } else {
                *reinterpret_cast<float4*>(out + (int64_t)e_io * ldo + 4 * e_oq) = make_float4(r[0], r[1], r[2], r[3]);
            }
            if (has_skip) {
                const float4 sv = *reinterpret_cast<const float4*>(Cs + (4 * L1H_NODES + e_nl) * L1H_LDC + 4 * e_oq);
                const float4 ssc = *reinterpret_cast<const float4*>(epc + 32 + 4 * e_oq);
                const float4 ssh = *reinterpret_cast<const float4*>(epc + 48 + 4 * e_oq);
                float q[4] = {__fadd_rn(__fmul_rn(sv.x, ssc.x), ssh.x), __fadd_rn(__fmul_rn(sv.y, ssc.y), ssh.y),
                              __fadd_rn(__fmul_rn(sv.z, ssc.z), ssh.z), __fadd_rn(__fmul_rn(sv.w, ssc.w), ssh.w)};
#pragma unroll
                for (int b = 0; b < 4; b++) {
                    if (ES.addend) q[b] = __fadd_rn(q[b], ES.addend[(int64_t)e_io * ES.ld_add + 4 * e_oq + b]);
                    if (ES.relu) q[b] = fmaxf(q[b], 0.f);
                }
                *reinterpret_cast<float4*>(skip_out + (int64_t)e_io * ld_skip + 4 * e_oq) = make_float4(q[0], q[1], q[2], q[3]);
            }
        }
    }
}

template <int CP>
static int launch_l1_h16(const int32_t* count, const int32_t* src, const uint8_t* cls, const int32_t* perm, int64_t n,
                         int64_t n_rows, const uint32_t* x, int ldx, const float* tapw, int ncls, const float* Whi, const float* Wlo,
                         const float* Shi, const float* Slo, const evad_epilogue_t& E, float* out, int ldo, int out_h2,
                         const evad_epilogue_t& ES, float* skip_out, int ld_skip, uint32_t* status, cudaStream_t st) {
    const size_t smem = L1HCfg<CP>::smem_bytes(ncls);
    if (smem > 113 * 1024) return EVAD_ERR_UNSUPPORTED;      // two CTAs per SM
    static size_t attr_bytes = 0;
    if (smem > attr_bytes) {
        if (cudaFuncSetAttribute(spline_l1_h16_kernel<CP>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem) != cudaSuccess)
            return EVAD_ERR_LAUNCH;
        attr_bytes = smem;
    }
    const int64_t ntiles = (n + L1H_NODES - 1) / L1H_NODES;
    const int grid = (int)(ntiles < 2 * 148 ? ntiles : 2 * 148);
    spline_l1_h16_kernel<CP><<<grid, L1H_THREADS, smem, st>>>(count, src, cls, perm, n, n_rows, x, ldx, tapw, ncls, Whi, Wlo, Shi, Slo,
                                                               E, out, ldo, out_h2, ES, skip_out, ld_skip, status);
    EVAD_CHECK_LAUNCH();
    return EVAD_OK;
}

extern "C" int evad_spline_conv_l1_h16(const int32_t* count, const int32_t* nbr, const uint8_t* nbr_cls, const int32_t* perm,
                                       int64_t n, int64_t n_rows, const uint32_t* x_h2, int ldx, int cin_pad,
                                       const float* tapw, int ncls, const float* Whi, const float* Wlo, const float* Shi, const float* Slo,
                                       const evad_epilogue_t* epi_h, float* out, int ldo, int out_h2,
                                       const evad_epilogue_t* skip_epi_h, float* skip_out, int ld_skip, uint32_t* status,
                                       void* stream) {
    if (!count || !nbr || !nbr_cls || !x_h2 || !tapw || !Whi || !Wlo || !epi_h || !out) return EVAD_ERR_ARG;
    if (cin_pad != 16 && cin_pad != 20) return EVAD_ERR_UNSUPPORTED;
    if (n_rows < n || n_rows * 16 >= (1ll << 31) || n_rows * (int64_t)ldx >= (1ll << 32)) return EVAD_ERR_ARG;
    if (epi_h->addend && ((((uintptr_t)epi_h->addend) & 15) != 0 || epi_h->ld_add % 4 != 0)) return EVAD_ERR_ARG;
    if (ldx < cin_pad || ldx % 4 != 0 || ldo % 4 != 0 || ncls < 1 || ncls > 255) return EVAD_ERR_ARG;   // class ncls = the all-zero row, one byte
    if ((((uintptr_t)x_h2 | (uintptr_t)tapw | (uintptr_t)Whi | (uintptr_t)Wlo | (uintptr_t)out | (uintptr_t)nbr |
          (uintptr_t)nbr_cls) & 15) != 0)
        return EVAD_ERR_ARG;
    const bool has_skip = Shi && Slo && skip_epi_h && skip_out;
    if (has_skip && ((((uintptr_t)Shi | (uintptr_t)Slo | (uintptr_t)skip_out) & 15) != 0 || ld_skip % 4 != 0))
        return EVAD_ERR_ARG;
    if (n <= 0) return EVAD_OK;
    const evad_epilogue_t es = has_skip ? *skip_epi_h : *epi_h;
    cudaStream_t st = (cudaStream_t)stream;
    if (cin_pad == 20)
        return launch_l1_h16<20>(count, nbr, nbr_cls, perm, n, n_rows, x_h2, ldx, tapw, ncls, Whi, Wlo, has_skip ? Shi : nullptr,
                                 has_skip ? Slo : nullptr, *epi_h, out, ldo, out_h2, es, skip_out, ld_skip, status, st);
    return launch_l1_h16<16>(count, nbr, nbr_cls, perm, n, n_rows, x_h2, ldx, tapw, ncls, Whi, Wlo, has_skip ? Shi : nullptr,
                             has_skip ? Slo : nullptr, *epi_h, out, ldo, out_h2, es, skip_out, ld_skip, status, st);
}
