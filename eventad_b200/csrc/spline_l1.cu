// eventad_b200/csrc/spline_l1.cu -- full-resolution SplineConv (layer 1: 19->16 and 16->16 on the event graph) as
// tap aggregation in registers + tensor-core contraction, two persistent CTAs per SM.
//
// Reference path replaced per call (paths relative to /root/reference/src/dagr/model):
//   layers/spline_conv.py:39-47  message_lut (gather x_j, gather lut[ix,iy] = sum_s b_s W[k_s], einsum)
//   PyG propagate sum aggregation + layers/spline_conv.py:72-76 root weight
//   layers/conv.py:23-28,47-56   eval BatchNorm, activation, skip Linear + BatchNorm
//
// Why not the LUT gather (csrc/spline_small.cu): every edge needs its own Cin x Cout matrix, 1280 B of shared-memory
// reads per edge -- ncu (profiles/r01_summary.md) shows that kernel bound by the LDS pipe and instruction issue, with
// DRAM at 1 % of peak.  Here the sum is re-associated:
//     out_i = sum_e (sum_s b_{e,s} W[k_{e,s}])^T x_e  =  sum_k W[k]^T (sum_e b_{e,k} x_e)  =  W_cat^T P_i
// (1) P_i[k] = sum_e b_{e,k} x_e over the <= 15 kernel taps the integer pixel lattice can reach: one thread per
//     (node, channel quad), 60 accumulators in registers, per edge one 128-bit gather of the source row and the 16 tap
//     weights of the edge's lattice class (a 6.5 KB shared table); no per-edge weight matrices.
// (2) out tile [64 nodes x 16] = P tile [64 x 16*Cin_pad] @ W_cat on the tensor cores (mma.sync m16n8k8, 3xTF32 so
//     that fp32 accuracy is kept; W_cat hi/lo parts stay in shared memory for the life of the persistent CTA), root
//     weight as one more tap slot, the skip Linear as a second small product, BatchNorm / skip add / ReLU in the epilogue.
// The lattice class of every edge (the reference's (ix, iy) LUT index, computed with its fp32 arithmetic) is written
// once per graph by evad_edge_class and shared by both convolutions of the layer.
#include "spline_common.cuh"

#define L1_NODES 64                   // nodes per tile; two CTAs per SM so that gather and MMA phases overlap
#define L1_MMA_WARPS (L1_NODES / 16 * 2)   // (m16 tile, k half)
#define L1_SLOTS 16                   // 15 tap slots + root
#define L1_TW_LD 20                   // tap-weight table row stride (floats): consecutive classes hit distinct bank quads
#define L1_LDB 16                     // W_cat row stride; column n of row k is stored at n ^ (((k >> 1) & 1) << 3): the
                                      // four k rows of a B fragment then fall into four distinct 8-bank groups
#define L1_LDC 20                     // C tile row stride

// ------------------------------------------------------------------------------------------------ edge classes
// 16 lanes per destination node: lane k takes CSR slots r0+k, r0+k+16, ... so that src / cls accesses are coalesced
__global__ void __launch_bounds__(256) edge_class_kernel(const int32_t* __restrict__ rowptr, const int32_t* __restrict__ src,
                                                         int64_t n, EvadPseudo P, EvadSplineMode M,
                                                         uint8_t* __restrict__ cls, uint32_t* __restrict__ status) {
    const int64_t t = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    const int64_t i = t >> 4;
    const int k = (int)(t & 15);
    if (i >= n) return;
    const int r0 = rowptr[i], r1 = rowptr[i + 1];
    for (int e = r0 + k; e < r1; e += 16) {
        float ax, ay;
        evad_pseudo(P, e, src[e], (int)i, &ax, &ay);
        int ix, iy;
        evad_lut_index(M, ax, ay, &ix, &iy, status);
        cls[e] = (uint8_t)(ix * M.ny + iy);
    }
}

extern "C" int evad_edge_class(const int32_t* rowptr, const int32_t* src, int64_t n, const evad_pseudo_t* pseudo_h,
                               const evad_spline_mode_t* mode_h, uint8_t* cls, uint32_t* status, void* stream) {
    if (!rowptr || !src || !pseudo_h || !mode_h || !cls || !status) return EVAD_ERR_ARG;
    if (!mode_h->use_lut || mode_h->nx * mode_h->ny > 256) return EVAD_ERR_UNSUPPORTED;
    if (n <= 0) return EVAD_OK;
    edge_class_kernel<<<evad_cdiv(n * 16, 256), 256, 0, (cudaStream_t)stream>>>(rowptr, src, n, *pseudo_h, *mode_h, cls, status);
    EVAD_CHECK_LAUNCH();
    return EVAD_OK;
}

__device__ __forceinline__ uint32_t l1_tf32(float x) {
    uint32_t r;
    asm("cvt.rna.tf32.f32 %0, %1;" : "=r"(r) : "f"(x));
    return r;
}

__device__ __forceinline__ void l1_mma(float (&d)[4], const uint32_t (&a)[4], const uint32_t (&b)[2]) {
    asm volatile(
        "mma.sync.aligned.m16n8k8.row.col.f32.tf32.tf32.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
        : "+f"(d[0]), "+f"(d[1]), "+f"(d[2]), "+f"(d[3])
        : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(b[0]), "r"(b[1]));
}

// A fragment (rows g / g+8 of an m16 tile, columns kk+t / kk+t+4) split into TF32 hi / lo parts
__device__ __forceinline__ void l1_load_a(const float* __restrict__ As, int lda, int row, int kk, int g, int t,
                                          uint32_t (&ah)[4], uint32_t (&al)[4]) {
    const float* p = As + (row + g) * lda + kk + t;
    const float v[4] = {p[0], p[8 * lda], p[4], p[8 * lda + 4]};
#pragma unroll
    for (int q = 0; q < 4; q++) {                            // hi = upper 19 bits (what the tensor core reads); lo = exact rest,
        ah[q] = __float_as_uint(v[q]) & 0xffffe000u;         // of which the MMA again reads the upper 19 bits
        al[q] = __float_as_uint(__fsub_rn(v[q], __uint_as_float(ah[q])));
    }
}

__device__ __forceinline__ float4 l1_epi(const evad_epilogue_t& E, float4 v, int64_t row, int c0) {
    float r[4] = {v.x, v.y, v.z, v.w};
#pragma unroll
    for (int b = 0; b < 4; b++) {
        if (E.scale) r[b] = __fmul_rn(r[b], E.scale[c0 + b]);
        if (E.shift) r[b] = __fadd_rn(r[b], E.shift[c0 + b]);
        if (E.addend) r[b] = __fadd_rn(r[b], E.addend[row * E.ld_add + c0 + b]);
        if (E.relu) r[b] = fmaxf(r[b], 0.f);
    }
    return make_float4(r[0], r[1], r[2], r[3]);
}

template <int CP>
struct L1Cfg {
    static constexpr int QN = CP / 4;                        // channel quads per node = threads per node
    static constexpr int THREADS = L1_NODES * QN;
    static constexpr int KR = 8 * CP;                        // P columns staged per round (8 slots)
    static constexpr int LDA = KR + ((20 - KR % 32) + 32) % 32;   // == 20 (mod 32): conflict-free fragment loads
    static constexpr int KTOT = L1_SLOTS * CP;
    static constexpr int KS = (CP + 7) / 8 * 8;              // rows of the skip weight (zero padded)
    static constexpr int A_FLOATS = L1_NODES * LDA;
    static constexpr int B_FLOATS = KTOT * L1_LDB;
    static constexpr int S_FLOATS = KS * L1_LDB;
    static constexpr int C_FLOATS = 3 * L1_NODES * L1_LDC;
    static size_t smem_bytes(int ncls) {
        return (size_t)(A_FLOATS + 2 * B_FLOATS + 2 * S_FLOATS + C_FLOATS + ncls * L1_TW_LD) * sizeof(float);
    }
};

// one edge of the tap aggregation: acc[s] += tapw[class][s] * x_src for the 15 tap slots
__device__ __forceinline__ void l1_edge_fma(float4 (&acc)[L1_SLOTS - 1], const float* __restrict__ twr, const float4 xq) {
#pragma unroll
    for (int u = 0; u < 4; u++) {
        const float4 w4 = *reinterpret_cast<const float4*>(twr + 4 * u);
        const float w[4] = {w4.x, w4.y, w4.z, w4.w};
#pragma unroll
        for (int b = 0; b < 4; b++) {
            const int s = 4 * u + b;
            if (s < L1_SLOTS - 1) {
                acc[s].x = fmaf(w[b], xq.x, acc[s].x); acc[s].y = fmaf(w[b], xq.y, acc[s].y);
                acc[s].z = fmaf(w[b], xq.z, acc[s].z); acc[s].w = fmaf(w[b], xq.w, acc[s].w);
            }
        }
    }
}

__device__ __forceinline__ void l1_cp16(uint32_t dst, const float* src) {
    asm volatile("cp.async.ca.shared.global [%0], [%1], 16;" ::"r"(dst), "l"(src) : "memory");
}
__device__ __forceinline__ void l1_cp_commit() { asm volatile("cp.async.commit_group;" ::: "memory"); }
template <int N>
__device__ __forceinline__ void l1_cp_wait() { asm volatile("cp.async.wait_group %0;" ::"n"(N) : "memory"); }

// ROWS16: the graph comes as fixed-stride rows of 16 slots (evad_graph_build_rows): source ids are read four at a time
// (one 128-bit load per group) and the class bytes one word per group, instead of two scalar loads per edge and thread.
template <int CP, bool ROWS16>
__global__ void __launch_bounds__(L1Cfg<CP>::THREADS, 2) spline_l1_kernel(
    const int32_t* __restrict__ rowptr, const int32_t* __restrict__ src, const uint8_t* __restrict__ cls,
    const int32_t* __restrict__ perm, int ell_stride, int64_t n, const float* __restrict__ x, int ldx, const float* __restrict__ tapw, int ncls, const float* __restrict__ Whi,
    const float* __restrict__ Wlo, const float* __restrict__ Shi, const float* __restrict__ Slo, evad_epilogue_t E,
    float* __restrict__ out, int ldo, evad_epilogue_t ES, float* __restrict__ skip_out, int ld_skip) {
    using C = L1Cfg<CP>;
    constexpr int QN = C::QN, LDA = C::LDA, KR = C::KR;
    constexpr bool RING = ROWS16 && CP <= 16;               // cp.async row ring: pays where the register budget allows (128 regs)
    extern __shared__ __align__(16) float l1_smem[];
    float* As = l1_smem;
    float* Bh = As + C::A_FLOATS;
    float* Bl = Bh + C::B_FLOATS;
    float* Sh = Bl + C::B_FLOATS;
    float* Sl = Sh + C::S_FLOATS;
    float* Cs = Sl + C::S_FLOATS;
    float* tw = Cs + C::C_FLOATS;
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int g = lane >> 2, t = lane & 3;
    const bool has_skip = Shi != nullptr;

    // one-time staging: weights (hi / lo), skip weights, tap-weight table; A tile zeroed (its pad columns stay zero)
    for (int idx = tid; idx < C::KTOT * 4; idx += C::THREADS) {
        const int r = idx >> 2, c4 = idx & 3;
        const int cs = (4 * c4) ^ (((r >> 1) & 1) << 3);
        *reinterpret_cast<float4*>(Bh + r * L1_LDB + cs) = __ldg(reinterpret_cast<const float4*>(Whi) + idx);
        *reinterpret_cast<float4*>(Bl + r * L1_LDB + cs) = __ldg(reinterpret_cast<const float4*>(Wlo) + idx);
    }
    if (has_skip)
        for (int idx = tid; idx < C::KS * 4; idx += C::THREADS) {
            const int r = idx >> 2, c4 = idx & 3;
            const int cs = (4 * c4) ^ (((r >> 1) & 1) << 3);
            *reinterpret_cast<float4*>(Sh + r * L1_LDB + cs) = __ldg(reinterpret_cast<const float4*>(Shi) + idx);
            *reinterpret_cast<float4*>(Sl + r * L1_LDB + cs) = __ldg(reinterpret_cast<const float4*>(Slo) + idx);
        }
    for (int idx = tid; idx < ncls * 4; idx += C::THREADS) {
        const int r = idx >> 2, c4 = idx & 3;
        *reinterpret_cast<float4*>(tw + r * L1_TW_LD + 4 * c4) = __ldg(reinterpret_cast<const float4*>(tapw) + idx);
    }
    for (int idx = tid; idx < C::A_FLOATS; idx += C::THREADS) As[idx] = 0.f;
    __syncthreads();

    const int node_l = tid / QN, q = tid - node_l * QN;
    const int mt = warp & (L1_NODES / 16 - 1), kp = warp / (L1_NODES / 16);   // MMA role of the first warps: m16 tile, k half
    const int swz = ((t >> 1) & 1) << 3;                     // column swizzle of this lane's B-fragment rows
    const int64_t ntiles = (n + L1_NODES - 1) / L1_NODES;
    // ROWS16: the head of a node's row (its id, length, first four sources / classes and its own features) is a chain
    // of three dependent loads; it is fetched for the NEXT tile while this tile's contraction runs (hd_*).
    int hd_i = 0, hd_deg = 0;
    int4 hd_s0 = make_int4(0, 0, 0, 0);
    uint32_t hd_c0 = 0;
    float4 hd_xroot = make_float4(0.f, 0.f, 0.f, 0.f);
    if (ROWS16) {
        const int64_t ip0 = (int64_t)blockIdx.x * L1_NODES + node_l;
        if (ip0 < n) {
            hd_i = perm ? perm[ip0] : (int)ip0;
            hd_deg = rowptr[hd_i];
            hd_s0 = __ldg(reinterpret_cast<const int4*>(src) + (int64_t)hd_i * 4);
            hd_c0 = __ldg(reinterpret_cast<const uint32_t*>(cls) + (int64_t)hd_i * 4);
            hd_xroot = __ldg(reinterpret_cast<const float4*>(x + (int64_t)hd_i * ldx) + q);
        }
    }
    for (int64_t tile = blockIdx.x; tile < ntiles; tile += gridDim.x) {
        // ---------------------------------------------------------------- (1) tap aggregation in registers
        const int64_t ip = tile * L1_NODES + node_l;          // position in processing order (pixel order when perm)
        const bool valid = ip < n;
        const int64_t i = ROWS16 ? (int64_t)hd_i : valid ? (perm ? (int64_t)perm[ip] : ip) : 0;
        const int64_t ip_n = ip + (int64_t)gridDim.x * L1_NODES;   // this thread's node of the block's next tile
        float4 acc[L1_SLOTS - 1];
#pragma unroll
        for (int s = 0; s < L1_SLOTS - 1; s++) acc[s] = make_float4(0.f, 0.f, 0.f, 0.f);
        float4 xroot = make_float4(0.f, 0.f, 0.f, 0.f);
        if (valid && ROWS16) {
            const int deg = hd_deg;
            const int4* srow = reinterpret_cast<const int4*>(src) + i * 4;
            const uint32_t* crow = reinterpret_cast<const uint32_t*>(cls) + i * 4;
            xroot = hd_xroot;
            if (deg > 0 && RING) {
                // Source rows travel global -> shared by cp.async into a 4-slot ring made of this thread's own cells of
                // the A tile (idle until the rounds below, and rewritten there), four edges ahead of their use: the
                // gather latency is covered without holding the rows in registers.
                float* ring = As + node_l * LDA + 4 * q;
                const uint32_t ring_s = (uint32_t)__cvta_generic_to_shared(ring);
                const float* xq_base = x + 4 * q;
                int4 s_next = hd_s0;
                uint32_t c_cur = hd_c0;
                {
                    const int s0[4] = {s_next.x, s_next.y, s_next.z, s_next.w};
#pragma unroll
                    for (int k = 0; k < 4; k++) {
                        if (k < deg) l1_cp16(ring_s + k * CP * 4, xq_base + (int64_t)s0[k] * ldx);
                        l1_cp_commit();
                    }
                }
                if (deg > 4) s_next = __ldg(srow + 1);
#pragma unroll 1
                for (int g = 0; 4 * g < deg; g++) {
                    int4 s_nn = s_next;
                    uint32_t c_next = c_cur;
                    if (4 * (g + 1) < deg) c_next = __ldg(crow + g + 1);
                    if (4 * (g + 2) < deg) s_nn = __ldg(srow + g + 2);
                    const int sj[4] = {s_next.x, s_next.y, s_next.z, s_next.w};
#pragma unroll
                    for (int k = 0; k < 4; k++) {
                        const int e = 4 * g + k;
                        if (e >= deg) break;
                        l1_cp_wait<3>();
                        const float4 xq = *reinterpret_cast<const float4*>(ring + k * CP);
                        l1_edge_fma(acc, tw + ((c_cur >> (8 * k)) & 0xffu) * L1_TW_LD, xq);
                        if (e + 4 < deg) l1_cp16(ring_s + k * CP * 4, xq_base + (int64_t)sj[k] * ldx);
                        l1_cp_commit();
                    }
                    s_next = s_nn;
                    c_cur = c_next;
                }
            }
            if (deg > 0 && !RING) {
                // rows prefetched one edge ahead in registers
                int4 s_cur = hd_s0;
                uint32_t c_cur = hd_c0;
                float4 x_next = __ldg(reinterpret_cast<const float4*>(x + (int64_t)s_cur.x * ldx) + q);
#pragma unroll 1
                for (int g = 0; 4 * g < deg; g++) {
                    int4 s_next = s_cur;
                    uint32_t c_next = c_cur;
                    if (4 * (g + 1) < deg) {                  // next group of source ids / class bytes: in flight for 4 edges
                        s_next = __ldg(srow + g + 1);
                        c_next = __ldg(crow + g + 1);
                    }
                    const int sj[5] = {s_cur.x, s_cur.y, s_cur.z, s_cur.w, s_next.x};
#pragma unroll
                    for (int k = 0; k < 4; k++) {
                        const int e = 4 * g + k;
                        if (e >= deg) break;
                        const float4 xq = x_next;
                        if (e + 1 < deg) x_next = __ldg(reinterpret_cast<const float4*>(x + (int64_t)sj[k + 1] * ldx) + q);
                        l1_edge_fma(acc, tw + ((c_cur >> (8 * k)) & 0xffu) * L1_TW_LD, xq);
                    }
                    s_cur = s_next;
                    c_cur = c_next;
                }
            }
        } else if (valid) {
            // CSR (rowptr[n+1]) or fixed-stride rows (ell_stride > 0: rowptr[i] = length of row i at i * ell_stride)
            const int r0 = ell_stride > 0 ? (int)i * ell_stride : rowptr[i];
            const int deg = ell_stride > 0 ? rowptr[i] : rowptr[i + 1] - r0;
            xroot = __ldg(reinterpret_cast<const float4*>(x + i * ldx) + q);
            // software pipeline without branches: the class and row of edge e+1 and the source id of edge e+2 are in
            // flight while the 60 FMAs of edge e run; indices are clamped to the last edge (a redundant, valid load)
            const int last = r0 + max(deg - 1, 0);
            int c_next = 0, j_n2 = 0;
            float4 x_next = make_float4(0.f, 0.f, 0.f, 0.f);
            if (deg > 0) {
                c_next = cls[r0];
                x_next = __ldg(reinterpret_cast<const float4*>(x + (int64_t)src[r0] * ldx) + q);
                j_n2 = src[min(r0 + 1, last)];
            }
            for (int e = 0; e < deg; e++) {
                const float4 xq = x_next;
                const float* twr = tw + c_next * L1_TW_LD;
                c_next = cls[min(r0 + e + 1, last)];
                x_next = __ldg(reinterpret_cast<const float4*>(x + (int64_t)j_n2 * ldx) + q);
                j_n2 = src[min(r0 + e + 2, last)];
                l1_edge_fma(acc, twr, xq);
            }
        }
        if (ROWS16) hd_i = ip_n < n ? (perm ? perm[ip_n] : (int)ip_n) : -1;   // next tile's node: in flight over round 0
        // ---------------------------------------------------------------- (2) contraction, two rounds of 8 slots
        float cacc[2][4] = {{0.f, 0.f, 0.f, 0.f}, {0.f, 0.f, 0.f, 0.f}};
        float sacc[2][4] = {{0.f, 0.f, 0.f, 0.f}, {0.f, 0.f, 0.f, 0.f}};
#pragma unroll
        for (int round = 0; round < 2; round++) {
            float* arow = As + node_l * LDA + 4 * q;
#pragma unroll
            for (int s = 0; s < 8; s++) {
                const int slot = round * 8 + s;
                *reinterpret_cast<float4*>(arow + s * CP) = slot < L1_SLOTS - 1 ? acc[slot < L1_SLOTS - 1 ? slot : 0] : xroot;
            }
            if (ROWS16 && round == 1) {                      // next tile's row head: in flight over round 1 (accumulators are dead now)
                hd_deg = 0;
                if (hd_i >= 0) {
                    hd_deg = rowptr[hd_i];
                    hd_s0 = __ldg(reinterpret_cast<const int4*>(src) + (int64_t)hd_i * 4);
                    hd_c0 = __ldg(reinterpret_cast<const uint32_t*>(cls) + (int64_t)hd_i * 4);
                    hd_xroot = __ldg(reinterpret_cast<const float4*>(x + (int64_t)hd_i * ldx) + q);
                } else {
                    hd_i = 0;
                }
            }
            __syncthreads();
            if (warp < L1_MMA_WARPS) {
                float part[2][4] = {{0.f, 0.f, 0.f, 0.f}, {0.f, 0.f, 0.f, 0.f}};
#pragma unroll 2
                for (int s = 0; s < CP / 2; s++) {
                    const int kk = (kp * (CP / 2) + s) * 8;
                    uint32_t ah[4], al[4];
                    l1_load_a(As, LDA, mt * 16, kk, g, t, ah, al);
                    const int krow = round * KR + kk;
#pragma unroll
                    for (int nt = 0; nt < 2; nt++) {
                        const int o0 = (krow + t) * L1_LDB + ((nt * 8 + g) ^ swz), o1 = o0 + 4 * L1_LDB;
                        const uint32_t bh[2] = {__float_as_uint(Bh[o0]), __float_as_uint(Bh[o1])};
                        const uint32_t bl[2] = {__float_as_uint(Bl[o0]), __float_as_uint(Bl[o1])};
                        l1_mma(part[nt], al, bh);
                        l1_mma(part[nt], ah, bl);
                        l1_mma(part[nt], ah, bh);
                    }
                }
#pragma unroll
                for (int nt = 0; nt < 2; nt++)
#pragma unroll
                    for (int b = 0; b < 4; b++) cacc[nt][b] += part[nt][b];
                if (round == 1 && has_skip && kp == 1) {     // skip Linear on the root slot (slot 7 of this round)
#pragma unroll
                    for (int s = 0; s < C::KS / 8; s++) {
                        uint32_t ah[4], al[4];
                        l1_load_a(As, LDA, mt * 16, 7 * CP + 8 * s, g, t, ah, al);
#pragma unroll
                        for (int nt = 0; nt < 2; nt++) {
                            const int o0 = (8 * s + t) * L1_LDB + ((nt * 8 + g) ^ swz), o1 = o0 + 4 * L1_LDB;
                            const uint32_t bh[2] = {__float_as_uint(Sh[o0]), __float_as_uint(Sh[o1])};
                            const uint32_t bl[2] = {__float_as_uint(Sl[o0]), __float_as_uint(Sl[o1])};
                            l1_mma(sacc[nt], al, bh);
                            l1_mma(sacc[nt], ah, bl);
                            l1_mma(sacc[nt], ah, bh);
                        }
                    }
                }
                if (round == 1) {
#pragma unroll
                    for (int nt = 0; nt < 2; nt++)
#pragma unroll
                        for (int h = 0; h < 2; h++) {
                            const int row = mt * 16 + g + 8 * h, col = nt * 8 + 2 * t;
                            *reinterpret_cast<float2*>(Cs + (kp * L1_NODES + row) * L1_LDC + col) =
                                make_float2(cacc[nt][2 * h], cacc[nt][2 * h + 1]);
                            if (has_skip && kp == 1)
                                *reinterpret_cast<float2*>(Cs + (2 * L1_NODES + row) * L1_LDC + col) =
                                    make_float2(sacc[nt][2 * h], sacc[nt][2 * h + 1]);
                        }
                }
            }
            __syncthreads();
        }
        // ---------------------------------------------------------------- (3) epilogue: BN / skip add / ReLU
        if (tid < 4 * L1_NODES) {
            const int nl = tid >> 2, oq = tid & 3;
            const int64_t iop = tile * L1_NODES + nl;
            if (iop < n) {
                const int64_t io = perm ? (int64_t)perm[iop] : iop;
                const float4 c0 = *reinterpret_cast<const float4*>(Cs + nl * L1_LDC + 4 * oq);
                const float4 c1 = *reinterpret_cast<const float4*>(Cs + (L1_NODES + nl) * L1_LDC + 4 * oq);
                const float4 v = make_float4(c0.x + c1.x, c0.y + c1.y, c0.z + c1.z, c0.w + c1.w);
                *reinterpret_cast<float4*>(out + io * ldo + 4 * oq) = l1_epi(E, v, io, 4 * oq);
                if (has_skip) {
                    const float4 sv = *reinterpret_cast<const float4*>(Cs + (2 * L1_NODES + nl) * L1_LDC + 4 * oq);
                    *reinterpret_cast<float4*>(skip_out + io * ld_skip + 4 * oq) = l1_epi(ES, sv, io, 4 * oq);
                }
            }
        }
    }
}

template <int CP, bool ROWS16>
static int launch_l1(const int32_t* rowptr, const int32_t* src, const uint8_t* cls, const int32_t* perm, int ell_stride, int64_t n,
                     const float* x, int ldx,
                     const float* tapw, int ncls, const float* Whi, const float* Wlo, const float* Shi, const float* Slo,
                     const evad_epilogue_t& E, float* out, int ldo, const evad_epilogue_t& ES, float* skip_out,
                     int ld_skip, cudaStream_t st) {
    const size_t smem = L1Cfg<CP>::smem_bytes(ncls);
    if (smem > 113 * 1024) return EVAD_ERR_UNSUPPORTED;      // two CTAs per SM
    static size_t attr_bytes = 0;
    if (smem > attr_bytes) {
        if (cudaFuncSetAttribute(spline_l1_kernel<CP, ROWS16>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem) != cudaSuccess)
            return EVAD_ERR_LAUNCH;
        attr_bytes = smem;
    }
    const int64_t ntiles = (n + L1_NODES - 1) / L1_NODES;
    const int grid = (int)(ntiles < 2 * 148 ? ntiles : 2 * 148);
    spline_l1_kernel<CP, ROWS16><<<grid, L1Cfg<CP>::THREADS, smem, st>>>(rowptr, src, cls, perm, ell_stride, n, x, ldx, tapw, ncls, Whi, Wlo, Shi, Slo,
                                                                E, out, ldo, ES, skip_out, ld_skip);
    EVAD_CHECK_LAUNCH();
    return EVAD_OK;
}

extern "C" int evad_spline_conv_l1(const int32_t* rowptr, const int32_t* src, const uint8_t* cls, const int32_t* perm,
                                   int ell_stride, int64_t n, const float* x, int ldx, int cin_pad, const float* tapw, int ncls, const float* Whi,
                                   const float* Wlo, const float* Shi, const float* Slo, const evad_epilogue_t* epi_h,
                                   float* out, int ldo, const evad_epilogue_t* skip_epi_h, float* skip_out, int ld_skip,
                                   void* stream) {
    if (!rowptr || !src || !cls || !x || !tapw || !Whi || !Wlo || !epi_h || !out) return EVAD_ERR_ARG;
    if (cin_pad != 16 && cin_pad != 20) return EVAD_ERR_UNSUPPORTED;
    if (ell_stride < 0 || (ell_stride > 0 && n * ell_stride >= (1ll << 31))) return EVAD_ERR_ARG;
    if (ldx < cin_pad || ldx % 4 != 0 || ldo % 4 != 0 || ncls < 1 || ncls > 256) return EVAD_ERR_ARG;
    if ((((uintptr_t)x | (uintptr_t)tapw | (uintptr_t)Whi | (uintptr_t)Wlo | (uintptr_t)out) & 15) != 0) return EVAD_ERR_ARG;
    const bool has_skip = Shi && Slo && skip_epi_h && skip_out;
    if (has_skip && ((((uintptr_t)Shi | (uintptr_t)Slo | (uintptr_t)skip_out) & 15) != 0 || ld_skip % 4 != 0))
        return EVAD_ERR_ARG;
    if (n <= 0) return EVAD_OK;
    const evad_epilogue_t es = has_skip ? *skip_epi_h : *epi_h;
    cudaStream_t st = (cudaStream_t)stream;
    const bool rows16 = ell_stride == 16 && (((uintptr_t)src | (uintptr_t)cls) & 15) == 0;
#define L1_LAUNCH(CP_, R_)                                                                                              \
    launch_l1<CP_, R_>(rowptr, src, cls, perm, ell_stride, n, x, ldx, tapw, ncls, Whi, Wlo, has_skip ? Shi : nullptr,   \
                       has_skip ? Slo : nullptr, *epi_h, out, ldo, es, skip_out, ld_skip, st)
    if (cin_pad == 20) return rows16 ? L1_LAUNCH(20, true) : L1_LAUNCH(20, false);
    return rows16 ? L1_LAUNCH(16, true) : L1_LAUNCH(16, false);
#undef L1_LAUNCH
}

// ================================================================================================================
// Variant with the tap aggregation on the tensor cores as well (fixed-stride rows of 16 only).
//
// Per destination node the aggregation P_i[slot][c] = sum_e tapw[class_e][slot] * x_{src_e}[c] is itself a small
// product: [16 slots x 16 row positions] @ [16 row positions x Cin].  One warp computes it per node with
// mma.sync.m16n8k8 (M = slots, K = row positions in two steps of 8, N = channels in tiles of 8):
//   A fragment = tap weights of the lane's four row positions (t, t+4, t+8, t+12), read from a shared table laid out
//     as (slot g, slot g+8) pairs; both operands are split into TF32 hi / lo parts in registers (3xTF32): three MMAs
//     per (k step, n tile) -- 12 (Cin 16) or 18 (Cin 20) per node where the register variant above issues
//     16 x QN x 60 FFMA;
//   B fragment = the source rows gathered straight from global memory.  The MMA column -> channel map is free, so
//     column j of tiles 0 / 1 is channel 2j / 2j+1: a lane fetches both with one 64-bit load and a row is 64
//     contiguous bytes per request; tile 2 (Cin 20) is channels 16..19.
//   D fragment: lane (g, t) ends up with slots g and g+8 of channels 4t..4t+3, i.e. one 128-bit store each into the
//     [32 nodes x 16*Cin] shared P tile; slot 15 is overwritten with the node's own row (root weight).
// The tile is 32 nodes (four per warp) with all 16 slots resident, so the contraction is a single pass split four ways
// along K over the eight warps (two m16 tiles x four k quarters); partial sums are added in the epilogue.
// Shared-memory layouts are chosen for the fragment loads of the contraction (ncu: the L1 data pipe bounds this kernel):
//   P tile columns: [16 slots][channels 0..15] then [16 slots][channels 16..19] (slot stride 16 / 4: the D-fragment
//     stores are conflict free), row stride == 8 (mod 32); a lane's two k columns of a k step are ADJACENT (k order
//     inside a step is free), so an A fragment is two 64-bit loads;
//   weights: fragment major, one float4 (hi k0, hi k1, lo k0, lo k1) per (k step, n tile, lane): a B fragment with both
//     TF32 parts is one conflict-free 128-bit load.
#define L1M_NODES 32
#define L1M_THREADS 256
#define L1M_TW_LD 24                  // floats per class row of the paired tap table: 8 x (slot g, slot g+8) + pad
#define L1M_AUX 128                   // words: node ids of the tile [2 parities][32] + epilogue scale / shift [4][16]

template <int CP>
struct L1MCfg {
    static constexpr int KTOT = L1_SLOTS * CP;
    static constexpr int LDA = KTOT + ((8 - KTOT % 32) + 32) % 32;    // == 8 (mod 32)
    static constexpr int KSTEPS = KTOT / 8;
    static constexpr int SSTEPS = (CP + 7) / 8;              // k steps of the skip Linear
    static constexpr int NT = (CP + 7) / 8;                  // channel tiles of the aggregation MMA
    static constexpr int A_FLOATS = L1M_NODES * LDA;
    static constexpr int B_FLOATS = KSTEPS * 2 * 32 * 4;
    static constexpr int S_FLOATS = SSTEPS * 2 * 32 * 4;
    static constexpr int C_FLOATS = 5 * L1M_NODES * L1_LDC;  // four k-quarter partials + skip
    static size_t smem_bytes(int ncls) {
        return (size_t)(A_FLOATS + B_FLOATS + S_FLOATS + C_FLOATS + (ncls + 1) * L1M_TW_LD + L1M_AUX) * sizeof(float);
    }
    // P tile column -> row of W_cat ([slot][channel], channel stride CP)
    __host__ __device__ static constexpr int wrow(int k) {
        return k < 256 ? (k >> 4) * CP + (k & 15) : ((k - 256) >> 2) * CP + 16 + ((k - 256) & 3);
    }
};

struct L1MHead {                      // what a lane keeps of one node's row: id, length, its 4 sources and their classes
    int i, deg;
    int s[4];
    uint32_t c;                       // class bytes of row positions t, t+4, t+8, t+12
};

__device__ __forceinline__ void l1m_mma(float (&d)[4], const uint32_t (&a)[4], uint32_t b0, uint32_t b1) {
    asm("mma.sync.aligned.m16n8k8.row.col.f32.tf32.tf32.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
        : "+f"(d[0]), "+f"(d[1]), "+f"(d[2]), "+f"(d[3])
        : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(b0), "r"(b1));
}

// A fragment of the contraction: rows g / g+8, the lane's two adjacent columns of the k step at kk
__device__ __forceinline__ void l1m_load_a(const float* __restrict__ As, int lda, int row, int kk, int g, int t,
                                           uint32_t (&ah)[4], uint32_t (&al)[4]) {
    const float2 p0 = *reinterpret_cast<const float2*>(As + (row + g) * lda + kk + 2 * t);
    const float2 p1 = *reinterpret_cast<const float2*>(As + (row + g + 8) * lda + kk + 2 * t);
    const float v[4] = {p0.x, p1.x, p0.y, p1.y};
#pragma unroll
    for (int q = 0; q < 4; q++) {
        ah[q] = __float_as_uint(v[q]) & 0xffffe000u;
        al[q] = __float_as_uint(__fsub_rn(v[q], __uint_as_float(ah[q])));
    }
}

template <int CP>
__global__ void __launch_bounds__(L1M_THREADS, 2) spline_l1_mma_kernel(
    const int32_t* __restrict__ count, const int32_t* __restrict__ src, const uint8_t* __restrict__ cls,
    const int32_t* __restrict__ perm, int64_t n, const float* __restrict__ x, int ldx, const float* __restrict__ tapw,
    int ncls, const float* __restrict__ Whi, const float* __restrict__ Wlo, const float* __restrict__ Shi,
    const float* __restrict__ Slo, evad_epilogue_t E, float* __restrict__ out, int ldo, evad_epilogue_t ES,
    float* __restrict__ skip_out, int ld_skip) {
    using C = L1MCfg<CP>;
    constexpr int LDA = C::LDA, NT = C::NT;
    extern __shared__ __align__(16) float l1_smem[];
    float* As = l1_smem;
    float4* Wf = reinterpret_cast<float4*>(As + C::A_FLOATS);
    float4* Sf = Wf + C::B_FLOATS / 4;
    float* Cs = reinterpret_cast<float*>(Sf + C::S_FLOATS / 4);
    float* tw = Cs + C::C_FLOATS;
    int* ids = reinterpret_cast<int*>(tw + (ncls + 1) * L1M_TW_LD);
    float* epc = reinterpret_cast<float*>(ids + 64);
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int g = lane >> 2, t = lane & 3;
    const bool has_skip = Shi != nullptr;

    if (tid < 64) {                                          // epilogue constants: scale (1 when absent), shift (0 when absent)
        const int c = tid & 15, w = tid >> 4;
        const float* pq = w == 0 ? E.scale : w == 1 ? E.shift : w == 2 ? ES.scale : ES.shift;
        epc[tid] = pq ? pq[c] : (w & 1) ? 0.f : 1.f;
    }
    // one-time staging.  Weight fragments: entry (ks, nt, lane) = W_cat rows of P columns 8ks+2t, 8ks+2t+1, column 8nt+g
    for (int idx = tid; idx < C::KSTEPS * 64; idx += L1M_THREADS) {
        const int ks = idx >> 6, nt = (idx >> 5) & 1, l = idx & 31;
        const int r0 = C::wrow(8 * ks + 2 * (l & 3)), r1 = C::wrow(8 * ks + 2 * (l & 3) + 1), col = 8 * nt + (l >> 2);
        Wf[idx] = make_float4(__ldg(Whi + r0 * 16 + col), __ldg(Whi + r1 * 16 + col), __ldg(Wlo + r0 * 16 + col),
                              __ldg(Wlo + r1 * 16 + col));
    }
    if (has_skip)
        for (int idx = tid; idx < C::SSTEPS * 64; idx += L1M_THREADS) {
            // skip k steps 0, 1 read the root slot's channels 0..15 (P columns 240..255); step 2 (Cin 20) reads columns
            // 312..319 = slots 14, 15 of the second block: only the slot-15 half carries skip weights
            const int s = idx >> 6, nt = (idx >> 5) & 1, l = idx & 31, col = 8 * nt + (l >> 2);
            float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
            const int k0 = 2 * (l & 3);
            if (s < 2) {
                const int c0 = 8 * s + k0;
                v = make_float4(__ldg(Shi + c0 * 16 + col), __ldg(Shi + (c0 + 1) * 16 + col), __ldg(Slo + c0 * 16 + col),
                                __ldg(Slo + (c0 + 1) * 16 + col));
            } else if (k0 >= 4) {
                const int c0 = 16 + k0 - 4;
                v = make_float4(__ldg(Shi + c0 * 16 + col), __ldg(Shi + (c0 + 1) * 16 + col), __ldg(Slo + c0 * 16 + col),
                                __ldg(Slo + (c0 + 1) * 16 + col));
            }
            Sf[idx] = v;
        }
    // paired tap table: row c = (tapw[c][0], tapw[c][8]), (tapw[c][1], tapw[c][9]), ...; row ncls = zeros (empty slots)
    for (int idx = tid; idx < (ncls + 1) * 8; idx += L1M_THREADS) {
        const int r = idx >> 3, k = idx & 7;
        float2 v = make_float2(0.f, 0.f);
        if (r < ncls) v = make_float2(__ldg(tapw + r * L1_SLOTS + k), __ldg(tapw + r * L1_SLOTS + k + 8));
        *reinterpret_cast<float2*>(tw + r * L1M_TW_LD + 2 * k) = v;
    }
    for (int idx = tid; idx < C::A_FLOATS; idx += L1M_THREADS) As[idx] = 0.f;
    __syncthreads();

    const int mt = warp & 1, kq = warp >> 1;                 // contraction role: m16 tile, k quarter
    const int64_t ntiles = (n + L1M_NODES - 1) / L1M_NODES;
    const uint32_t tw_zero = (uint32_t)ncls;
    const float* xg = x + 2 * g;                             // lane bases of the row gathers (offsets fit 32 bits)
    const float* xg16 = x + 16 + g;
    const float* xt = x + 4 * t;
    const uint32_t uldx = (uint32_t)ldx;

    // Row heads travel in three steps so that no load is waited for where it is issued: node ids (perm) two tiles
    // ahead, row length / sources / class words one tile ahead (issued after the aggregation, from the ids that are
    // already there), class bytes packed and the first node's rows gathered in the middle of the contraction.
    L1MHead hd[4];
    uint4 craw[4];
    int inext[4];
    auto load_ids = [&](int64_t tile) {
#pragma unroll
        for (int k = 0; k < 4; k++) {
            const int64_t ip = tile * L1M_NODES + warp * 4 + k;
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
            for (int j = 0; j < 4; j++) hd[k].s[j] = 0;
            if (i >= 0) {
                hd[k].deg = count[i];
                craw[k] = __ldg(reinterpret_cast<const uint4*>(cls) + i);
                const int32_t* srow = src + ((uint32_t)i * 16u + (uint32_t)t);
#pragma unroll
                for (int j = 0; j < 4; j++) hd[k].s[j] = __ldg(srow + 4 * j);
            }
        }
    };
    auto pack_classes = [&]() {
#pragma unroll
        for (int k = 0; k < 4; k++)
            hd[k].c = __byte_perm(__byte_perm(craw[k].x, craw[k].y, t | ((4 + t) << 4)),
                                  __byte_perm(craw[k].z, craw[k].w, t | ((4 + t) << 4)), 0x5410);
    };
    float2 xa[2][4];                                         // [buffer][row position j]: channels 2g, 2g+1
    float xb[2][4];                                          // channel 16 + g (Cin 20)
    float4 xr[2];                                            // the node's own row (lanes g == 7): channels 4t..4t+3
    float2 xr2[2];                                           //   and 16+2t, 17+2t
    auto gather = [&](int k, int buf) {
        const int i = hd[k].i < 0 ? 0 : hd[k].i;
#pragma unroll
        for (int j = 0; j < 4; j++) {
            const int sj = (4 * j + t) < hd[k].deg ? hd[k].s[j] : i;   // empty slots read the node's own (valid) row
            const uint32_t off = (uint32_t)sj * uldx;
            xa[buf][j] = __ldg(reinterpret_cast<const float2*>(xg + off));
            if (CP > 16) xb[buf][j] = g < 4 ? __ldg(xg16 + off) : 0.f;
        }
        xr[buf] = make_float4(0.f, 0.f, 0.f, 0.f);
        xr2[buf] = make_float2(0.f, 0.f);
        if (g == 7 && hd[k].i >= 0) {
            const uint32_t off = (uint32_t)i * uldx;
            xr[buf] = __ldg(reinterpret_cast<const float4*>(xt + off));
            if (CP > 16 && t < 2) xr2[buf] = __ldg(reinterpret_cast<const float2*>(x + off + 16 + 2 * t));
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
#pragma unroll
        for (int k = 0; k < 4; k++) {
            const int buf = k & 1;
            if (k + 1 < 4) gather(k + 1, buf ^ 1);
            float d[NT][4];
#pragma unroll
            for (int nt = 0; nt < NT; nt++)
#pragma unroll
                for (int b = 0; b < 4; b++) d[nt][b] = 0.f;
#pragma unroll
            for (int ks = 0; ks < 2; ks++) {
                uint32_t ah[4], al[4];
#pragma unroll
                for (int h = 0; h < 2; h++) {
                    const int j = 2 * ks + h;
                    const uint32_t c = (4 * j + t) < hd[k].deg ? ((hd[k].c >> (8 * j)) & 0xffu) : tw_zero;
                    const float2 w = *reinterpret_cast<const float2*>(tw + c * L1M_TW_LD + 2 * g);
                    ah[2 * h] = __float_as_uint(w.x) & 0xffffe000u;
                    ah[2 * h + 1] = __float_as_uint(w.y) & 0xffffe000u;
                    al[2 * h] = __float_as_uint(__fsub_rn(w.x, __uint_as_float(ah[2 * h])));
                    al[2 * h + 1] = __float_as_uint(__fsub_rn(w.y, __uint_as_float(ah[2 * h + 1])));
                }
                uint32_t bh[NT][2], bl[NT][2];
#pragma unroll
                for (int nt = 0; nt < NT; nt++)
#pragma unroll
                    for (int h = 0; h < 2; h++) {
                        const int j = 2 * ks + h;
                        const float v = nt == 0 ? xa[buf][j].x : nt == 1 ? xa[buf][j].y : xb[buf][j];
                        bh[nt][h] = __float_as_uint(v) & 0xffffe000u;
                        bl[nt][h] = __float_as_uint(__fsub_rn(v, __uint_as_float(bh[nt][h])));
                    }
                // product by product over the channel tiles: consecutive MMAs are independent
#pragma unroll
                for (int nt = 0; nt < NT; nt++) l1m_mma(d[nt], al, bh[nt][0], bh[nt][1]);
#pragma unroll
                for (int nt = 0; nt < NT; nt++) l1m_mma(d[nt], ah, bl[nt][0], bl[nt][1]);
#pragma unroll
                for (int nt = 0; nt < NT; nt++) l1m_mma(d[nt], ah, bh[nt][0], bh[nt][1]);
            }
            float* arow = As + (warp * 4 + k) * LDA;
            float4 v0 = make_float4(d[0][0], d[1][0], d[0][1], d[1][1]);
            float4 v1 = make_float4(d[0][2], d[1][2], d[0][3], d[1][3]);
            if (g == 7) v1 = xr[buf];
            *reinterpret_cast<float4*>(arow + g * 16 + 4 * t) = v0;
            *reinterpret_cast<float4*>(arow + (g + 8) * 16 + 4 * t) = v1;
            if (CP > 16 && t < 2) {
                float2 u0 = make_float2(d[NT - 1][0], d[NT - 1][1]);
                float2 u1 = make_float2(d[NT - 1][2], d[NT - 1][3]);
                if (g == 7) u1 = xr2[buf];
                *reinterpret_cast<float2*>(arow + 256 + g * 4 + 2 * t) = u0;
                *reinterpret_cast<float2*>(arow + 256 + (g + 8) * 4 + 2 * t) = u1;
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
        if (tid < 4 * L1M_NODES) {
            e_io = ids[par + e_nl];
            if (e_io >= 0 && E.addend) e_add = __ldg(reinterpret_cast<const float4*>(E.addend + (int64_t)e_io * E.ld_add) + e_oq);
        }
        // ---------------------------------------------------------------- (2) contraction P tile @ W_cat, k quarter per warp
        {
            float part[2][4] = {{0.f, 0.f, 0.f, 0.f}, {0.f, 0.f, 0.f, 0.f}};
            constexpr int KQ = C::KSTEPS / 4;
#pragma unroll
            for (int s = 0; s < KQ; s++) {
                if (s == KQ / 2 && more) {                   // first node of the next tile: rows in flight over the rest
                    pack_classes();
                    gather(0, 0);
                }
                const int ks = kq * KQ + s;
                uint32_t ah[4], al[4];
                l1m_load_a(As, LDA, mt * 16, ks * 8, g, t, ah, al);
                const float4 w0 = Wf[(ks * 2 + 0) * 32 + lane], w1 = Wf[(ks * 2 + 1) * 32 + lane];
                l1m_mma(part[0], al, __float_as_uint(w0.x), __float_as_uint(w0.y));
                l1m_mma(part[1], al, __float_as_uint(w1.x), __float_as_uint(w1.y));
                l1m_mma(part[0], ah, __float_as_uint(w0.z), __float_as_uint(w0.w));
                l1m_mma(part[1], ah, __float_as_uint(w1.z), __float_as_uint(w1.w));
                l1m_mma(part[0], ah, __float_as_uint(w0.x), __float_as_uint(w0.y));
                l1m_mma(part[1], ah, __float_as_uint(w1.x), __float_as_uint(w1.y));
            }
#pragma unroll
            for (int nt = 0; nt < 2; nt++)
#pragma unroll
                for (int h = 0; h < 2; h++) {
                    const int row = mt * 16 + g + 8 * h, col = nt * 8 + 2 * t;
                    *reinterpret_cast<float2*>(Cs + (kq * L1M_NODES + row) * L1_LDC + col) =
                        make_float2(part[nt][2 * h], part[nt][2 * h + 1]);
                }
            if (has_skip && kq == 3) {                       // skip Linear on the root slot
                float sacc[2][4] = {{0.f, 0.f, 0.f, 0.f}, {0.f, 0.f, 0.f, 0.f}};
#pragma unroll
                for (int s = 0; s < C::SSTEPS; s++) {
                    uint32_t ah[4], al[4];
                    l1m_load_a(As, LDA, mt * 16, s < 2 ? 240 + 8 * s : 312, g, t, ah, al);
                    const float4 w0 = Sf[(s * 2 + 0) * 32 + lane], w1 = Sf[(s * 2 + 1) * 32 + lane];
                    l1m_mma(sacc[0], al, __float_as_uint(w0.x), __float_as_uint(w0.y));
                    l1m_mma(sacc[1], al, __float_as_uint(w1.x), __float_as_uint(w1.y));
                    l1m_mma(sacc[0], ah, __float_as_uint(w0.z), __float_as_uint(w0.w));
                    l1m_mma(sacc[1], ah, __float_as_uint(w1.z), __float_as_uint(w1.w));
                    l1m_mma(sacc[0], ah, __float_as_uint(w0.x), __float_as_uint(w0.y));
                    l1m_mma(sacc[1], ah, __float_as_uint(w1.x), __float_as_uint(w1.y));
                }
#pragma unroll
                for (int nt = 0; nt < 2; nt++)
#pragma unroll
                    for (int h = 0; h < 2; h++) {
                        const int row = mt * 16 + g + 8 * h, col = nt * 8 + 2 * t;
                        *reinterpret_cast<float2*>(Cs + (4 * L1M_NODES + row) * L1_LDC + col) =
                            make_float2(sacc[nt][2 * h], sacc[nt][2 * h + 1]);
                    }
            }
        }
        __syncthreads();
        // ---------------------------------------------------------------- (3) epilogue: BN / skip add / ReLU
        if (e_io >= 0) {
            float4 v = *reinterpret_cast<const float4*>(Cs + e_nl * L1_LDC + 4 * e_oq);
#pragma unroll
            for (int p = 1; p < 4; p++) {
                const float4 c = *reinterpret_cast<const float4*>(Cs + (p * L1M_NODES + e_nl) * L1_LDC + 4 * e_oq);
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
            *reinterpret_cast<float4*>(out + (int64_t)e_io * ldo + 4 * e_oq) = make_float4(r[0], r[1], r[2], r[3]);
            if (has_skip) {
                const float4 sv = *reinterpret_cast<const float4*>(Cs + (4 * L1M_NODES + e_nl) * L1_LDC + 4 * e_oq);
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
static int launch_l1_mma(const int32_t* count, const int32_t* src, const uint8_t* cls, const int32_t* perm, int64_t n,
                         const float* x, int ldx, const float* tapw, int ncls, const float* Whi, const float* Wlo,
                         const float* Shi, const float* Slo, const evad_epilogue_t& E, float* out, int ldo,
                         const evad_epilogue_t& ES, float* skip_out, int ld_skip, cudaStream_t st) {
    const size_t smem = L1MCfg<CP>::smem_bytes(ncls);
    if (smem > 113 * 1024) return EVAD_ERR_UNSUPPORTED;      // two CTAs per SM
    static size_t attr_bytes = 0;
    if (smem > attr_bytes) {
        if (cudaFuncSetAttribute(spline_l1_mma_kernel<CP>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem) != cudaSuccess)
            return EVAD_ERR_LAUNCH;
        attr_bytes = smem;
    }
    const int64_t ntiles = (n + L1M_NODES - 1) / L1M_NODES;
    const int grid = (int)(ntiles < 2 * 148 ? ntiles : 2 * 148);
    spline_l1_mma_kernel<CP><<<grid, L1M_THREADS, smem, st>>>(count, src, cls, perm, n, x, ldx, tapw, ncls, Whi, Wlo, Shi, Slo,
                                                               E, out, ldo, ES, skip_out, ld_skip);
    EVAD_CHECK_LAUNCH();
    return EVAD_OK;
}

extern "C" int evad_spline_conv_l1_mma(const int32_t* count, const int32_t* nbr, const uint8_t* nbr_cls, const int32_t* perm,
                                       int64_t n, const float* x, int ldx, int cin_pad, const float* tapw, int ncls,
                                       const float* Whi, const float* Wlo, const float* Shi, const float* Slo,
                                       const evad_epilogue_t* epi_h, float* out, int ldo,
                                       const evad_epilogue_t* skip_epi_h, float* skip_out, int ld_skip, void* stream) {
    if (!count || !nbr || !nbr_cls || !x || !tapw || !Whi || !Wlo || !epi_h || !out) return EVAD_ERR_ARG;
    if (cin_pad != 16 && cin_pad != 20) return EVAD_ERR_UNSUPPORTED;
    if (n * 16 >= (1ll << 31) || n * (int64_t)ldx >= (1ll << 32)) return EVAD_ERR_ARG;   // 32-bit row offsets in the kernel
    if (epi_h->addend && ((((uintptr_t)epi_h->addend) & 15) != 0 || epi_h->ld_add % 4 != 0)) return EVAD_ERR_ARG;
    if (ldx < cin_pad || ldx % 4 != 0 || ldo % 4 != 0 || ncls < 1 || ncls > 256) return EVAD_ERR_ARG;
    if ((((uintptr_t)x | (uintptr_t)tapw | (uintptr_t)Whi | (uintptr_t)Wlo | (uintptr_t)out | (uintptr_t)nbr |
          (uintptr_t)nbr_cls) & 15) != 0)
        return EVAD_ERR_ARG;
    const bool has_skip = Shi && Slo && skip_epi_h && skip_out;
    if (has_skip && ((((uintptr_t)Shi | (uintptr_t)Slo | (uintptr_t)skip_out) & 15) != 0 || ld_skip % 4 != 0))
        return EVAD_ERR_ARG;
    if (n <= 0) return EVAD_OK;
    const evad_epilogue_t es = has_skip ? *skip_epi_h : *epi_h;
    cudaStream_t st = (cudaStream_t)stream;
    if (cin_pad == 20)
        return launch_l1_mma<20>(count, nbr, nbr_cls, perm, n, x, ldx, tapw, ncls, Whi, Wlo, has_skip ? Shi : nullptr,
                                 has_skip ? Slo : nullptr, *epi_h, out, ldo, es, skip_out, ld_skip, st);
    return launch_l1_mma<16>(count, nbr, nbr_cls, perm, n, x, ldx, tapw, ncls, Whi, Wlo, has_skip ? Shi : nullptr,
                             has_skip ? Slo : nullptr, *epi_h, out, ldo, es, skip_out, ld_skip, st);
}
