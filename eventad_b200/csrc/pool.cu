// eventad_b200/csrc/pool.cu -- voxel-grid pooling for sm_100a.
//
// Reference path replaced per call of Pooling.forward (paths relative to /root/reference/src/dagr/model):
//   layers/pooling.py:55-57   torch_cluster.grid_cluster over [x,y,t,batch] + consecutive_cluster (torch.unique)
//   layers/pooling.py:58-64   edge_index = cluster[edge_index]; drop self loops; unique(dim=-1)
//   layers/pooling.py:66-77   batch[perm], pool_pos (scatter mean), scatter_max / _avg_pool_x
//   layers/pooling.py:47-49,86 round_to_pixel
//   networks/net.py:15-17,200-228 sampling_skip: grid_sample of the CNN map at the nodes + concat, done right
//                              before every pooling call -- fused here so the [n, C+64] concat never exists.
// Design: nodes are grouped per voxel by one stable radix sort of (voxel id, node id); one CTA owns one voxel:
// thread 0 sums the positions serially in ascending node order (bit-identical to the CPU scatter_add the oracle
// models), the threads reduce the channels, and the coarse in-edges are collected in a shared-memory bitmap, which
// yields them de-duplicated and sorted with no global sort and no atomics on global memory.
#include "common.cuh"
#include "sample.cuh"
#include <stdlib.h>

#define PL_THREADS 128
#define PL_BITS 4096                                        // bitmap chunk: one 32-bit word per thread

struct PoolGeom {
    float sx, sy, st;        // voxel size (x, y, t); the batch dimension has size 1
    int nvx, nvy, nvt;       // voxels per dimension = ((end-start)/size).long() + 1  (computed by the host like torch_cluster)
    uint32_t key_bound;      // nvx*nvy*nvt*(B+1): valid keys are < key_bound
};

// torch_cluster grid_cluster, D = 4, start = 0: c_d = (int64)(p_d / size_d) in fp32
__device__ __forceinline__ uint32_t pool_voxel_key(const float* __restrict__ pos, int ld_pos, const int32_t* __restrict__ batch,
                                                   int64_t i, const PoolGeom& G, uint32_t* __restrict__ status) {
    const long long cx = (long long)__fdiv_rn(pos[i * ld_pos], G.sx);
    const long long cy = (long long)__fdiv_rn(pos[i * ld_pos + 1], G.sy);
    const long long ct = (long long)__fdiv_rn(pos[i * ld_pos + 2], G.st);
    const long long cb = (long long)__fdiv_rn((float)batch[i], 1.0f);
    long long key = cx + (long long)G.nvx * (cy + (long long)G.nvy * (ct + (long long)G.nvt * cb));
    if (cx < 0 || cy < 0 || ct < 0 || cb < 0 || key >= (long long)G.key_bound) {
        atomicOr(status, EVAD_STATUS_CLUSTER_OOB);
        key = min(max(key, 0ll), (long long)G.key_bound - 1);
    }
    return (uint32_t)key;
}

__global__ void __launch_bounds__(256) pool_keys_kernel(const float* __restrict__ pos, int ld_pos,
                                                        const int32_t* __restrict__ batch, int64_t n_cap,
                                                        const int32_t* __restrict__ n_dev, PoolGeom G,
                                                        uint32_t* __restrict__ keys, uint32_t* __restrict__ status) {
    int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n_cap) return;
    const int64_t n = n_dev ? min((int64_t)*n_dev, n_cap) : n_cap;
    if (i >= n) { keys[i] = G.key_bound; return; }          // padding sorts to the end
    keys[i] = pool_voxel_key(pos, ld_pos, batch, i, G, status);
}

__global__ void __launch_bounds__(256) pool_flag_kernel(const uint32_t* __restrict__ skey, int64_t n_cap,
                                                        uint32_t key_bound, uint32_t* __restrict__ flag) {
    int64_t p = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (p >= n_cap) return;
    const uint32_t k = skey[p];
    flag[p] = (p > 0 && k < key_bound && k != skey[p - 1]) ? 1u : 0u;
}

// dense ids (consecutive_cluster), segment starts, cluster[] in node order
__global__ void __launch_bounds__(256) pool_relabel_kernel(const uint32_t* __restrict__ skey,
                                                           const uint32_t* __restrict__ sidx,
                                                           const uint32_t* __restrict__ flag,
                                                           const uint32_t* __restrict__ excl, int64_t n_cap,
                                                           uint32_t key_bound, int32_t* __restrict__ cluster,
                                                           int32_t* __restrict__ seg_start,
                                                           int32_t* __restrict__ n_out) {
    int64_t p = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (p == 0 && (n_cap == 0 || skey[0] >= key_bound)) { *n_out = 0; seg_start[0] = 0; }
    if (p >= n_cap) return;
    if (skey[p] >= key_bound) return;
    const uint32_t d = excl[p] + flag[p];
    cluster[sidx[p]] = (int32_t)d;
    if (p == 0 || flag[p]) seg_start[d] = (int32_t)p;
    if (p == n_cap - 1 || skey[p + 1] >= key_bound) { seg_start[d + 1] = (int32_t)(p + 1); *n_out = (int32_t)(d + 1); }
}

// ------------------------------------------------------------------------------------------------ small inputs
// Poolings 2-4 see a few thousand nodes: keys, stable sort, flag, scan and relabel (ten ~3 us launches above) run as ONE
// CTA here, by counting in shared memory: voxel histogram -> block scan (segment starts, dense ids) -> placement ->
// each segment put into ascending node order (segments are a handful of nodes: the children of one voxel).  Outputs are
// identical to the general path: skey / sidx (sorted pairs, padding at the end), cluster, seg_start, n_out.
#define PFS_THREADS 1024
#define PFS_NMAX 4608
#define PFS_KMAX 8192

__global__ void __launch_bounds__(PFS_THREADS) pool_front_small_kernel(
    const float* __restrict__ pos, int ld_pos, const int32_t* __restrict__ batch, int n_cap, const int32_t* __restrict__ n_dev,
    PoolGeom G, uint32_t* __restrict__ skey, uint32_t* __restrict__ sidx, int32_t* __restrict__ cluster,
    int32_t* __restrict__ seg_start, int32_t* __restrict__ n_out, uint32_t* __restrict__ status) {
    extern __shared__ uint32_t pfs_smem[];
    uint32_t* key = pfs_smem;                                // [PFS_NMAX]
    uint32_t* sid = key + PFS_NMAX;                          // [PFS_NMAX] node ids in sorted order
    uint32_t* cur = sid + PFS_NMAX;                          // [PFS_KMAX] counts, then placement cursors
    uint32_t* beg = cur + PFS_KMAX;                          // [PFS_KMAX] segment starts
    uint32_t* den = beg + PFS_KMAX;                          // [PFS_KMAX] dense ids
    __shared__ uint32_t s_wc[32], s_wn[32];
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int n = n_dev ? min(*n_dev, n_cap) : n_cap;
    const int kb = (int)G.key_bound;
    for (int b = tid; b < kb; b += PFS_THREADS) cur[b] = 0;
    __syncthreads();
    for (int i = tid; i < n; i += PFS_THREADS) {
        const uint32_t k = pool_voxel_key(pos, ld_pos, batch, i, G, status);
        key[i] = k;
        atomicAdd(&cur[k], 1u);
    }
    __syncthreads();
    // block scan over the voxel bins: every thread owns a contiguous run of bins
    const int per = (kb + PFS_THREADS - 1) / PFS_THREADS;
    const int b0 = min(tid * per, kb), b1 = min(b0 + per, kb);
    uint32_t csum = 0, nsum = 0;
    for (int b = b0; b < b1; b++) { const uint32_t c = cur[b]; csum += c; nsum += c > 0; }
    uint32_t ic = csum, in = nsum;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        const uint32_t uc = __shfl_up_sync(0xffffffffu, ic, o), un = __shfl_up_sync(0xffffffffu, in, o);
        if (lane >= o) { ic += uc; in += un; }
    }
    if (lane == 31) { s_wc[warp] = ic; s_wn[warp] = in; }
    __syncthreads();
    if (warp == 0) {
        uint32_t wc = s_wc[lane], wn = s_wn[lane];
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            const uint32_t uc = __shfl_up_sync(0xffffffffu, wc, o), un = __shfl_up_sync(0xffffffffu, wn, o);
            if (lane >= o) { wc += uc; wn += un; }
        }
        s_wc[lane] = wc; s_wn[lane] = wn;                    // inclusive over warps
    }
    __syncthreads();
    uint32_t ec = ic - csum + (warp ? s_wc[warp - 1] : 0u), en = in - nsum + (warp ? s_wn[warp - 1] : 0u);
    const uint32_t total_n = s_wn[31];
    for (int b = b0; b < b1; b++) {
        const uint32_t c = cur[b];
        beg[b] = ec; den[b] = en; cur[b] = ec;
        if (c) seg_start[en] = (int32_t)ec;
        ec += c; en += c > 0;
    }
    if (tid == 0) { *n_out = (int32_t)total_n; seg_start[total_n] = (int32_t)n; }
    __syncthreads();
    for (int i = tid; i < n; i += PFS_THREADS) sid[atomicAdd(&cur[key[i]], 1u)] = (uint32_t)i;
    __syncthreads();
    for (int b = tid; b < kb; b += PFS_THREADS) {            // ascending node ids inside every segment (= the stable sort)
        const int s0 = (int)beg[b], s1 = (int)cur[b];
        for (int a = s0 + 1; a < s1; a++) {
            const uint32_t v = sid[a];
            int q = a - 1;
            while (q >= s0 && sid[q] > v) { sid[q + 1] = sid[q]; q--; }
            sid[q + 1] = v;
        }
    }
    __syncthreads();
    for (int p = tid; p < n_cap; p += PFS_THREADS) {
        if (p < n) {
            const uint32_t i = sid[p], k = key[i];
            skey[p] = k; sidx[p] = i;
            cluster[i] = (int32_t)den[k];
        } else {
            skey[p] = G.key_bound; sidx[p] = (uint32_t)p;    // padding keeps its place, as in the stable sort
        }
    }
}

// c10::div_floor_floating (torch.div(..., rounding_mode='floor') on floats), used by round_to_pixel (pooling.py:47-49)
__device__ __forceinline__ float div_floor_f(float a, float b) {
    if (b == 0.f) return __fdiv_rn(a, b);
    const float mod = fmodf(a, b);
    float div = __fdiv_rn(__fsub_rn(a, mod), b);
    if (mod != 0.f && ((b < 0.f) != (mod < 0.f))) div = __fsub_rn(div, 1.f);
    float fl;
    if (div != 0.f) {
        fl = floorf(div);
        if (__fsub_rn(div, fl) > 0.5f) fl = __fadd_rn(fl, 1.f);
    } else {
        fl = copysignf(0.f, __fdiv_rn(a, b));
    }
    return fl;
}

// ------------------------------------------------------------------------------------------------ per-voxel reduce
// One CTA per output cluster.  x_out row = [reduced x (cx) | reduced sampled map (S.C) | pos_x, pos_y (if append_pos)].
// Members are staged in chunks of PL_CHUNK: one thread per member resolves the node id and the (up to 8) trilinear
// taps of the CNN map once; the channel threads then only do coalesced loads + FMAs.
#define PL_CHUNK 32

__global__ void __launch_bounds__(PL_THREADS) pool_reduce_kernel(
    const uint32_t* __restrict__ sidx, const int32_t* __restrict__ seg_start, const int32_t* __restrict__ n_out,
    const float* __restrict__ pos, int ld_pos, const int32_t* __restrict__ batch, const float* __restrict__ x, int ldx,
    int cx, PoolSample S, int aggr_mean, float inv_w, float inv_h, int append_pos, float* __restrict__ x_out, int ldo,
    float* __restrict__ pos_out, int32_t* __restrict__ batch_out) {
    const int c = blockIdx.x;
    if (c >= *n_out) return;
    const int s0 = seg_start[c], s1 = seg_start[c + 1];
    __shared__ float s_pos[2];
    __shared__ int s_node[PL_CHUNK];
    __shared__ int s_tap[PL_CHUNK][8];
    __shared__ float s_wgt[PL_CHUNK][8];
    const int ctot = cx + (S.feat ? S.C : 0);
    const float cntf = fmaxf((float)(s1 - s0), 1.f);
    float sx = 0.f, sy = 0.f, st = 0.f;                      // thread 0 only: serial position sums
    float acc[2] = {0.f, 0.f};                               // up to 2 channels per thread (ctot <= 2*PL_THREADS)
    for (int p0 = s0; p0 < s1; p0 += PL_CHUNK) {
        const int cnt = min(PL_CHUNK, s1 - p0);
        __syncthreads();
        if (threadIdx.x < cnt) {
            const int m = threadIdx.x;
            const int64_t i = sidx[p0 + m];
            s_node[m] = (int)i;
            if (S.feat) sample_taps(S, pos[i * ld_pos], pos[i * ld_pos + 1], (float)batch[i], s_tap[m], s_wgt[m]);
        }
        __syncthreads();
        if (threadIdx.x == 0) {
            // pool_pos = scatter mean: fp32 sums in ascending node order, then / count.clamp(min=1)
            for (int m = 0; m < cnt; m++) {
                const int64_t i = s_node[m];
                sx = __fadd_rn(sx, pos[i * ld_pos]);
                sy = __fadd_rn(sy, pos[i * ld_pos + 1]);
                st = __fadd_rn(st, pos[i * ld_pos + 2]);
            }
        }
#pragma unroll
        for (int u = 0; u < 2; u++) {
            const int ch = threadIdx.x + u * PL_THREADS;
            if (ch >= ctot) continue;
            float a = acc[u];
            for (int m = 0; m < cnt; m++) {
                float v;
                if (ch < cx) {
                    v = x[(int64_t)s_node[m] * ldx + ch];
                } else {
                    v = 0.f;
#pragma unroll
                    for (int t = 0; t < 8; t++) {
                        const int idx = s_tap[m][t];
                        if (idx >= 0) v = fmaf(S.feat[(int64_t)idx * S.C + (ch - cx)], s_wgt[m][t], v);
                    }
                }
                if (aggr_mean) a = __fadd_rn(a, v);
                else a = (p0 == s0 && m == 0) ? v : fmaxf(a, v);
            }
            acc[u] = a;
        }
    }
    if (threadIdx.x == 0) {
        float mx = __fdiv_rn(sx, cntf), my = __fdiv_rn(sy, cntf);
        const float mt = __fdiv_rn(st, cntf);
        // round_to_pixel: floor_divide(p + 1e-5, wh_inv) * wh_inv
        mx = __fmul_rn(div_floor_f(__fadd_rn(mx, 1e-5f), inv_w), inv_w);
        my = __fmul_rn(div_floor_f(__fadd_rn(my, 1e-5f), inv_h), inv_h);
        pos_out[(int64_t)c * 3] = mx;
        pos_out[(int64_t)c * 3 + 1] = my;
        pos_out[(int64_t)c * 3 + 2] = mt;
        batch_out[c] = batch[sidx[s1 - 1]];                 // perm: last writer wins = largest node id (pooling.py:15,66)
        s_pos[0] = mx;
        s_pos[1] = my;
    }
#pragma unroll
    for (int u = 0; u < 2; u++) {
        const int ch = threadIdx.x + u * PL_THREADS;
        if (ch < ctot) x_out[(int64_t)c * ldo + ch] = aggr_mean ? __fdiv_rn(acc[u], cntf) : acc[u];
    }
    if (append_pos) {
        __syncthreads();
        if (threadIdx.x < 2) x_out[(int64_t)c * ldo + ctot + threadIdx.x] = s_pos[threadIdx.x];
    }
    for (int ch = ctot + (append_pos ? 2 : 0) + threadIdx.x; ch < ldo; ch += PL_THREADS) x_out[(int64_t)c * ldo + ch] = 0.f;
}

// Warp-per-voxel variant (cx % 4 == 0, S.C % 4 == 0, 16-byte aligned rows).
//  * members of a chunk (<= 32) are resolved one per lane: node id, position, trilinear taps;
//  * node features: lane = member, one 128-bit row load per channel quad (all in flight together), then a butterfly
//    max / sum over the lanes -- no per-member memory latency;
//  * sampled CNN features: all members of a voxel sample the same few map cells (a voxel spans ~1.5 cells of the map
//    at its stride), so the bounding window of the chunk (<= PRW_WIN cells) is staged once in warp-private shared
//    memory and the member loop runs out of shared memory; lanes own channel quads, the two half-warps take alternate
//    members.  Per channel the taps are accumulated dz,dy,dx with fmaf exactly like sample_feat, and max pooling is
//    order independent, so max-pooled rows are bit-identical to pool_reduce_kernel; mean rows agree to rounding;
//  * positions: serial fp32 sums in ascending node order (bit-exact with the oracle's CPU scatter_add).
#define PRW_WARPS 8
#define PRW_WIN 16                    // window cells staged per warp: a voxel spans ~1.5 map cells per axis, single plane

__device__ __forceinline__ float4 prw_comb(float4 a, float4 b, int mean) {
    if (mean) return make_float4(__fadd_rn(a.x, b.x), __fadd_rn(a.y, b.y), __fadd_rn(a.z, b.z), __fadd_rn(a.w, b.w));
    return make_float4(fmaxf(a.x, b.x), fmaxf(a.y, b.y), fmaxf(a.z, b.z), fmaxf(a.w, b.w));
}

__device__ __forceinline__ float4 prw_shfl_xor(float4 v, int d) {
    return make_float4(__shfl_xor_sync(0xffffffffu, v.x, d), __shfl_xor_sync(0xffffffffu, v.y, d),
                       __shfl_xor_sync(0xffffffffu, v.z, d), __shfl_xor_sync(0xffffffffu, v.w, d));
}

// member loop of one lane group: channel quad q of members grp, grp + groups, ... of the chunk; NT taps per member
// (dz, dy, dx order, fmaf chain as in sample_feat), from the staged window (WIN) or straight from the map
template <bool WIN, int NT>
__device__ __forceinline__ float4 prw_members(float4 a, const int* __restrict__ s_tap, const float* __restrict__ s_wgt,
                                              const float4* __restrict__ s_win, const float4* __restrict__ f4, int q, int grp,
                                              int cnt, int groups, int aggr_mean) {
    for (int m = grp; m < cnt; m += groups) {
        int tp[8];
        float wg[8];
        {
            const int4 t0 = *reinterpret_cast<const int4*>(s_tap + m * 8);
            const float4 w0 = *reinterpret_cast<const float4*>(s_wgt + m * 8);
            tp[0] = t0.x; tp[1] = t0.y; tp[2] = t0.z; tp[3] = t0.w;
            wg[0] = w0.x; wg[1] = w0.y; wg[2] = w0.z; wg[3] = w0.w;
        }
        if (NT == 8) {
            const int4 t1 = *reinterpret_cast<const int4*>(s_tap + m * 8 + 4);
            const float4 w1 = *reinterpret_cast<const float4*>(s_wgt + m * 8 + 4);
            tp[4] = t1.x; tp[5] = t1.y; tp[6] = t1.z; tp[7] = t1.w;
            wg[4] = w1.x; wg[5] = w1.y; wg[6] = w1.z; wg[7] = w1.w;
        }
        float4 fv[NT];
#pragma unroll
        for (int k = 0; k < NT; k++) fv[k] = WIN ? s_win[tp[k] + q] : __ldg(f4 + tp[k] + q);
        float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
#pragma unroll
        for (int k = 0; k < NT; k++) {
            v.x = fmaf(fv[k].x, wg[k], v.x); v.y = fmaf(fv[k].y, wg[k], v.y);
            v.z = fmaf(fv[k].z, wg[k], v.z); v.w = fmaf(fv[k].w, wg[k], v.w);
        }
        a = prw_comb(a, v, aggr_mean);
    }
    return a;
}

__global__ void __launch_bounds__(PRW_WARPS * 32, 3) pool_reduce_warp_kernel(
    const uint32_t* __restrict__ sidx, const int32_t* __restrict__ seg_start, const int32_t* __restrict__ n_out,
    const float* __restrict__ pos, int ld_pos, const int32_t* __restrict__ batch, const float* __restrict__ x, int ldx,
    int cx, PoolSample S, int aggr_mean, float inv_w, float inv_h, int append_pos, float* __restrict__ x_out, int ldo,
    float* __restrict__ pos_out, int32_t* __restrict__ batch_out) {
    extern __shared__ __align__(16) float prw_smem[];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int c = blockIdx.x * PRW_WARPS + warp;
    if (c >= *n_out) return;
    const int qx = cx >> 2, qf = S.feat ? (S.C >> 2) : 0;
    int qxp = 1;
    while (qxp < qx) qxp <<= 1;                              // lanes per member in the node-feature pass (qx <= 32)
    float* wbase = prw_smem + (size_t)warp * (PRW_WIN * 4 * qf + 640);
    float4* s_win = reinterpret_cast<float4*>(wbase);                       // [PRW_WIN][qf] float4
    int* s_tap = reinterpret_cast<int*>(wbase + PRW_WIN * 4 * qf);          // [32][8]
    float* s_wgt = wbase + PRW_WIN * 4 * qf + 256;                          // [32][8]
    int* s_cell = reinterpret_cast<int*>(wbase + PRW_WIN * 4 * qf + 512);   // [32] flat map index of window cells
    float* s_pos = wbase + PRW_WIN * 4 * qf + 544;                          // [32][3] member positions of the chunk
    const int s0 = seg_start[c], s1 = seg_start[c + 1];
    const float cntf = fmaxf((float)(s1 - s0), 1.f);
    const float4* f4 = reinterpret_cast<const float4*>(S.feat);
    const float ident = aggr_mean ? 0.f : -INFINITY;
    const int groups = (qf > 0 && qf <= 16) ? 2 : 1, lg = 32 / groups;      // sampled part: lane groups over members
    const int grp = lane / lg, ql = lane - grp * lg;
    float psum = 0.f;                                        // lane d < 3: running sum of position component d
    float4 accx[2], accf[2];
#pragma unroll
    for (int u = 0; u < 2; u++) accx[u] = accf[u] = make_float4(ident, ident, ident, ident);
    for (int p0 = s0; p0 < s1; p0 += 32) {
        const int cnt = min(32, s1 - p0);
        float px = 0.f, py = 0.f, pt = 0.f;
        int node = 0, bx = 0x7fffffff, by = 0x7fffffff, bz = 0x7fffffff;
        int tap[8];
        float wgt[8];
        __syncwarp();
        if (lane < cnt) {
            node = (int)sidx[p0 + lane];
            px = pos[(int64_t)node * ld_pos]; py = pos[(int64_t)node * ld_pos + 1]; pt = pos[(int64_t)node * ld_pos + 2];
            if (qf) sample_taps_base(S, px, py, (float)batch[node], tap, wgt, &bx, &by, &bz);
        }
        // ---- node features: lane = (member slot, channel quad): qxp = qx rounded up to a power of two lanes per member,
        //      32 / qxp members per pass, one 128-bit row load each, running max / sum in registers; the member slots
        //      are folded by a butterfly once per voxel (below).  More than 32 quads: lane = member, butterfly per quad.
        if (qx <= 32) {
            const int ms = lane / qxp, quad = lane - ms * qxp, mpp = 32 / qxp;
            for (int m0 = 0; m0 < cnt; m0 += mpp) {
                const int m = m0 + ms;
                const int nd = __shfl_sync(0xffffffffu, node, m & 31);
                if (m < cnt && quad < qx)
                    accx[0] = prw_comb(accx[0], __ldg(reinterpret_cast<const float4*>(x + (int64_t)nd * ldx) + quad), aggr_mean);
            }
        } else {
            for (int q0 = 0; q0 < qx; q0 += 4) {
                float4 v[4];
#pragma unroll
                for (int u = 0; u < 4; u++) {
                    v[u] = make_float4(ident, ident, ident, ident);
                    if (q0 + u < qx && lane < cnt) v[u] = __ldg(reinterpret_cast<const float4*>(x + (int64_t)node * ldx) + q0 + u);
                }
#pragma unroll
                for (int u = 0; u < 4; u++) {
                    if (q0 + u >= qx) break;
#pragma unroll
                    for (int d = 16; d >= 1; d >>= 1) v[u] = prw_comb(v[u], prw_shfl_xor(v[u], d), aggr_mean);
                    const int q = q0 + u;
                    if (lane == (q & 31)) {
                        if (q < 32) accx[0] = prw_comb(accx[0], v[u], aggr_mean);
                        else accx[1] = prw_comb(accx[1], v[u], aggr_mean);
                    }
                }
            }
        }
        // ---- positions: serial fp32 sums in ascending node order; lane d < 3 sums component d from shared memory
        //      (one load + one add per member instead of three shuffles + three adds)
        if (lane < cnt) { s_pos[3 * lane] = px; s_pos[3 * lane + 1] = py; s_pos[3 * lane + 2] = pt; }
        __syncwarp();
        if (lane < 3)
            for (int m = 0; m < cnt; m++) psum = __fadd_rn(psum, s_pos[3 * m + lane]);
        __syncwarp();
        if (!qf) continue;
        // ---- sampled CNN features: window of the chunk in shared memory
        const int xmin = __reduce_min_sync(0xffffffffu, bx), ymin = __reduce_min_sync(0xffffffffu, by),
                  zmin = __reduce_min_sync(0xffffffffu, bz);
        const int xmax = __reduce_max_sync(0xffffffffu, lane < cnt ? bx : (int)0x80000000),
                  ymax = __reduce_max_sync(0xffffffffu, lane < cnt ? by : (int)0x80000000),
                  zmax = __reduce_max_sync(0xffffffffu, lane < cnt ? bz : (int)0x80000000);
        // taps 4..7 are the dz = 1 plane: with the batch index as depth coordinate their weights are exactly 0 for every
        // member (fmaf(x, 0, v) == v), so the common case stages one plane less and runs a 4-tap loop -- bit-identical
        bool hi0 = true;
        if (lane < cnt) hi0 = wgt[4] == 0.f && wgt[5] == 0.f && wgt[6] == 0.f && wgt[7] == 0.f;
        const bool four = __all_sync(0xffffffffu, hi0);
        const int wx = xmax - xmin + 2, wy = ymax - ymin + 2, wz = zmax - zmin + (four ? 1 : 2);
        const bool use_win = (long long)wx * wy * wz <= PRW_WIN;
        if (use_win) {
            // lane c resolves window cell c to its flat map index once (-1 outside the map); the copy loop then walks
            // (cell, quad) incrementally -- no per-element divisions
            const int ncell = wx * wy * wz;
            if (lane < ncell) {
                const int cz = lane / (wx * wy), r = lane - cz * wx * wy, cy = r / wx, cxx = r - cy * wx;
                const int gx = xmin + cxx, gy = ymin + cy, gz = zmin + cz;
                s_cell[lane] = (gx >= 0 && gx < S.w && gy >= 0 && gy < S.h && gz >= 0 && gz < S.Bimg)
                                   ? (gz * S.h + gy) * S.w + gx : -1;
            }
            __syncwarp();
            int cell = lane / qf, q = lane - cell * qf;
            const int dcell = 32 / qf, dq = 32 - dcell * qf;
            for (int idx = lane; idx < ncell * qf; idx += 32) {
                const int gcell = s_cell[cell];
                float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
                if (gcell >= 0) v = __ldg(f4 + (int64_t)gcell * qf + q);
                s_win[idx] = v;
                q += dq; cell += dcell;
                if (q >= qf) { q -= qf; cell++; }
            }
        }
        if (lane < cnt) {
            // taps outside the map get weight 0 on a valid cell: fmaf(x, 0, v) == v, so the member loop is branch free
#pragma unroll
            for (int k = 0; k < 8; k++) {
                int tl = tap[k];
                float wk = wgt[k];
                if (tl < 0) { tl = 0; wk = 0.f; }
                else if (use_win)
                    tl = ((bz - zmin + (k >> 2)) * wy + (by - ymin + ((k >> 1) & 1))) * wx + (bx - xmin + (k & 1));
                s_tap[lane * 8 + k] = tl * qf;
                s_wgt[lane * 8 + k] = wk;
            }
        }
        __syncwarp();
#pragma unroll
        for (int u = 0; u < 2; u++) {
            const int q = ql + u * lg;
            if (q >= qf) continue;
            if (four) {
                if (use_win) accf[u] = prw_members<true, 4>(accf[u], s_tap, s_wgt, s_win, f4, q, grp, cnt, groups, aggr_mean);
                else accf[u] = prw_members<false, 4>(accf[u], s_tap, s_wgt, s_win, f4, q, grp, cnt, groups, aggr_mean);
            } else if (use_win) {
                accf[u] = prw_members<true, 8>(accf[u], s_tap, s_wgt, s_win, f4, q, grp, cnt, groups, aggr_mean);
            } else {
                accf[u] = prw_members<false, 8>(accf[u], s_tap, s_wgt, s_win, f4, q, grp, cnt, groups, aggr_mean);
            }
        }
    }
    if (qx <= 32)
        for (int d = qxp; d < 32; d <<= 1) accx[0] = prw_comb(accx[0], prw_shfl_xor(accx[0], d), aggr_mean);
    if (groups == 2) accf[0] = prw_comb(accf[0], prw_shfl_xor(accf[0], 16), aggr_mean);
    const float sx = __shfl_sync(0xffffffffu, psum, 0), sy = __shfl_sync(0xffffffffu, psum, 1),
                st = __shfl_sync(0xffffffffu, psum, 2);
    float mx = __fdiv_rn(sx, cntf), my = __fdiv_rn(sy, cntf);
    const float mt = __fdiv_rn(st, cntf);
    mx = __fmul_rn(div_floor_f(__fadd_rn(mx, 1e-5f), inv_w), inv_w);
    my = __fmul_rn(div_floor_f(__fadd_rn(my, 1e-5f), inv_h), inv_h);
    if (lane == 0) {
        pos_out[(int64_t)c * 3] = mx;
        pos_out[(int64_t)c * 3 + 1] = my;
        pos_out[(int64_t)c * 3 + 2] = mt;
        batch_out[c] = batch[sidx[s1 - 1]];
    }
    float* o = x_out + (int64_t)c * ldo;
#pragma unroll
    for (int u = 0; u < 2; u++) {
        const int q = lane + 32 * u;
        if (q < qx) {
            float4 a = accx[u];
            if (aggr_mean) { a.x = __fdiv_rn(a.x, cntf); a.y = __fdiv_rn(a.y, cntf); a.z = __fdiv_rn(a.z, cntf); a.w = __fdiv_rn(a.w, cntf); }
            *reinterpret_cast<float4*>(o + 4 * q) = a;
        }
        const int qq = ql + u * lg;
        if (grp == 0 && qq < qf) {
            float4 a = accf[u];
            if (aggr_mean) { a.x = __fdiv_rn(a.x, cntf); a.y = __fdiv_rn(a.y, cntf); a.z = __fdiv_rn(a.z, cntf); a.w = __fdiv_rn(a.w, cntf); }
            *reinterpret_cast<float4*>(o + 4 * (qx + qq)) = a;
        }
    }
    const int nq = qx + qf;
    if (append_pos && lane < 2) o[4 * nq + lane] = lane ? my : mx;
    for (int ch = 4 * nq + (append_pos ? 2 : 0) + lane; ch < ldo; ch += 32) o[ch] = 0.f;     // row padding
}

// ------------------------------------------------------------------------------------------------ coarse edges
// One CTA per destination cluster: mark cluster[src] of every in-edge of every member (self loops dropped) in a
// shared-memory bitmap; FILL=false counts the distinct sources, FILL=true writes them (ascending).
// Fast path (the common case): the sources of cluster c lie in the 4096-id window centred on c (dense ids follow the
// voxel order, neighbours are a few voxel rows away), so ONE pass over the fine edges yields the bitmap; rows of at most
// PL_ROWCAP sources are parked in `tmp` and the FILL launch only copies them to their CSR position.  Anything else
// (sources outside the window, longer rows) takes the general path: id range first, then 4096-id chunks, in both launches.
#define PL_ROWCAP 64

// fine-graph rows: CSR (in_stride == 0, rowptr[n+1]) or fixed-stride rows (in_stride > 0, rowptr[i] = row length)
__device__ __forceinline__ int pl_row_begin(const int32_t* __restrict__ rowptr, int in_stride, int i) {
    return in_stride > 0 ? i * in_stride : rowptr[i];
}
__device__ __forceinline__ int pl_row_end(const int32_t* __restrict__ rowptr, int in_stride, int i) {
    return in_stride > 0 ? i * in_stride + rowptr[i] : rowptr[i + 1];
}

// exclusive prefix of per-thread popcounts over the CTA (PL_THREADS = 128); returns the CTA total
__device__ __forceinline__ uint32_t pool_popc_scan(uint32_t pc, uint32_t* s_warp, uint32_t* excl) {
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    uint32_t inc = pc;
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) {
        uint32_t t = __shfl_up_sync(0xffffffffu, inc, d);
        if (lane >= d) inc += t;
    }
    if (lane == 31) s_warp[warp] = inc;
    __syncthreads();
    uint32_t woff = 0, total = 0;
#pragma unroll
    for (int w = 0; w < PL_THREADS / 32; w++) {
        if (w < warp) woff += s_warp[w];
        total += s_warp[w];
    }
    *excl = woff + inc - pc;
    __syncthreads();
    return total;
}

template <bool FILL>
__global__ void __launch_bounds__(PL_THREADS) pool_edges_kernel(
    const uint32_t* __restrict__ sidx, const int32_t* __restrict__ seg_start, const int32_t* __restrict__ n_out,
    const int32_t* __restrict__ rowptr_in, int in_stride, int keep_self, const int32_t* __restrict__ src_in,
    const int32_t* __restrict__ cluster, uint32_t* __restrict__ deg, const uint32_t* __restrict__ rowptr_out, int32_t* __restrict__ src_out, int64_t e_cap,
    int32_t* __restrict__ tmp, uint8_t* __restrict__ parked, uint32_t* __restrict__ status,
    int32_t* __restrict__ rowptr_final, int32_t* __restrict__ num_edges) {
    const int c = blockIdx.x;
    const int nc = *n_out;
    if (FILL && threadIdx.x == 0 && c <= nc) {               // the caller's int32 row pointers and edge count
        rowptr_final[c] = (int32_t)rowptr_out[c];
        if (c == nc) *num_edges = (int32_t)rowptr_out[c];
    }
    if (c >= nc) {
        if (!FILL && c == nc) deg[c] = 0;
        return;
    }
    __shared__ uint32_t bits[PL_BITS / 32];
    __shared__ int s_lo, s_hi, s_oow;
    __shared__ uint32_t s_warp[PL_THREADS / 32 + 1];
    if (FILL && parked[c]) {                                 // row parked by the counting launch: copy it into place
        const uint32_t base = rowptr_out[c], cnt = rowptr_out[c + 1] - base;
        for (uint32_t k = threadIdx.x; k < cnt; k += PL_THREADS) {
            if ((int64_t)(base + k) < e_cap) src_out[base + k] = tmp[(int64_t)c * PL_ROWCAP + k];
            else atomicOr(status, EVAD_STATUS_EDGE_OVERFLOW);
        }
        return;
    }
    const int s0 = seg_start[c], s1 = seg_start[c + 1];
    if (!FILL) {
        bits[threadIdx.x] = 0u;
        if (threadIdx.x == 0) s_oow = 0;
        __syncthreads();
        const int w0 = max(0, c - PL_BITS / 2);
        bool oow = false;
        for (int p = s0 + threadIdx.x; p < s1; p += PL_THREADS) {
            const int i = (int)sidx[p];
            if (in_stride == 16) {
                // fixed-stride rows of 16 slots: four source ids per 128-bit load, their four cluster gathers in flight
                const int deg = rowptr_in[i];
                const int4* row = reinterpret_cast<const int4*>(src_in) + (int64_t)i * 4;
                for (int g = 0; 4 * g < deg; g++) {
                    const int4 s4 = __ldg(row + g);
                    const int sj[4] = {s4.x, s4.y, s4.z, s4.w};
                    int scs[4];
#pragma unroll
                    for (int k = 0; k < 4; k++) scs[k] = 4 * g + k < deg ? cluster[sj[k]] : c;
#pragma unroll
                    for (int k = 0; k < 4; k++) {
                        if (4 * g + k >= deg) break;
                        const int sc = scs[k], r = sc - w0;
                        if (sc == c && !keep_self) continue;
                        if (r >= 0 && r < PL_BITS) atomicOr(&bits[r >> 5], 1u << (r & 31));
                        else oow = true;
                    }
                }
                continue;
            }
            for (int e = pl_row_begin(rowptr_in, in_stride, i), e1 = pl_row_end(rowptr_in, in_stride, i); e < e1; e++) {
                const int sc = cluster[src_in[e]];
                const int r = sc - w0;
                if (sc == c && !keep_self) continue;             // Pooling(self_loop=False) drops them (pooling.py:60-63)
                if (r >= 0 && r < PL_BITS) atomicOr(&bits[r >> 5], 1u << (r & 31));
                else oow = true;
            }
        }
        if (oow) s_oow = 1;
        __syncthreads();
        if (!s_oow) {
            uint32_t word = bits[threadIdx.x];
            uint32_t off;
            const uint32_t total = pool_popc_scan(__popc(word), s_warp, &off);
            const bool park = total <= PL_ROWCAP;
            if (park) {
                while (word) {
                    const int b = __ffs(word) - 1;
                    word &= word - 1;
                    tmp[(int64_t)c * PL_ROWCAP + off++] = w0 + threadIdx.x * 32 + b;
                }
            }
            if (threadIdx.x == 0) { deg[c] = total; parked[c] = park ? 1 : 0; }
            return;
        }
        if (threadIdx.x == 0) parked[c] = 0;
        __syncthreads();
    }
    // ---- general path
    if (threadIdx.x == 0) { s_lo = 0x7fffffff; s_hi = -1; }
    __syncthreads();
    int lo = 0x7fffffff, hi = -1;
    for (int p = s0 + threadIdx.x; p < s1; p += PL_THREADS) {
        const int i = (int)sidx[p];
        for (int e = pl_row_begin(rowptr_in, in_stride, i), e1 = pl_row_end(rowptr_in, in_stride, i); e < e1; e++) {
            const int sc = cluster[src_in[e]];
            if (sc != c || keep_self) { lo = min(lo, sc); hi = max(hi, sc); }
        }
    }
    atomicMin(&s_lo, lo);
    atomicMax(&s_hi, hi);
    __syncthreads();
    lo = s_lo; hi = s_hi;
    uint32_t total = 0;
    const uint32_t base_out = FILL ? rowptr_out[c] : 0u;
    for (int w0 = lo; w0 <= hi; w0 += PL_BITS) {
        bits[threadIdx.x] = 0u;
        __syncthreads();
        for (int p = s0 + threadIdx.x; p < s1; p += PL_THREADS) {
            const int i = (int)sidx[p];
            for (int e = pl_row_begin(rowptr_in, in_stride, i), e1 = pl_row_end(rowptr_in, in_stride, i); e < e1; e++) {
                const int sc = cluster[src_in[e]];
                const int r = sc - w0;
                if ((sc != c || keep_self) && r >= 0 && r < PL_BITS) atomicOr(&bits[r >> 5], 1u << (r & 31));
            }
        }
        __syncthreads();
        uint32_t word = bits[threadIdx.x];
        uint32_t off;
        const uint32_t chunk_total = pool_popc_scan(__popc(word), s_warp, &off);
        if (FILL) {
            uint32_t o = base_out + total + off;
            while (word) {
                const int b = __ffs(word) - 1;
                word &= word - 1;
                if ((int64_t)o < e_cap) src_out[o] = w0 + threadIdx.x * 32 + b;
                else atomicOr(status, EVAD_STATUS_EDGE_OVERFLOW);
                o++;
            }
        }
        total += chunk_total;
    }
    if (!FILL && threadIdx.x == 0) deg[c] = total;
}

// The counting launch (pool_edges_kernel<false>) with one WARP per destination cluster: a cluster has a few dozen members, so
// a 128-thread CTA leaves most threads idle in the member loop and spends the rest of its time in block barriers; here
// lanes take the members, the 4096-bit window is private to the warp (four words per lane, scanned with shuffles) and
// nothing waits on another warp.  Writes deg / parked / tmp exactly like the CTA version.
#define PEW_WARPS 4
__global__ void __launch_bounds__(PEW_WARPS * 32) pool_edges_count_warp_kernel(
    const uint32_t* __restrict__ sidx, const int32_t* __restrict__ seg_start, const int32_t* __restrict__ n_out,
    const int32_t* __restrict__ rowptr_in, int in_stride, int keep_self, const int32_t* __restrict__ src_in,
    const int32_t* __restrict__ cluster, uint32_t* __restrict__ deg, int32_t* __restrict__ tmp, uint8_t* __restrict__ parked) {
    __shared__ uint32_t s_bits[PEW_WARPS][PL_BITS / 32];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int c = blockIdx.x * PEW_WARPS + warp;
    const int nc = *n_out;
    if (c > nc) return;
    if (c == nc) {
        if (lane == 0) deg[c] = 0;
        return;
    }
    uint32_t* bits = s_bits[warp];
    const int s0 = seg_start[c], s1 = seg_start[c + 1];
#pragma unroll
    for (int k = 0; k < PL_BITS / 1024; k++) bits[lane + 32 * k] = 0u;
    __syncwarp();
    const int w0 = max(0, c - PL_BITS / 2);
    bool oow = false;
    for (int p = s0 + lane; p < s1; p += 32) {
        const int i = (int)sidx[p];
        if (in_stride == 16) {
            // fixed-stride rows of 16 slots: four source ids per 128-bit load, their four cluster gathers in flight
            const int dg = rowptr_in[i];
            const int4* row = reinterpret_cast<const int4*>(src_in) + (int64_t)i * 4;
            for (int g = 0; 4 * g < dg; g++) {
                const int4 s4 = __ldg(row + g);
                const int sj[4] = {s4.x, s4.y, s4.z, s4.w};
                int scs[4];
#pragma unroll
                for (int k = 0; k < 4; k++) scs[k] = 4 * g + k < dg ? cluster[sj[k]] : c;
#pragma unroll
                for (int k = 0; k < 4; k++) {
                    if (4 * g + k >= dg) break;
                    const int sc = scs[k], r = sc - w0;
                    if (sc == c && !keep_self) continue;
                    if (r >= 0 && r < PL_BITS) atomicOr(&bits[r >> 5], 1u << (r & 31));
                    else oow = true;
                }
            }
            continue;
        }
        for (int e = pl_row_begin(rowptr_in, in_stride, i), e1 = pl_row_end(rowptr_in, in_stride, i); e < e1; e++) {
            const int sc = cluster[src_in[e]];
            const int r = sc - w0;
            if (sc == c && !keep_self) continue;             // Pooling(self_loop=False) drops them (pooling.py:60-63)
            if (r >= 0 && r < PL_BITS) atomicOr(&bits[r >> 5], 1u << (r & 31));
            else oow = true;
        }
    }
    __syncwarp();
    if (!__any_sync(0xffffffffu, oow)) {
        constexpr int WPL = PL_BITS / 1024;                  // words per lane: lane l owns ids w0 + 32 WPL l ..
        uint32_t word[WPL];
        uint32_t pc = 0;
#pragma unroll
        for (int k = 0; k < WPL; k++) { word[k] = bits[lane * WPL + k]; pc += __popc(word[k]); }
        uint32_t inc = pc;
#pragma unroll
        for (int d = 1; d < 32; d <<= 1) {
            const uint32_t t = __shfl_up_sync(0xffffffffu, inc, d);
            if (lane >= d) inc += t;
        }
        const uint32_t total = __shfl_sync(0xffffffffu, inc, 31);
        const bool park = total <= PL_ROWCAP;
        if (park) {
            uint32_t off = inc - pc;
#pragma unroll
            for (int k = 0; k < WPL; k++)
                for (uint32_t wd = word[k]; wd; wd &= wd - 1)
                    tmp[(int64_t)c * PL_ROWCAP + off++] = w0 + (lane * WPL + k) * 32 + __ffs((int)wd) - 1;
        }
        if (lane == 0) { deg[c] = total; parked[c] = park ? 1 : 0; }
        return;
    }
    if (lane == 0) parked[c] = 0;
    // ---- general path: id range of the sources, then 4096-id chunks (count only; the FILL launch repeats the walk)
    int lo = 0x7fffffff, hi = -1;
    for (int p = s0 + lane; p < s1; p += 32) {
        const int i = (int)sidx[p];
        for (int e = pl_row_begin(rowptr_in, in_stride, i), e1 = pl_row_end(rowptr_in, in_stride, i); e < e1; e++) {
            const int sc = cluster[src_in[e]];
            if (sc != c || keep_self) { lo = min(lo, sc); hi = max(hi, sc); }
        }
    }
    lo = __reduce_min_sync(0xffffffffu, lo);
    hi = __reduce_max_sync(0xffffffffu, hi);
    uint32_t total = 0;
    for (int64_t wb = lo; wb <= hi; wb += PL_BITS) {
        __syncwarp();
#pragma unroll
        for (int k = 0; k < PL_BITS / 1024; k++) bits[lane + 32 * k] = 0u;
        __syncwarp();
        for (int p = s0 + lane; p < s1; p += 32) {
            const int i = (int)sidx[p];
            for (int e = pl_row_begin(rowptr_in, in_stride, i), e1 = pl_row_end(rowptr_in, in_stride, i); e < e1; e++) {
                const int sc = cluster[src_in[e]];
                const int64_t r = (int64_t)sc - wb;
                if ((sc != c || keep_self) && r >= 0 && r < PL_BITS) atomicOr(&bits[r >> 5], 1u << (r & 31));
            }
        }
        __syncwarp();
        uint32_t pc = 0;
#pragma unroll
        for (int k = 0; k < PL_BITS / 1024; k++) pc += __popc(bits[lane + 32 * k]);
        total += __reduce_add_sync(0xffffffffu, pc);
    }
    if (lane == 0) deg[c] = total;
}

// ------------------------------------------------------------------------------------------------ host side
struct PoolWs {
    uint32_t *k0, *v0, *k1, *v1, *flag, *excl, *deg, *rowptr_u, *sort_tmp, *scan_tmp;
    int32_t *seg_start, *row_tmp;
    uint8_t* parked;
};

static size_t pool_ws_layout(EvadArena& a, int64_t n_cap, int64_t c_cap, PoolWs* w) {
    const int64_t n = n_cap > 0 ? n_cap : 1;
    w->k0 = a.take<uint32_t>(n); w->v0 = a.take<uint32_t>(n);
    w->k1 = a.take<uint32_t>(n); w->v1 = a.take<uint32_t>(n);
    w->flag = a.take<uint32_t>(n);
    w->excl = a.take<uint32_t>(n);
    w->seg_start = a.take<int32_t>(c_cap + 2);
    w->deg = a.take<uint32_t>(c_cap + 2);
    w->rowptr_u = a.take<uint32_t>(c_cap + 2);
    w->row_tmp = a.take<int32_t>((size_t)(c_cap + 1) * PL_ROWCAP);
    w->parked = a.take<uint8_t>(c_cap + 2);
    w->sort_tmp = a.take<uint32_t>(evad_sort_tmp_elems(n));
    const int64_t m = n > c_cap + 2 ? n : c_cap + 2;
    w->scan_tmp = a.take<uint32_t>(evad_scan_tmp_elems(m));
    return a.off;
}

extern "C" size_t evad_voxel_pool_ws_bytes(int64_t n_cap, int64_t c_cap) {
    EvadArena a(nullptr, 0);
    PoolWs w;
    return pool_ws_layout(a, n_cap, c_cap, &w) + 256;
}

extern "C" int evad_voxel_pool(const evad_pool_params_t* p, int64_t n_cap, const int32_t* n_dev, const float* pos,
                               int ld_pos, const int32_t* batch, const float* x, int ldx, int cx,
                               const int32_t* rowptr_in, const int32_t* src_in, int64_t c_cap, int64_t e_cap,
                               int32_t* cluster, int32_t* n_out, float* x_out, int ldo, float* pos_out,
                               int32_t* batch_out, int32_t* rowptr_out, int32_t* src_out, int32_t* num_edges_out,
                               uint32_t* status, void* ws, size_t ws_bytes, void* stream) {
    cudaStream_t st = (cudaStream_t)stream;
    if (!p || !pos || !batch || !x || !cluster || !n_out || !x_out || !pos_out || !batch_out || !status || !ws)
        return EVAD_ERR_ARG;
    if (n_cap < 0 || c_cap < 1 || ld_pos < 3 || cx < 0 || p->batch_size < 1) return EVAD_ERR_ARG;
    const bool want_edges = rowptr_in && src_in && rowptr_out && src_out && num_edges_out;
    const int ctot = cx + (p->feat ? p->feat_c : 0);
    if (ldo < ctot + (p->append_pos ? 2 : 0)) return EVAD_ERR_ARG;
    if (ctot > 2 * PL_THREADS) return EVAD_ERR_UNSUPPORTED;
    const long long kb = (long long)p->nvx * p->nvy * p->nvt * ((long long)p->batch_size + 1);
    if (kb >= (1ll << 31)) return EVAD_ERR_UNSUPPORTED;
    EvadArena a(ws, ws_bytes);
    PoolWs w;
    pool_ws_layout(a, n_cap, c_cap, &w);
    if (!a.ok) return EVAD_ERR_WORKSPACE;
    PoolGeom G{p->size_x, p->size_y, p->size_t, p->nvx, p->nvy, p->nvt, (uint32_t)kb};
    PoolSample S{p->feat, p->feat_b, p->feat_h, p->feat_w, p->feat_c, p->width, p->height};
    const int64_t n = n_cap > 0 ? n_cap : 1;
    if (n_cap == 0) {
        cudaMemsetAsync(n_out, 0, sizeof(int32_t), st);
        if (want_edges) { cudaMemsetAsync(num_edges_out, 0, sizeof(int32_t), st); cudaMemsetAsync(rowptr_out, 0, sizeof(int32_t), st); }
        return EVAD_OK;
    }
    uint32_t *skey, *sidx;
    if (n_cap <= PFS_NMAX && kb <= PFS_KMAX) {
        const size_t fs_smem = (size_t)(2 * PFS_NMAX + 3 * PFS_KMAX) * sizeof(uint32_t);
        static bool fs_attr = false;
        if (!fs_attr) {
            if (cudaFuncSetAttribute(pool_front_small_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)fs_smem) !=
                cudaSuccess)
                return EVAD_ERR_LAUNCH;
            fs_attr = true;
        }
        skey = w.k1; sidx = w.v1;
        pool_front_small_kernel<<<1, PFS_THREADS, fs_smem, st>>>(pos, ld_pos, batch, (int)n_cap, n_dev, G, skey, sidx, cluster,
                                                                 w.seg_start, n_out, status);
        EVAD_CHECK_LAUNCH();
    } else {
        pool_keys_kernel<<<evad_cdiv(n, 256), 256, 0, st>>>(pos, ld_pos, batch, n_cap, n_dev, G, w.k0, status);
        EVAD_CHECK_LAUNCH();
        int key_bits = 1;
        while ((1ll << key_bits) <= kb) key_bits++;         // keys take values 0..kb (kb = padding)
        EVAD_TRY(evad_sort_pairs_u32(w.k0, w.v0, w.k1, w.v1, 1, n_cap, key_bits, w.sort_tmp, &skey, &sidx, st));
        pool_flag_kernel<<<evad_cdiv(n, 256), 256, 0, st>>>(skey, n_cap, G.key_bound, w.flag);
        EVAD_CHECK_LAUNCH();
        EVAD_TRY(evad_exclusive_scan_u32(w.flag, w.excl, n_cap, nullptr, w.scan_tmp, st));
        pool_relabel_kernel<<<evad_cdiv(n, 256), 256, 0, st>>>(skey, sidx, w.flag, w.excl, n_cap, G.key_bound, cluster,
                                                             w.seg_start, n_out);
        EVAD_CHECK_LAUNCH();
    }
    const int qf_h = p->feat ? p->feat_c / 4 : 0;
    const size_t prw_smem = (size_t)PRW_WARPS * (PRW_WIN * 4 * qf_h + 640) * sizeof(float);
    const bool quad_ok = cx % 4 == 0 && ldx % 4 == 0 && ldo % 4 == 0 && (((uintptr_t)x | (uintptr_t)x_out) & 15) == 0 &&
                         (!p->feat || (p->feat_c % 4 == 0 && ((uintptr_t)p->feat & 15) == 0)) && cx <= 256 &&
                         qf_h <= 64 && prw_smem <= 100 * 1024;
    if (quad_ok) {
        static size_t attr_bytes = 0;
        if (prw_smem > attr_bytes) {
            if (cudaFuncSetAttribute(pool_reduce_warp_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)prw_smem) !=
                cudaSuccess)
                return EVAD_ERR_LAUNCH;
            attr_bytes = prw_smem;
        }
        pool_reduce_warp_kernel<<<evad_cdiv(c_cap, PRW_WARPS), PRW_WARPS * 32, prw_smem, st>>>(
            sidx, w.seg_start, n_out, pos, ld_pos, batch, x, ldx, cx, S, p->aggr_mean, p->inv_w, p->inv_h, p->append_pos,
            x_out, ldo, pos_out, batch_out);
    } else {
        pool_reduce_kernel<<<(unsigned)c_cap, PL_THREADS, 0, st>>>(sidx, w.seg_start, n_out, pos, ld_pos, batch, x, ldx, cx,
                                                                  S, p->aggr_mean, p->inv_w, p->inv_h, p->append_pos,
                                                                  x_out, ldo, pos_out, batch_out);
    }
    EVAD_CHECK_LAUNCH();
    if (want_edges) {
        static const bool count_by_warp = [] { const char* e = getenv("EVAD_POOL_EDGES_WARP"); return !(e && e[0] == '0'); }();
        if (count_by_warp)
            pool_edges_count_warp_kernel<<<(unsigned)evad_cdiv(c_cap + 1, PEW_WARPS), PEW_WARPS * 32, 0, st>>>(
                sidx, w.seg_start, n_out, rowptr_in, p->ell_stride, p->self_loop, src_in, cluster, w.deg, w.row_tmp, w.parked);
        else
            pool_edges_kernel<false><<<(unsigned)(c_cap + 1), PL_THREADS, 0, st>>>(sidx, w.seg_start, n_out, rowptr_in, p->ell_stride, p->self_loop, src_in,
                                                                             cluster, w.deg, nullptr, nullptr, e_cap,
                                                                             w.row_tmp, w.parked, status, nullptr, nullptr);
        EVAD_CHECK_LAUNCH();
        EVAD_TRY(evad_exclusive_scan_u32(w.deg, w.rowptr_u, c_cap + 1, nullptr, w.scan_tmp, st));
        pool_edges_kernel<true><<<(unsigned)(c_cap + 1), PL_THREADS, 0, st>>>(sidx, w.seg_start, n_out, rowptr_in, p->ell_stride, p->self_loop, src_in,
                                                                            cluster, w.deg, w.rowptr_u, src_out, e_cap,
                                                                            w.row_tmp, w.parked, status, rowptr_out, num_edges_out);
        EVAD_CHECK_LAUNCH();
    }
    return EVAD_OK;
}

// ================================================================================================================
// Incremental voxel pooling (reference: src/dagr/asynchronous/max_pool.py:123-243 keeps, per voxel, the
// maximum with its arg-max and the homogeneous position sums, and updates the voxels that receive nodes).
//
// State per voxel KEY (dense key space, key_bound entries; caller-owned memory): running maximum of
// [x | sampled CNN map] (max is order independent: exact), serial position sums and member count (appended nodes have
// the highest ids, so "old sum + new members in ascending id" IS the dense kernel's ascending-node order: bit-identical),
// and one bitmap row of coarse in-edges (key of the source voxel).  evad_pool_inc_update folds appended nodes and their
// in-edges into the state; evad_pool_inc_emit writes the pooled graph exactly as evad_voxel_pool would for all nodes
// seen so far (dense ids = rank of the key among the occupied keys, sources ascending).  Max aggregation, no coarse
// self loops, fixed-stride fine rows.
struct PincState {
    float* vmax;      // [kb][ctot]
    float* vsum;      // [kb][3]
    int32_t* vcnt;    // [kb]
    int32_t* vbatch;  // [kb]
    uint32_t* adj;    // [kb][kw]
    int kw;
};

static size_t pinc_layout(void* state, long long kb, int ctot, PincState* s) {
    EvadArena a(state, (size_t)1 << 62);
    s->kw = (int)((kb + 31) / 32);
    s->vmax = a.take<float>((size_t)kb * ctot);
    s->vsum = a.take<float>((size_t)kb * 3);
    s->vcnt = a.take<int32_t>((size_t)kb);
    s->vbatch = a.take<int32_t>((size_t)kb);
    s->adj = a.take<uint32_t>((size_t)kb * s->kw);
    return a.off;
}

__global__ void __launch_bounds__(256) pinc_reset_kernel(float* __restrict__ vmax, int64_t n_max, uint32_t* __restrict__ rest,
                                                         int64_t n_rest) {
    const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n_max) vmax[i] = -INFINITY;
    if (i < n_rest) rest[i] = 0u;
}

__global__ void __launch_bounds__(256) pinc_keys_kernel(const float* __restrict__ pos, int ld_pos,
                                                        const int32_t* __restrict__ batch, int64_t n0, int64_t n_new,
                                                        PoolGeom G, uint32_t* __restrict__ nodekey, uint32_t* __restrict__ keys,
                                                        uint32_t* __restrict__ status) {
    const int64_t p = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (p >= n_new) return;
    const uint32_t k = pool_voxel_key(pos, ld_pos, batch, n0 + p, G, status);
    nodekey[n0 + p] = k;
    keys[p] = k;
}

// one thread per run of equal keys in the (stably) sorted new nodes: serial position sums in ascending node order
__global__ void __launch_bounds__(128) pinc_pos_kernel(const uint32_t* __restrict__ skey, const uint32_t* __restrict__ sidx,
                                                       int64_t n0, int64_t n_new, const float* __restrict__ pos, int ld_pos,
                                                       const int32_t* __restrict__ batch, PincState S) {
    const int64_t q0 = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (q0 >= n_new) return;
    const uint32_t key = skey[q0];
    if (q0 > 0 && skey[q0 - 1] == key) return;
    float sx = S.vsum[3 * (int64_t)key], sy = S.vsum[3 * (int64_t)key + 1], st = S.vsum[3 * (int64_t)key + 2];
    int cnt = S.vcnt[key], last = 0;
    for (int64_t q = q0; q < n_new && skey[q] == key; q++) {
        const int64_t i = n0 + sidx[q];
        sx = __fadd_rn(sx, pos[i * ld_pos]);
        sy = __fadd_rn(sy, pos[i * ld_pos + 1]);
        st = __fadd_rn(st, pos[i * ld_pos + 2]);
        cnt++;
        last = batch[i];
    }
    S.vsum[3 * (int64_t)key] = sx; S.vsum[3 * (int64_t)key + 1] = sy; S.vsum[3 * (int64_t)key + 2] = st;
    S.vcnt[key] = cnt;
    S.vbatch[key] = last;
}

__device__ __forceinline__ void pinc_atomic_max(float* addr, float v) {
    if (v >= 0.f) atomicMax(reinterpret_cast<int*>(addr), __float_as_int(v));
    else atomicMin(reinterpret_cast<unsigned int*>(addr), __float_as_uint(v));
}

// one thread per (new node, channel quad): the node's [x | sampled map] quad folded into its voxel's running maximum
__global__ void __launch_bounds__(256) pinc_feat_kernel(const uint32_t* __restrict__ nodekey, int64_t n0, int64_t n_new,
                                                        const float* __restrict__ pos, int ld_pos,
                                                        const int32_t* __restrict__ batch, const float* __restrict__ x,
                                                        int ldx, int cx, PoolSample Sm, PincState S, int ctot) {
    const int qx = cx >> 2, qf = Sm.feat ? (Sm.C >> 2) : 0, qt = qx + qf;
    const int64_t t = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    const int64_t p = t / qt;
    const int q = (int)(t - p * qt);
    if (p >= n_new) return;
    const int64_t i = n0 + p;
    float4 v;
    if (q < qx) {
        v = __ldg(reinterpret_cast<const float4*>(x + i * ldx) + q);
    } else {
        int tap[8];
        float wgt[8];
        sample_taps(Sm, pos[i * ld_pos], pos[i * ld_pos + 1], (float)batch[i], tap, wgt);
        v = make_float4(0.f, 0.f, 0.f, 0.f);
        const float4* f4 = reinterpret_cast<const float4*>(Sm.feat);
#pragma unroll
        for (int k = 0; k < 8; k++) {
            if (tap[k] < 0 || wgt[k] == 0.f) continue;
            const float4 f = __ldg(f4 + (int64_t)tap[k] * qf + (q - qx));
            v.x = fmaf(f.x, wgt[k], v.x); v.y = fmaf(f.y, wgt[k], v.y);
            v.z = fmaf(f.z, wgt[k], v.z); v.w = fmaf(f.w, wgt[k], v.w);
        }
    }
    float* dst = S.vmax + (int64_t)nodekey[i] * ctot + 4 * q;
    pinc_atomic_max(dst, v.x); pinc_atomic_max(dst + 1, v.y); pinc_atomic_max(dst + 2, v.z); pinc_atomic_max(dst + 3, v.w);
}

// one thread per slot of a new node's fixed-stride row: coarse in-edge (voxel of the source -> voxel of the node)
__global__ void __launch_bounds__(256) pinc_adj_kernel(const uint32_t* __restrict__ nodekey, int64_t n0, int64_t n_new,
                                                       const int32_t* __restrict__ count, const int32_t* __restrict__ nbr,
                                                       int K, PincState S) {
    const int64_t t = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    const int64_t p = t / K;
    const int s = (int)(t - p * K);
    if (p >= n_new) return;
    const int64_t i = n0 + p;
    if (s >= count[i]) return;
    const uint32_t kd = nodekey[i], ks = nodekey[nbr[i * K + s]];
    if (ks != kd) atomicOr(S.adj + (int64_t)kd * S.kw + (ks >> 5), 1u << (ks & 31));
}

// ---- emit: the pooled graph of everything folded in so far
__global__ void __launch_bounds__(256) pinc_flag_kernel(const int32_t* __restrict__ vcnt, int64_t kb, uint32_t* __restrict__ flag) {
    const int64_t k = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (k < kb) flag[k] = vcnt[k] > 0 ? 1u : 0u;
}

// one warp per occupied key: node row [max | pos_x pos_y | 0], position (mean + round_to_pixel), batch, in-degree
__global__ void __launch_bounds__(256) pinc_emit_nodes_kernel(PincState S, int64_t kb, int ctot, const uint32_t* __restrict__ flag,
                                                              const uint32_t* __restrict__ rank, int64_t c_cap, float inv_w,
                                                              float inv_h, int append_pos, float* __restrict__ x_out, int ldo,
                                                              float* __restrict__ pos_out, int32_t* __restrict__ batch_out,
                                                              uint32_t* __restrict__ deg, uint32_t* __restrict__ status) {
    const int lane = threadIdx.x & 31;
    const int64_t key = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    if (key >= kb || !flag[key]) return;
    const int64_t c = rank[key];
    if (c >= c_cap) {
        if (lane == 0) atomicOr(status, EVAD_STATUS_CLUSTER_OOB);
        return;
    }
    const float cntf = fmaxf((float)S.vcnt[key], 1.f);
    float mx = __fdiv_rn(S.vsum[3 * key], cntf), my = __fdiv_rn(S.vsum[3 * key + 1], cntf);
    const float mt = __fdiv_rn(S.vsum[3 * key + 2], cntf);
    mx = __fmul_rn(div_floor_f(__fadd_rn(mx, 1e-5f), inv_w), inv_w);
    my = __fmul_rn(div_floor_f(__fadd_rn(my, 1e-5f), inv_h), inv_h);
    if (lane == 0) {
        pos_out[c * 3] = mx; pos_out[c * 3 + 1] = my; pos_out[c * 3 + 2] = mt;
        batch_out[c] = S.vbatch[key];
    }
    float* o = x_out + c * ldo;
    const float* v = S.vmax + key * ctot;
    for (int ch = lane; ch < ldo; ch += 32) {
        float val = 0.f;
        if (ch < ctot) val = v[ch];
        else if (append_pos && ch < ctot + 2) val = ch == ctot ? mx : my;
        o[ch] = val;
    }
    uint32_t d = 0;
    const uint32_t* row = S.adj + key * S.kw;
    for (int w = lane; w < S.kw; w += 32) d += __popc(row[w]);
#pragma unroll
    for (int s = 16; s > 0; s >>= 1) d += __shfl_xor_sync(0xffffffffu, d, s);
    if (lane == 0) deg[c] = d;
}

// one warp per occupied key: the set bits of its row, ascending, as dense source ids
__global__ void __launch_bounds__(256) pinc_emit_edges_kernel(PincState S, int64_t kb, const uint32_t* __restrict__ flag,
                                                              const uint32_t* __restrict__ rank, int64_t c_cap,
                                                              const uint32_t* __restrict__ rowptr_u, int64_t e_cap,
                                                              int32_t* __restrict__ rowptr_out, int32_t* __restrict__ src_out,
                                                              int32_t* __restrict__ num_edges_out, const int32_t* __restrict__ n_out,
                                                              uint32_t* __restrict__ status) {
    const int lane = threadIdx.x & 31;
    const int64_t key = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    if (key == 0 && lane == 0) {
        const int n = min((int64_t)*n_out, c_cap);
        const uint32_t tot = rowptr_u[n];
        if ((int64_t)tot > e_cap) atomicOr(status, EVAD_STATUS_EDGE_OVERFLOW);
        *num_edges_out = (int32_t)min((int64_t)tot, e_cap);
        rowptr_out[n] = (int32_t)tot;
    }
    if (key >= kb || !flag[key]) return;
    const int64_t c = rank[key];
    if (c >= c_cap) return;
    uint32_t base = rowptr_u[c];
    if (lane == 0) rowptr_out[c] = (int32_t)base;
    const uint32_t* row = S.adj + key * S.kw;
    for (int w0 = 0; w0 < S.kw; w0 += 32) {
        const int w = w0 + lane;
        uint32_t bits = w < S.kw ? row[w] : 0u;
        const uint32_t cntw = __popc(bits);
        uint32_t inc = cntw;                                  // inclusive prefix over the lanes
#pragma unroll
        for (int s = 1; s < 32; s <<= 1) {
            const uint32_t t = __shfl_up_sync(0xffffffffu, inc, s);
            if (lane >= s) inc += t;
        }
        uint32_t o = base + inc - cntw;
        while (bits) {
            const int b = __ffs(bits) - 1;
            bits &= bits - 1;
            if ((int64_t)o < e_cap) src_out[o] = (int32_t)rank[(int64_t)w * 32 + b];
            o++;
        }
        base += __shfl_sync(0xffffffffu, inc, 31);
    }
}

static int pinc_check(const evad_pool_params_t* p, int cx, long long* kb, int* ctot) {
    if (!p || cx < 0 || cx % 4 != 0 || p->batch_size < 1) return EVAD_ERR_ARG;
    if (p->aggr_mean || p->self_loop || p->ell_stride < 1) return EVAD_ERR_UNSUPPORTED;   // max pooling over fixed-stride rows
    if (p->feat && (p->feat_c % 4 != 0 || ((uintptr_t)p->feat & 15) != 0)) return EVAD_ERR_UNSUPPORTED;
    *kb = (long long)p->nvx * p->nvy * p->nvt * ((long long)p->batch_size + 1);
    if (*kb >= (1ll << 24)) return EVAD_ERR_UNSUPPORTED;      // the coarse adjacency is a dense bitmap: kb^2 / 8 bytes
    *ctot = cx + (p->feat ? p->feat_c : 0);
    return EVAD_OK;
}

extern "C" size_t evad_pool_inc_state_bytes(const evad_pool_params_t* p, int cx) {
    long long kb;
    int ctot;
    if (pinc_check(p, cx, &kb, &ctot) != EVAD_OK) return 0;
    PincState s;
    return pinc_layout(nullptr, kb, ctot, &s) + 256;
}

extern "C" int evad_pool_inc_reset(void* state, const evad_pool_params_t* p, int cx, void* stream) {
    long long kb;
    int ctot;
    EVAD_TRY(pinc_check(p, cx, &kb, &ctot));
    if (!state || ((uintptr_t)state & 255) != 0) return EVAD_ERR_ARG;
    PincState s;
    const size_t total = pinc_layout(state, kb, ctot, &s);
    const int64_t n_max = kb * ctot;
    uint32_t* rest = reinterpret_cast<uint32_t*>(s.vsum);
    const int64_t n_rest = (int64_t)((reinterpret_cast<char*>(state) + total) - reinterpret_cast<char*>(rest)) / 4;
    const int64_t n = n_max > n_rest ? n_max : n_rest;
    pinc_reset_kernel<<<evad_cdiv(n, 256), 256, 0, (cudaStream_t)stream>>>(s.vmax, n_max, rest, n_rest);
    EVAD_CHECK_LAUNCH();
    return EVAD_OK;
}

extern "C" size_t evad_pool_inc_update_ws_bytes(int64_t n_new) {
    const size_t n = (size_t)(n_new > 0 ? n_new : 1);
    return evad_align(n * 4) * 4 + evad_align(evad_sort_tmp_elems(n_new) * 4) + 256;
}

// folds the nodes n0 .. n0 + n_new - 1 (positions, batch, features x [., ldx] with cx channels, fixed-stride rows
// count / nbr of K = p->ell_stride) into the state; node_key [>= n0 + n_new] is the caller's per-node key array
extern "C" int evad_pool_inc_update(void* state, const evad_pool_params_t* p, int64_t n0, int64_t n_new, const float* pos,
                                    int ld_pos, const int32_t* batch, const float* x, int ldx, int cx, const int32_t* count,
                                    const int32_t* nbr, uint32_t* node_key, uint32_t* status, void* ws, size_t ws_bytes,
                                    void* stream) {
    long long kb;
    int ctot;
    EVAD_TRY(pinc_check(p, cx, &kb, &ctot));
    if (!state || !pos || !batch || !x || !count || !nbr || !node_key || !status || !ws || n0 < 0 || n_new < 0 || ld_pos < 3 ||
        ldx % 4 != 0 || ((uintptr_t)x & 15) != 0)
        return EVAD_ERR_ARG;
    if (ws_bytes < evad_pool_inc_update_ws_bytes(n_new)) return EVAD_ERR_WORKSPACE;
    if (n_new == 0) return EVAD_OK;
    cudaStream_t st = (cudaStream_t)stream;
    PincState s;
    pinc_layout(state, kb, ctot, &s);
    EvadArena a(ws, ws_bytes);
    uint32_t* k0 = a.take<uint32_t>(n_new);
    uint32_t* v0 = a.take<uint32_t>(n_new);
    uint32_t* k1 = a.take<uint32_t>(n_new);
    uint32_t* v1 = a.take<uint32_t>(n_new);
    uint32_t* sort_tmp = a.take<uint32_t>(evad_sort_tmp_elems(n_new));
    if (!a.ok) return EVAD_ERR_WORKSPACE;
    PoolGeom G{p->size_x, p->size_y, p->size_t, p->nvx, p->nvy, p->nvt, (uint32_t)kb};
    PoolSample S{p->feat, p->feat_b, p->feat_h, p->feat_w, p->feat_c, p->width, p->height};
    pinc_keys_kernel<<<evad_cdiv(n_new, 256), 256, 0, st>>>(pos, ld_pos, batch, n0, n_new, G, node_key, k0, status);
    EVAD_CHECK_LAUNCH();
    int key_bits = 1;
    while ((1ll << key_bits) < kb) key_bits++;
    uint32_t *skey, *sidx;
    EVAD_TRY(evad_sort_pairs_u32(k0, v0, k1, v1, 1, n_new, key_bits, sort_tmp, &skey, &sidx, st));
    pinc_pos_kernel<<<evad_cdiv(n_new, 128), 128, 0, st>>>(skey, sidx, n0, n_new, pos, ld_pos, batch, s);
    EVAD_CHECK_LAUNCH();
    const int qt = ctot / 4;
    pinc_feat_kernel<<<evad_cdiv(n_new * qt, 256), 256, 0, st>>>(node_key, n0, n_new, pos, ld_pos, batch, x, ldx, cx, S, s, ctot);
    EVAD_CHECK_LAUNCH();
    pinc_adj_kernel<<<evad_cdiv(n_new * p->ell_stride, 256), 256, 0, st>>>(node_key, n0, n_new, count, nbr, p->ell_stride, s);
    EVAD_CHECK_LAUNCH();
    return EVAD_OK;
}

extern "C" size_t evad_pool_inc_emit_ws_bytes(const evad_pool_params_t* p, int64_t c_cap) {
    long long kb = (long long)p->nvx * p->nvy * p->nvt * ((long long)p->batch_size + 1);
    return evad_align((size_t)kb * 4) * 2 + evad_align((size_t)(c_cap + 2) * 4) * 2 +
           evad_align(evad_scan_tmp_elems(kb > c_cap + 1 ? kb : c_cap + 1) * 4) + 256;
}

// writes the pooled graph of all folded nodes in the layout of evad_voxel_pool (same argument meaning)
extern "C" int evad_pool_inc_emit(void* state, const evad_pool_params_t* p, int cx, int64_t c_cap, int64_t e_cap,
                                  int32_t* n_out, float* x_out, int ldo, float* pos_out, int32_t* batch_out,
                                  int32_t* rowptr_out, int32_t* src_out, int32_t* num_edges_out, uint32_t* status, void* ws,
                                  size_t ws_bytes, void* stream) {
    long long kb;
    int ctot;
    EVAD_TRY(pinc_check(p, cx, &kb, &ctot));
    if (!state || !n_out || !x_out || !pos_out || !batch_out || !rowptr_out || !src_out || !num_edges_out || !status || !ws ||
        c_cap < 1 || e_cap < 1 || ldo < ctot + (p->append_pos ? 2 : 0))
        return EVAD_ERR_ARG;
    if (ws_bytes < evad_pool_inc_emit_ws_bytes(p, c_cap)) return EVAD_ERR_WORKSPACE;
    cudaStream_t st = (cudaStream_t)stream;
    PincState s;
    pinc_layout(state, kb, ctot, &s);
    EvadArena a(ws, ws_bytes);
    uint32_t* flag = a.take<uint32_t>(kb);
    uint32_t* rank = a.take<uint32_t>(kb);
    uint32_t* deg = a.take<uint32_t>(c_cap + 2);
    uint32_t* rowptr_u = a.take<uint32_t>(c_cap + 2);
    uint32_t* scan_tmp = a.take<uint32_t>(evad_scan_tmp_elems(kb > c_cap + 1 ? kb : c_cap + 1));
    if (!a.ok) return EVAD_ERR_WORKSPACE;
    pinc_flag_kernel<<<evad_cdiv(kb, 256), 256, 0, st>>>(s.vcnt, kb, flag);
    EVAD_CHECK_LAUNCH();
    EVAD_TRY(evad_exclusive_scan_u32(flag, rank, kb, reinterpret_cast<uint32_t*>(n_out), scan_tmp, st));
    if (cudaMemsetAsync(deg, 0, (size_t)(c_cap + 2) * 4, st) != cudaSuccess) return EVAD_ERR_LAUNCH;
    pinc_emit_nodes_kernel<<<evad_cdiv(kb * 32, 256), 256, 0, st>>>(s, kb, ctot, flag, rank, c_cap, p->inv_w, p->inv_h,
                                                                    p->append_pos, x_out, ldo, pos_out, batch_out, deg, status);
    EVAD_CHECK_LAUNCH();
    EVAD_TRY(evad_exclusive_scan_u32(deg, rowptr_u, c_cap + 1, nullptr, scan_tmp, st));
    pinc_emit_edges_kernel<<<evad_cdiv(kb * 32, 256), 256, 0, st>>>(s, kb, flag, rank, c_cap, rowptr_u, e_cap, rowptr_out,
                                                                    src_out, num_edges_out, n_out, status);
    EVAD_CHECK_LAUNCH();
    return EVAD_OK;
}
