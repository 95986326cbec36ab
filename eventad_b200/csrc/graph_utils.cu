// eventad_b200/csrc/graph_utils.cu -- format glue around the kernels: COO <-> CSR, Cartesian edge attributes,
// image-feature sampling + concat.
//
// Reference pieces replaced (paths relative to /root/reference/src/dagr/model):
//   layers/components.py:25-35 + networks/net.py:120-121  Cartesian edge attributes (+ clamp on the first layer)
//   networks/net.py:15-17,122-123,200-228                 sampling_skip (grid_sample + cat) and cat(x, pos[:, :2])
//   layers/pooling.py:64                                  `unique(dim=-1)` ordering (src-major) of coarse edges
#include "common.cuh"

// ------------------------------------------------------------------------------------------------ COO -> CSR
__global__ void __launch_bounds__(256) coo_keys_kernel(const int64_t* __restrict__ dst, int64_t e, int64_t n,
                                                       uint32_t* __restrict__ keys, uint32_t* __restrict__ status) {
    int64_t t = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (t >= e) return;
    int64_t d = dst[t];
    if (d < 0 || d >= n) { atomicOr(status, EVAD_STATUS_EVENT_OOB); d = 0; }
    keys[t] = (uint32_t)d;
}

__global__ void __launch_bounds__(256) coo_finish_kernel(const uint32_t* __restrict__ skey,
                                                         const uint32_t* __restrict__ seid,
                                                         const int64_t* __restrict__ src64, int64_t e, int64_t n,
                                                         int32_t* __restrict__ rowptr, int32_t* __restrict__ src,
                                                         int32_t* __restrict__ eid) {
    int64_t t = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (t <= n) {                                            // rowptr[t] = lower_bound(skey, t)
        int64_t lo = 0, hi = e;
        while (lo < hi) {
            int64_t mid = (lo + hi) >> 1;
            if ((int64_t)skey[mid] < t) lo = mid + 1; else hi = mid;
        }
        rowptr[t] = (int32_t)lo;
    }
    if (t < e) {
        const uint32_t id = seid[t];
        src[t] = (int32_t)src64[id];
        eid[t] = (int32_t)id;
    }
}

extern "C" size_t evad_coo_to_csr_ws_bytes(int64_t e) {
    const int64_t m = e > 0 ? e : 1;
    return 4 * evad_align((size_t)m * 4) + evad_sort_tmp_elems(m) * 4 + 512;
}

extern "C" int evad_coo_to_csr(const int64_t* edge_index, int64_t stride, int64_t e, int64_t n, int32_t* rowptr,
                               int32_t* src, int32_t* eid, uint32_t* status, void* ws, size_t ws_bytes, void* stream) {
    cudaStream_t st = (cudaStream_t)stream;
    if (!rowptr || !src || !eid || !status || e < 0 || n < 0 || n >= (1ll << 31) || e >= (1ll << 31)) return EVAD_ERR_ARG;
    if (e > 0 && (!edge_index || !ws)) return EVAD_ERR_ARG;
    if (ws_bytes < evad_coo_to_csr_ws_bytes(e)) return EVAD_ERR_WORKSPACE;
    EvadArena a(ws, ws_bytes);
    const int64_t m = e > 0 ? e : 1;
    uint32_t* k0 = a.take<uint32_t>(m); uint32_t* v0 = a.take<uint32_t>(m);
    uint32_t* k1 = a.take<uint32_t>(m); uint32_t* v1 = a.take<uint32_t>(m);
    uint32_t* tmp = a.take<uint32_t>(evad_sort_tmp_elems(m));
    uint32_t *sk = k0, *sv = v0;
    if (e > 0) {
        coo_keys_kernel<<<evad_cdiv(e, 256), 256, 0, st>>>(edge_index + stride, e, n, k0, status);
        EVAD_CHECK_LAUNCH();
        int bits = 1;
        while ((1ll << bits) < n) bits++;
        EVAD_TRY(evad_sort_pairs_u32(k0, v0, k1, v1, 1, e, bits, tmp, &sk, &sv, st));
    }
    const int64_t work = (n + 1 > e) ? n + 1 : e;
    coo_finish_kernel<<<evad_cdiv(work, 256), 256, 0, st>>>(sk, sv, edge_index, e, n, rowptr, src, eid);
    EVAD_CHECK_LAUNCH();
    return EVAD_OK;
}

// ------------------------------------------------------------------------------------------------ CSR -> COO
__global__ void __launch_bounds__(256) csr_expand_kernel(const int32_t* __restrict__ rowptr,
                                                         const int32_t* __restrict__ src, int64_t n_cap,
                                                         const int32_t* __restrict__ n_dev, int64_t e_cap,
                                                         int64_t* __restrict__ out_src, int64_t* __restrict__ out_dst,
                                                         uint32_t* __restrict__ keys, uint32_t* __restrict__ dsts) {
    int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    const int64_t n = n_dev ? min((int64_t)*n_dev, n_cap) : n_cap;
    if (i >= n) return;
    for (int e = rowptr[i]; e < rowptr[i + 1] && e < e_cap; e++) {
        if (out_src) { out_src[e] = src[e]; out_dst[e] = i; }
        if (keys) { keys[e] = (uint32_t)src[e]; dsts[e] = (uint32_t)i; }
    }
}

__global__ void __launch_bounds__(256) csr_pad_keys_kernel(const int32_t* __restrict__ num_edges, int64_t e_cap,
                                                           uint32_t pad, uint32_t* __restrict__ keys) {
    int64_t e = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (e < e_cap && e >= *num_edges) keys[e] = pad;
}

__global__ void __launch_bounds__(256) csr_srcmajor_kernel(const uint32_t* __restrict__ skey,
                                                           const uint32_t* __restrict__ sslot,
                                                           const uint32_t* __restrict__ dsts,
                                                           const int32_t* __restrict__ num_edges, int64_t e_cap,
                                                           int64_t* __restrict__ out_src, int64_t* __restrict__ out_dst,
                                                           int32_t* __restrict__ slot_of) {
    int64_t q = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (q >= e_cap || q >= *num_edges) return;
    out_src[q] = skey[q];
    out_dst[q] = dsts[sslot[q]];
    if (slot_of) slot_of[q] = (int32_t)sslot[q];
}

extern "C" size_t evad_csr_to_coo_ws_bytes(int64_t e_cap) {
    const int64_t m = e_cap > 0 ? e_cap : 1;
    return 5 * evad_align((size_t)m * 4) + evad_sort_tmp_elems(m) * 4 + 512;
}

extern "C" int evad_csr_to_coo(const int32_t* rowptr, const int32_t* src, int64_t n_cap, const int32_t* n_dev,
                               int64_t e_cap, const int32_t* num_edges, int src_major, int64_t* edge_index,
                               int64_t stride, int32_t* slot_of, void* ws, size_t ws_bytes, void* stream) {
    cudaStream_t st = (cudaStream_t)stream;
    if (!rowptr || !src || !num_edges || !edge_index || n_cap < 0 || e_cap < 0) return EVAD_ERR_ARG;
    if (n_cap == 0 || e_cap == 0) return EVAD_OK;
    if (!src_major) {
        csr_expand_kernel<<<evad_cdiv(n_cap, 256), 256, 0, st>>>(rowptr, src, n_cap, n_dev, e_cap, edge_index,
                                                               edge_index + stride, nullptr, nullptr);
        EVAD_CHECK_LAUNCH();
        return EVAD_OK;
    }
    if (!ws || ws_bytes < evad_csr_to_coo_ws_bytes(e_cap)) return EVAD_ERR_WORKSPACE;
    EvadArena a(ws, ws_bytes);
    uint32_t* k0 = a.take<uint32_t>(e_cap); uint32_t* v0 = a.take<uint32_t>(e_cap);
    uint32_t* k1 = a.take<uint32_t>(e_cap); uint32_t* v1 = a.take<uint32_t>(e_cap);
    uint32_t* dsts = a.take<uint32_t>(e_cap);
    uint32_t* tmp = a.take<uint32_t>(evad_sort_tmp_elems(e_cap));
    csr_expand_kernel<<<evad_cdiv(n_cap, 256), 256, 0, st>>>(rowptr, src, n_cap, n_dev, e_cap, nullptr, nullptr, k0, dsts);
    EVAD_CHECK_LAUNCH();
    csr_pad_keys_kernel<<<evad_cdiv(e_cap, 256), 256, 0, st>>>(num_edges, e_cap, (uint32_t)n_cap, k0);
    EVAD_CHECK_LAUNCH();
    int bits = 1;
    while ((1ll << bits) <= n_cap) bits++;
    uint32_t *sk, *sv;
    EVAD_TRY(evad_sort_pairs_u32(k0, v0, k1, v1, 1, e_cap, bits, tmp, &sk, &sv, st));
    csr_srcmajor_kernel<<<evad_cdiv(e_cap, 256), 256, 0, st>>>(sk, sv, dsts, num_edges, e_cap, edge_index,
                                                             edge_index + stride, slot_of);
    EVAD_CHECK_LAUNCH();
    return EVAD_OK;
}

// ------------------------------------------------------------------------------------------------ Cartesian
__global__ void __launch_bounds__(256) cartesian_kernel(const float* __restrict__ pos, int ld_pos, int dims,
                                                        const int64_t* __restrict__ row, const int64_t* __restrict__ col,
                                                        int64_t e_cap, const int32_t* __restrict__ num_edges,
                                                        float two_m, int clamp01, float* __restrict__ out) {
    int64_t e = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    const int64_t ne = num_edges ? min((int64_t)*num_edges, e_cap) : e_cap;
    if (e >= ne) return;
    const int64_t j = row[e], i = col[e];
    const float inv = __fdiv_rn(1.0f, two_m);      // tensor / Python scalar on CUDA = multiply by the fp32 reciprocal
    for (int d = 0; d < dims; d++) {
        float a = __fadd_rn(__fmul_rn(__fsub_rn(pos[j * ld_pos + d], pos[i * ld_pos + d]), inv), 0.5f);
        if (clamp01) a = fminf(fmaxf(a, 0.f), 1.f);
        out[e * dims + d] = a;
    }
}

extern "C" int evad_cartesian(const float* pos, int ld_pos, int dims, const int64_t* edge_index, int64_t stride,
                              int64_t e_cap, const int32_t* num_edges, float two_m, int clamp01, float* out,
                              void* stream) {
    if (!pos || !edge_index || !out || dims < 1 || ld_pos < dims || e_cap < 0) return EVAD_ERR_ARG;
    if (e_cap == 0) return EVAD_OK;
    cartesian_kernel<<<evad_cdiv(e_cap, 256), 256, 0, (cudaStream_t)stream>>>(pos, ld_pos, dims, edge_index,
                                                                            edge_index + stride, e_cap, num_edges,
                                                                            two_m, clamp01, out);
    EVAD_CHECK_LAUNCH();
    return EVAD_OK;
}

// ------------------------------------------------------------------------------------------------ sampling + concat
#include "sample.cuh"

// out[i, :] = [ x[i, 0:cx] | sample(feat)(C) | pos_x, pos_y (append_pos) | 0 ... ]  up to ldo columns
__global__ void __launch_bounds__(256) sample_concat_kernel(const float* __restrict__ x, int ldx, int cx,
                                                            const float* __restrict__ pos, int ld_pos,
                                                            const int32_t* __restrict__ batch, int64_t n_cap,
                                                            const int32_t* __restrict__ n_dev, PoolSample S,
                                                            int append_pos, float* __restrict__ out, int ldo) {
    const int64_t t = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    const int64_t n = n_dev ? min((int64_t)*n_dev, n_cap) : n_cap;
    const int64_t i = t / ldo;
    const int c = (int)(t - i * ldo);
    if (i >= n) return;
    const int cs = S.feat ? S.C : 0;
    float v = 0.f;
    if (c < cx) v = x[i * ldx + c];
    else if (c < cx + cs) v = sample_feat(S, pos[i * ld_pos], pos[i * ld_pos + 1], (float)batch[i], c - cx);
    else if (append_pos && c < cx + cs + 2) v = pos[i * ld_pos + (c - cx - cs)];
    out[t] = v;
}

// Vectorised variant (C % 4 == 0, channel-last map 16-byte aligned): one thread per (node, channel quad).  The up to 8
// trilinear taps are resolved once per thread and each tap is one 128-bit load; accumulation order per channel is the
// order of sample_feat (dz, dy, dx; out-of-bounds taps skipped), so both kernels give identical bits.
__global__ void __launch_bounds__(256) sample_concat_quad_kernel(const float* __restrict__ x, int ldx, int cx,
                                                                 const float* __restrict__ pos, int ld_pos,
                                                                 const int32_t* __restrict__ batch, int64_t n_cap,
                                                                 const int32_t* __restrict__ n_dev, PoolSample S,
                                                                 int append_pos, float* __restrict__ out, int ldo) {
    const int quads = S.C >> 2;
    const int64_t t = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    const int64_t n = n_dev ? min((int64_t)*n_dev, n_cap) : n_cap;
    const int64_t i = t / quads;
    const int q = (int)(t - i * quads);
    if (i >= n) return;
    const float px = pos[i * ld_pos], py = pos[i * ld_pos + 1];
    int tap[8];
    float wgt[8];
    sample_taps(S, px, py, (float)batch[i], tap, wgt);
    float4 acc = make_float4(0.f, 0.f, 0.f, 0.f);
    const float4* f4 = reinterpret_cast<const float4*>(S.feat);
#pragma unroll
    for (int k = 0; k < 8; k++) {
        // zero weights (the whole dz = 1 plane when the depth coordinate is the integer batch index): fmaf(v, 0, acc) == acc
        if (tap[k] < 0 || wgt[k] == 0.f) continue;
        const float4 v = __ldg(f4 + (int64_t)tap[k] * quads + q);
        acc.x = fmaf(v.x, wgt[k], acc.x); acc.y = fmaf(v.y, wgt[k], acc.y);
        acc.z = fmaf(v.z, wgt[k], acc.z); acc.w = fmaf(v.w, wgt[k], acc.w);
    }
    float* o = out + i * ldo;
    o[cx + 4 * q] = acc.x; o[cx + 4 * q + 1] = acc.y; o[cx + 4 * q + 2] = acc.z; o[cx + 4 * q + 3] = acc.w;
    if (q == 0) {
        for (int c = 0; c < cx; c++) o[c] = x[i * ldx + c];
        int c = cx + S.C;
        if (append_pos) { o[c] = px; o[c + 1] = py; c += 2; }
        for (; c < ldo; c++) o[c] = 0.f;
    }
}

// Layer-1 shape (one polarity channel + 16 sampled channels + position, rows of 20 floats): one thread per node builds
// the whole row in registers -- taps resolved once per node instead of once per channel quad, and the row leaves as five
// 128-bit stores instead of twenty 32-bit ones.
// H2: the row leaves as "h2 words" (evad_h2_split), the input format of evad_spline_conv_l1_h16.
template <int CX, int C, int LDO, bool H2>
__global__ void __launch_bounds__(256) sample_concat_row_kernel(const float* __restrict__ x, int ldx,
                                                                const float* __restrict__ pos, int ld_pos,
                                                                const int32_t* __restrict__ batch, int64_t n_cap,
                                                                const int32_t* __restrict__ n_dev, PoolSample S,
                                                                int append_pos, float* __restrict__ out,
                                                                uint32_t* __restrict__ status,
                                                                const int32_t* __restrict__ visit) {
    const int64_t tix = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    const int64_t n = n_dev ? min((int64_t)*n_dev, n_cap) : n_cap;
    if (tix >= n) return;
    const int64_t i = visit ? visit[tix] : tix;               // visiting order only: row i is written either way
    const float px = pos[i * ld_pos], py = pos[i * ld_pos + 1];
    int tap[8];
    float wgt[8];
    sample_taps(S, px, py, (float)batch[i], tap, wgt);
    float row[LDO];
#pragma unroll
    for (int c = 0; c < LDO; c++) row[c] = 0.f;
#pragma unroll
    for (int c = 0; c < CX; c++) row[c] = x[i * ldx + c];
    const float4* f4 = reinterpret_cast<const float4*>(S.feat);
#pragma unroll
    for (int k = 0; k < 8; k++) {
        if (tap[k] < 0 || wgt[k] == 0.f) continue;            // fmaf(v, 0, acc) == acc: same bits as the chain with it
#pragma unroll
        for (int q = 0; q < C / 4; q++) {
            const float4 v = __ldg(f4 + (int64_t)tap[k] * (C / 4) + q);
            row[CX + 4 * q] = fmaf(v.x, wgt[k], row[CX + 4 * q]);
            row[CX + 4 * q + 1] = fmaf(v.y, wgt[k], row[CX + 4 * q + 1]);
            row[CX + 4 * q + 2] = fmaf(v.z, wgt[k], row[CX + 4 * q + 2]);
            row[CX + 4 * q + 3] = fmaf(v.w, wgt[k], row[CX + 4 * q + 3]);
        }
    }
    if (append_pos) { row[CX + C] = px; row[CX + C + 1] = py; }
    if (H2) {
        bool ok = true;
#pragma unroll
        for (int c = 0; c < LDO; c++) ok = ok && fabsf(row[c]) < EVAD_H2_RANGE;
        if (!ok && status) atomicOr(status, EVAD_STATUS_H16_RANGE);
        uint4* o = reinterpret_cast<uint4*>(out + i * LDO);
#pragma unroll
        for (int c = 0; c < LDO / 4; c++)
            o[c] = make_uint4(evad_h2_split(row[4 * c]), evad_h2_split(row[4 * c + 1]), evad_h2_split(row[4 * c + 2]),
                              evad_h2_split(row[4 * c + 3]));
        return;
    }
    float4* o = reinterpret_cast<float4*>(out + i * LDO);
#pragma unroll
    for (int c = 0; c < LDO / 4; c++) o[c] = make_float4(row[4 * c], row[4 * c + 1], row[4 * c + 2], row[4 * c + 3]);
}

extern "C" int evad_sample_concat(const float* x, int ldx, int cx, const float* pos, int ld_pos, const int32_t* batch,
                                  int64_t n_cap, const int32_t* n_dev, const float* feat, int feat_b, int feat_h,
                                  int feat_w, int feat_c, float width, float height, int append_pos, float* out,
                                  int ldo, void* stream) {
    if (!pos || !out || n_cap < 0 || cx < 0 || (cx > 0 && !x) || (feat && !batch)) return EVAD_ERR_ARG;
    const int cs = feat ? feat_c : 0;
    if (ldo < cx + cs + (append_pos ? 2 : 0)) return EVAD_ERR_ARG;
    if (n_cap == 0) return EVAD_OK;
    PoolSample S{feat, feat_b, feat_h, feat_w, feat_c, width, height};
    if (feat && cx == 1 && feat_c == 16 && ldo == 20 && (((uintptr_t)feat | (uintptr_t)out) & 15) == 0) {
        sample_concat_row_kernel<1, 16, 20, false><<<evad_cdiv(n_cap, 256), 256, 0, (cudaStream_t)stream>>>(
            x, ldx, pos, ld_pos, batch, n_cap, n_dev, S, append_pos, out, nullptr, nullptr);
        EVAD_CHECK_LAUNCH();
        return EVAD_OK;
    }
    if (feat && feat_c >= 4 && feat_c % 4 == 0 && ((uintptr_t)feat & 15) == 0) {
        const int64_t threads = n_cap * (feat_c / 4);
        sample_concat_quad_kernel<<<evad_cdiv(threads, 256), 256, 0, (cudaStream_t)stream>>>(
            x, ldx, cx, pos, ld_pos, batch, n_cap, n_dev, S, append_pos, out, ldo);
        EVAD_CHECK_LAUNCH();
        return EVAD_OK;
    }
    sample_concat_kernel<<<evad_cdiv(n_cap * ldo, 256), 256, 0, (cudaStream_t)stream>>>(x, ldx, cx, pos, ld_pos, batch,
                                                                                      n_cap, n_dev, S, append_pos, out,
                                                                                      ldo);
    EVAD_CHECK_LAUNCH();
    return EVAD_OK;
}

// the layer-1 row shape of evad_sample_concat (one polarity channel + 16 sampled channels + position, rows of 20) with the
// row written as h2 words for evad_spline_conv_l1_h16; EVAD_ERR_UNSUPPORTED for any other shape (caller: evad_sample_concat
// + evad_split_h2).  visit (optional, a permutation of 0..n-1): the order in which threads take the nodes, e.g. the graph
// build's cell_perm so that neighbouring threads sample neighbouring map cells; the output is the same either way
extern "C" int evad_sample_concat_h2(const float* x, int ldx, int cx, const float* pos, int ld_pos, const int32_t* batch,
                                     int64_t n_cap, const int32_t* n_dev, const float* feat, int feat_b, int feat_h,
                                     int feat_w, int feat_c, float width, float height, int append_pos, uint32_t* out,
                                     int ldo, const int32_t* visit, uint32_t* status, void* stream) {
    if (!pos || !out || n_cap < 0 || cx < 0 || (cx > 0 && !x) || (feat && !batch)) return EVAD_ERR_ARG;
    if (!(feat && cx == 1 && feat_c == 16 && ldo == 20 && (((uintptr_t)feat | (uintptr_t)out) & 15) == 0))
        return EVAD_ERR_UNSUPPORTED;
    if (n_cap == 0) return EVAD_OK;
    PoolSample S{feat, feat_b, feat_h, feat_w, feat_c, width, height};
    sample_concat_row_kernel<1, 16, 20, true><<<evad_cdiv(n_cap, 256), 256, 0, (cudaStream_t)stream>>>(
        x, ldx, pos, ld_pos, batch, n_cap, n_dev, S, append_pos, reinterpret_cast<float*>(out), status, visit);
    EVAD_CHECK_LAUNCH();
    return EVAD_OK;
}

// ------------------------------------------------------------------------------------------------ format_events
// utils/buffers.py:33-44 (format_data) on the raw Data layout of data/utils.py:6-20: pos int16 [N,2], t int32 [N],
// x int8 [N,1] (polarity), batch int64 [N]  ->  x float32, pos = cat([pos, t]) / [W, H, T] (integer true division = fp32
// divide of the converted operands), batch int32.  One launch instead of six torch kernels.
__global__ void __launch_bounds__(256) format_events_kernel(const int16_t* __restrict__ pos_xy, const int32_t* __restrict__ t,
                                                            const int8_t* __restrict__ pol, const int64_t* __restrict__ batch64,
                                                            int64_t n, float w, float h, float tw, float* __restrict__ x,
                                                            float* __restrict__ pos, int32_t* __restrict__ batch) {
    const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const short2 xy = reinterpret_cast<const short2*>(pos_xy)[i];
    pos[3 * i] = __fdiv_rn((float)xy.x, w);
    pos[3 * i + 1] = __fdiv_rn((float)xy.y, h);
    pos[3 * i + 2] = __fdiv_rn((float)t[i], tw);
    x[i] = (float)pol[i];
    batch[i] = batch64 ? (int32_t)batch64[i] : 0;
}

extern "C" int evad_format_events(const int16_t* pos_xy, const int32_t* t, const int8_t* polarity, const int64_t* batch64,
                                  int64_t n, int width, int height, int time_window, float* x, float* pos, int32_t* batch,
                                  void* stream) {
    if (!pos_xy || !t || !polarity || !x || !pos || !batch || n < 0 || width <= 0 || height <= 0 || time_window <= 0)
        return EVAD_ERR_ARG;
    if (((uintptr_t)pos_xy & 3) != 0) return EVAD_ERR_ARG;
    if (n == 0) return EVAD_OK;
    format_events_kernel<<<evad_cdiv(n, 256), 256, 0, (cudaStream_t)stream>>>(pos_xy, t, polarity, batch64, n, (float)width,
                                                                          (float)height, (float)time_window, x, pos, batch);
    EVAD_CHECK_LAUNCH();
    return EVAD_OK;
}

// ------------------------------------------------------------------------------------------------ denormalize_pos
// ev_tgn.py:11-16: (pos * [W, H, T] + 1e-3).int() -- fp32 multiply, fp32 add, truncation.
__global__ void __launch_bounds__(256) denormalize_kernel(const float* __restrict__ pos, int ld_pos, int64_t n, float w,
                                                          float h, float t, int32_t* __restrict__ out) {
    const int64_t k = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (k >= 3 * n) return;
    const int64_t i = k / 3;
    const int d = (int)(k - 3 * i);
    const float s = d == 0 ? w : (d == 1 ? h : t);
    out[k] = (int32_t)__fadd_rn(__fmul_rn(s, pos[i * ld_pos + d]), 1e-3f);
}

extern "C" int evad_denormalize_pos(const float* pos, int ld_pos, int64_t n, int width, int height, int time_window,
                                    int32_t* pos_int, void* stream) {
    if (!pos || !pos_int || n < 0 || ld_pos < 3) return EVAD_ERR_ARG;
    if (n == 0) return EVAD_OK;
    denormalize_kernel<<<evad_cdiv(3 * n, 256), 256, 0, (cudaStream_t)stream>>>(pos, ld_pos, n, (float)width, (float)height,
                                                                              (float)time_window, pos_int);
    EVAD_CHECK_LAUNCH();
    return EVAD_OK;
}
