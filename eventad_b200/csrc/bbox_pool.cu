// eventad_b200/csrc/bbox_pool.cu -- per-bounding-box pooling of the DAGR output graph (EventAD head input).
//
// Reference path replaced (paths relative to /root/reference): models/EventAD.py:382-397 (python loop over the boxes of
// every sample, current frame `bbox` and previous frame `bbox0`) calling _process_bbox (:406-498) once per box, each
// call a dozen tiny kernels and four .item() host synchronisations.  Here: one launch, one warp per box.
//   * track id = (int)box[5]; ids outside 1..max_boxes are skipped (:427-428)
//   * box corners and centre are normalised in DOUBLE (python floats from .item(), :421,431-442) and rounded to fp32
//     where torch compares them with / subtracts them from the fp32 node positions (:456-457,463,473)
//   * nodes of the box's sample inside the box (inclusive bounds): inverse-distance weights 1 / (d + 1e-6), normalised
//     by their sum, weighted feature sum (:459-470)
//   * no node inside: mean feature of the (up to) 5 nodes of the sample closest to the centre (:472-482); ties in the
//     distance take the smaller node index (torch.topk leaves the order of ties unspecified)
//   * sample without nodes: the global feature = mean over no nodes = zeros (:366-373,447-449)
//   * the reference writes box after box into batch_features[b, frame, track]: of several boxes with the same
//     (sample, frame, track) the LAST one wins; a warp whose key re-appears later in the list does not write.
#include "common.cuh"

#define BP_WARPS 4

__global__ void __launch_bounds__(BP_WARPS * 32) bbox_pool_kernel(
    const float* __restrict__ feat, int ldf, int C, const float* __restrict__ pos, int ldp,
    const int32_t* __restrict__ node_batch, int64_t n_cap, const int32_t* __restrict__ n_dev,
    const float* __restrict__ bbox, int ldb, const int64_t* __restrict__ bbox_batch, int m, int frame,
    const float* __restrict__ width, const float* __restrict__ height, int n_frames, int max_boxes,
    float* __restrict__ out /* [B][n_frames][max_boxes + 1][C] */) {
    const int lane = threadIdx.x & 31;
    const int idx = blockIdx.x * BP_WARPS + (threadIdx.x >> 5);
    if (idx >= m) return;
    const float* bx = bbox + (int64_t)idx * ldb;
    const int track = ldb > 5 ? (int)bx[5] : 1;
    if (track <= 0 || track > max_boxes) return;
    const int b = (int)bbox_batch[idx];
    // last box with this (sample, track) wins
    bool later = false;
    for (int j = idx + 1 + lane; j < m; j += 32)
        if ((int)bbox_batch[j] == b && (ldb > 5 ? (int)bbox[(int64_t)j * ldb + 5] : 1) == track) later = true;
    if (__any_sync(0xffffffffu, later)) return;
    const int64_t n = n_dev ? min((int64_t)*n_dev, n_cap) : n_cap;
    const double W = (double)width[b], H = (double)height[b];
    const double x = (double)bx[0], y = (double)bx[1], w = (double)bx[2], h = (double)bx[3];
    const double x1 = x / W, y1 = y / H, x2 = (x + w) / W, y2 = (y + h) / H;
    const float fx1 = (float)x1, fy1 = (float)y1, fx2 = (float)x2, fy2 = (float)y2;
    const float cx = (float)((x1 + x2) / 2.0), cy = (float)((y1 + y2) / 2.0);

    // sweep 1: nodes of the sample, nodes in the box, sum of the inverse-distance weights
    int nb = 0, nin = 0;
    float wsum = 0.f;
    for (int64_t i0 = 0; i0 < n; i0 += 32) {
        const int64_t i = i0 + lane;
        bool mine = false, in = false;
        float wt = 0.f;
        if (i < n && node_batch[i] == b) {
            mine = true;
            const float px = pos[i * ldp], py = pos[i * ldp + 1];
            in = px >= fx1 && px <= fx2 && py >= fy1 && py <= fy2;
            if (in) {
                const float dx = __fsub_rn(px, cx), dy = __fsub_rn(py, cy);
                const float d = __fsqrt_rn(__fadd_rn(__fmul_rn(dx, dx), __fmul_rn(dy, dy)));
                wt = __fdiv_rn(1.0f, __fadd_rn(d, 1e-6f));
            }
        }
        nb += __popc(__ballot_sync(0xffffffffu, mine));
        nin += __popc(__ballot_sync(0xffffffffu, in));
        wsum += wt;
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) wsum += __shfl_xor_sync(0xffffffffu, wsum, o);

    float* dst = out + (((int64_t)b * n_frames + frame) * (max_boxes + 1) + track) * C;
    if (nb == 0) {                                            // global feature of a sample without nodes: zeros
        for (int c = lane; c < C; c += 32) dst[c] = 0.f;
        return;
    }
    float acc[4] = {0.f, 0.f, 0.f, 0.f};                      // channels lane, lane + 32, ... (C <= 128)
    if (nin > 0) {
        for (int64_t i0 = 0; i0 < n; i0 += 32) {
            const int64_t i = i0 + lane;
            float wt = 0.f;
            bool in = false;
            if (i < n && node_batch[i] == b) {
                const float px = pos[i * ldp], py = pos[i * ldp + 1];
                in = px >= fx1 && px <= fx2 && py >= fy1 && py <= fy2;
                if (in) {
                    const float dx = __fsub_rn(px, cx), dy = __fsub_rn(py, cy);
                    const float d = __fsqrt_rn(__fadd_rn(__fmul_rn(dx, dx), __fmul_rn(dy, dy)));
                    wt = __fdiv_rn(__fdiv_rn(1.0f, __fadd_rn(d, 1e-6f)), wsum);   // weights / weights.sum()
                }
            }
            unsigned mask = __ballot_sync(0xffffffffu, in);
            while (mask) {
                const int l = __ffs(mask) - 1;
                mask &= mask - 1;
                const float wl = __shfl_sync(0xffffffffu, wt, l);
                const float* f = feat + (i0 + l) * ldf;
#pragma unroll
                for (int u = 0; u < 4; u++)
                    if (lane + 32 * u < C) acc[u] = __fadd_rn(acc[u], __fmul_rn(f[lane + 32 * u], wl));
            }
        }
    } else {
        // the k = min(5, nb) nodes of the sample closest to the centre, in ascending (distance, index) order
        const int k = nb < 5 ? nb : 5;
        float dprev = -1.f;
        int64_t iprev = -1;
        for (int r = 0; r < k; r++) {
            float dbest = 3.0e38f;
            int64_t ibest = -1;
            for (int64_t i = lane; i < n; i += 32) {
                if (node_batch[i] != b) continue;
                const float dx = __fsub_rn(pos[i * ldp], cx), dy = __fsub_rn(pos[i * ldp + 1], cy);
                const float d = __fsqrt_rn(__fadd_rn(__fmul_rn(dx, dx), __fmul_rn(dy, dy)));
                const bool after = d > dprev || (d == dprev && i > iprev);
                if (after && (ibest < 0 || d < dbest || (d == dbest && i < ibest))) { dbest = d; ibest = i; }
            }
#pragma unroll
            for (int o = 16; o > 0; o >>= 1) {
                const float d2 = __shfl_xor_sync(0xffffffffu, dbest, o);
                const int64_t i2 = __shfl_xor_sync(0xffffffffu, ibest, o);
                if (i2 >= 0 && (ibest < 0 || d2 < dbest || (d2 == dbest && i2 < ibest))) { dbest = d2; ibest = i2; }
            }
            dprev = dbest;
            iprev = ibest;
            const float* f = feat + ibest * ldf;
#pragma unroll
            for (int u = 0; u < 4; u++)
                if (lane + 32 * u < C) acc[u] = __fadd_rn(acc[u], f[lane + 32 * u]);
        }
#pragma unroll
        for (int u = 0; u < 4; u++) acc[u] = __fdiv_rn(acc[u], (float)k);
    }
#pragma unroll
    for (int u = 0; u < 4; u++)
        if (lane + 32 * u < C) dst[lane + 32 * u] = acc[u];
}

extern "C" int evad_bbox_pool(const float* feat, int ldf, int channels, const float* pos, int ldp,
                              const int32_t* node_batch, int64_t n_cap, const int32_t* n_dev, const float* bbox, int ldb,
                              const int64_t* bbox_batch, int num_boxes, int frame, const float* width, const float* height,
                              int n_frames, int max_boxes, float* batch_features, void* stream) {
    if (!feat || !pos || !node_batch || !width || !height || !batch_features) return EVAD_ERR_ARG;
    if (channels < 1 || channels > 128 || ldf < channels || ldp < 2 || frame < 0 || frame >= n_frames || max_boxes < 1)
        return EVAD_ERR_ARG;
    if (num_boxes <= 0) return EVAD_OK;
    if (!bbox || !bbox_batch || ldb < 4) return EVAD_ERR_ARG;
    bbox_pool_kernel<<<evad_cdiv(num_boxes, BP_WARPS), BP_WARPS * 32, 0, (cudaStream_t)stream>>>(
        feat, ldf, channels, pos, ldp, node_batch, n_cap, n_dev, bbox, ldb, bbox_batch, num_boxes, frame, width, height,
        n_frames, max_boxes, batch_features);
    EVAD_CHECK_LAUNCH();
    return EVAD_OK;
}
