// eventad_b200/csrc/sample.cuh -- CNN feature-map sampling at graph nodes (reference: sample_features /
// _sample_features, /root/reference/src/dagr/model/networks/net.py:200-228: 5-D F.grid_sample, bilinear,
// align_corners=True, zero padding, batch index used as the depth coordinate).
#pragma once
#include "common.cuh"

struct PoolSample {          // optional fused sampling_skip (net.py:200-228); feat == nullptr disables it
    const float* feat;       // channel-last map [Bimg, h, w, C]
    int Bimg, h, w, C;
    float width, height;     // data.width[0], data.height[0] as floats
};

// F.grid_sample(mode='bilinear', align_corners=True, padding zeros) on the 5-D view of net.py:216-224, one channel
__device__ __forceinline__ float sample_feat(const PoolSample& S, float px, float py, float b, int ch) {
    float gx = __fmul_rn(px, S.width), gy = __fmul_rn(py, S.height);
    gx = __fsub_rn(__fdiv_rn(__fmul_rn(2.f, gx), __fsub_rn(S.width, 1.f)), 1.f);
    gy = __fsub_rn(__fdiv_rn(__fmul_rn(2.f, gy), __fsub_rn(S.height, 1.f)), 1.f);
    const float bs = S.Bimg > 1 ? (float)S.Bimg : 2.f;
    const float gb = __fsub_rn(__fdiv_rn(__fmul_rn(2.f, b), __fsub_rn(bs, 1.f)), 1.f);
    // grid_sampler_unnormalize, align_corners: ((c + 1) / 2) * (size - 1)
    const float ix = __fmul_rn(__fdiv_rn(__fadd_rn(gx, 1.f), 2.f), (float)(S.w - 1));
    const float iy = __fmul_rn(__fdiv_rn(__fadd_rn(gy, 1.f), 2.f), (float)(S.h - 1));
    const float iz = __fmul_rn(__fdiv_rn(__fadd_rn(gb, 1.f), 2.f), (float)(S.Bimg - 1));
    const float x0 = floorf(ix), y0 = floorf(iy), z0 = floorf(iz);
    const float fx1 = __fsub_rn(ix, x0), fy1 = __fsub_rn(iy, y0), fz1 = __fsub_rn(iz, z0);
    const float fx0 = __fsub_rn(__fadd_rn(x0, 1.f), ix), fy0 = __fsub_rn(__fadd_rn(y0, 1.f), iy),
                fz0 = __fsub_rn(__fadd_rn(z0, 1.f), iz);
    float acc = 0.f;
#pragma unroll
    for (int dz = 0; dz < 2; dz++) {
        const int z = (int)z0 + dz;
        if (z < 0 || z >= S.Bimg) continue;
        const float wz = dz ? fz1 : fz0;
#pragma unroll
        for (int dy = 0; dy < 2; dy++) {
            const int y = (int)y0 + dy;
            if (y < 0 || y >= S.h) continue;
            const float wy = dy ? fy1 : fy0;
#pragma unroll
            for (int dx = 0; dx < 2; dx++) {
                const int x = (int)x0 + dx;
                if (x < 0 || x >= S.w) continue;
                const float wx = dx ? fx1 : fx0;
                const float wgt = __fmul_rn(__fmul_rn(wx, wy), wz);
                acc = fmaf(S.feat[(((int64_t)z * S.h + y) * S.w + x) * S.C + ch], wgt, acc);
            }
        }
    }
    return acc;
}


// the same sampling split in two: per node the (flat pixel index, weight) of the up to 8 in-bounds taps
// (index -1 = out of bounds, zero padding), in the order sample_feat accumulates them
__device__ __forceinline__ void sample_taps(const PoolSample& S, float px, float py, float b, int* tap, float* wgt) {
    float gx = __fmul_rn(px, S.width), gy = __fmul_rn(py, S.height);
    gx = __fsub_rn(__fdiv_rn(__fmul_rn(2.f, gx), __fsub_rn(S.width, 1.f)), 1.f);
    gy = __fsub_rn(__fdiv_rn(__fmul_rn(2.f, gy), __fsub_rn(S.height, 1.f)), 1.f);
    const float bs = S.Bimg > 1 ? (float)S.Bimg : 2.f;
    const float gb = __fsub_rn(__fdiv_rn(__fmul_rn(2.f, b), __fsub_rn(bs, 1.f)), 1.f);
    const float ix = __fmul_rn(__fdiv_rn(__fadd_rn(gx, 1.f), 2.f), (float)(S.w - 1));
    const float iy = __fmul_rn(__fdiv_rn(__fadd_rn(gy, 1.f), 2.f), (float)(S.h - 1));
    const float iz = __fmul_rn(__fdiv_rn(__fadd_rn(gb, 1.f), 2.f), (float)(S.Bimg - 1));
    const float x0 = floorf(ix), y0 = floorf(iy), z0 = floorf(iz);
    const float fx1 = __fsub_rn(ix, x0), fy1 = __fsub_rn(iy, y0), fz1 = __fsub_rn(iz, z0);
    const float fx0 = __fsub_rn(__fadd_rn(x0, 1.f), ix), fy0 = __fsub_rn(__fadd_rn(y0, 1.f), iy),
                fz0 = __fsub_rn(__fadd_rn(z0, 1.f), iz);
#pragma unroll
    for (int t = 0; t < 8; t++) {
        const int dz = t >> 2, dy = (t >> 1) & 1, dx = t & 1;
        const int z = (int)z0 + dz, y = (int)y0 + dy, x = (int)x0 + dx;
        const bool ok = z >= 0 && z < S.Bimg && y >= 0 && y < S.h && x >= 0 && x < S.w;
        tap[t] = ok ? (z * S.h + y) * S.w + x : -1;
        wgt[t] = __fmul_rn(__fmul_rn(dx ? fx1 : fx0, dy ? fy1 : fy0), dz ? fz1 : fz0);
    }
}


// sample_taps plus the integer base cell (x0, y0, z0) of the 2x2x2 tap cube: tap t sits at (x0 + (t&1), y0 + ((t>>1)&1),
// z0 + (t>>2)).  Used by the pooling kernel to stage the few map cells a voxel touches in shared memory.
__device__ __forceinline__ void sample_taps_base(const PoolSample& S, float px, float py, float b, int* tap, float* wgt,
                                                 int* bx, int* by, int* bz) {
    float gx = __fmul_rn(px, S.width), gy = __fmul_rn(py, S.height);
    gx = __fsub_rn(__fdiv_rn(__fmul_rn(2.f, gx), __fsub_rn(S.width, 1.f)), 1.f);
    gy = __fsub_rn(__fdiv_rn(__fmul_rn(2.f, gy), __fsub_rn(S.height, 1.f)), 1.f);
    const float bs = S.Bimg > 1 ? (float)S.Bimg : 2.f;
    const float gb = __fsub_rn(__fdiv_rn(__fmul_rn(2.f, b), __fsub_rn(bs, 1.f)), 1.f);
    const float ix = __fmul_rn(__fdiv_rn(__fadd_rn(gx, 1.f), 2.f), (float)(S.w - 1));
    const float iy = __fmul_rn(__fdiv_rn(__fadd_rn(gy, 1.f), 2.f), (float)(S.h - 1));
    const float iz = __fmul_rn(__fdiv_rn(__fadd_rn(gb, 1.f), 2.f), (float)(S.Bimg - 1));
    const float x0 = floorf(ix), y0 = floorf(iy), z0 = floorf(iz);
    const float fx1 = __fsub_rn(ix, x0), fy1 = __fsub_rn(iy, y0), fz1 = __fsub_rn(iz, z0);
    const float fx0 = __fsub_rn(__fadd_rn(x0, 1.f), ix), fy0 = __fsub_rn(__fadd_rn(y0, 1.f), iy),
                fz0 = __fsub_rn(__fadd_rn(z0, 1.f), iz);
    *bx = (int)x0; *by = (int)y0; *bz = (int)z0;
#pragma unroll
    for (int t = 0; t < 8; t++) {
        const int dz = t >> 2, dy = (t >> 1) & 1, dx = t & 1;
        const int z = (int)z0 + dz, y = (int)y0 + dy, x = (int)x0 + dx;
        const bool ok = z >= 0 && z < S.Bimg && y >= 0 && y < S.h && x >= 0 && x < S.w;
        tap[t] = ok ? (z * S.h + y) * S.w + x : -1;
        wgt[t] = __fmul_rn(__fmul_rn(dx ? fx1 : fx0, dy ? fy1 : fy0), dz ? fz1 : fz0);
    }
}
