// eventad_b200/csrc/spline_common.cuh -- per-edge pseudo-coordinates and degree-1 open B-spline taps.
//
// Reference arithmetic reproduced (paths relative to /root/reference/src/dagr/model):
//   PyG T.Cartesian(norm=True,cat=False,max_value=m): attr = (pos[src] - pos[dst]) / (2m) + 0.5   (layers/components.py:25-35)
//   clamp(attr, 0, 1) for the first layer only                                                    (networks/net.py:120-121)
//   message_lut: ix = (long)(attr_x * R00 + R02 + 1e-3), iy likewise                              (layers/spline_conv.py:39-42)
//   torch_spline_conv.spline_basis (degree 1, open spline) when no LUT is cached                   (layers/spline_conv.py:6)
// All fp32 steps use explicit round-to-nearest intrinsics so that nvcc cannot contract them into FMAs.
#pragma once
#include "common.cuh"

// the two parameter blocks are part of the C ABI: see evad_pseudo_t / evad_spline_mode_t in include/eventad_b200.h
typedef evad_pseudo_t EvadPseudo;
typedef evad_spline_mode_t EvadSplineMode;

__device__ __forceinline__ void evad_pseudo(const EvadPseudo& P, int e, int j, int i, float* ax, float* ay) {
    float a0, a1;
    if (P.edge_attr) {
        const int64_t id = P.eid ? P.eid[e] : e;
        a0 = P.edge_attr[id * P.ld_attr];
        a1 = P.edge_attr[id * P.ld_attr + 1];
    } else {
        const float dx = __fsub_rn(P.pos[(int64_t)j * P.ld_pos], P.pos[(int64_t)i * P.ld_pos]);
        const float dy = __fsub_rn(P.pos[(int64_t)j * P.ld_pos + 1], P.pos[(int64_t)i * P.ld_pos + 1]);
        // torch on CUDA divides a tensor by a Python scalar as a multiplication with the fp32 reciprocal
        // (BinaryDivTrueKernel.cu, is_cpu_scalar branch): T.Cartesian's `cart / (2 * max_value)` as the reference runs it
        const float inv = __fdiv_rn(1.0f, P.two_m);
        a0 = __fadd_rn(__fmul_rn(dx, inv), 0.5f);
        a1 = __fadd_rn(__fmul_rn(dy, inv), 0.5f);
        if (P.clamp01) {
            a0 = fminf(fmaxf(a0, 0.f), 1.f);
            a1 = fminf(fmaxf(a1, 0.f), 1.f);
        }
    }
    *ax = a0;
    *ay = a1;
}

// LUT lattice coordinates of an edge; returns false (and raises the status bit) when outside the table
__device__ __forceinline__ bool evad_lut_index(const EvadSplineMode& M, float ax, float ay, int* ix, int* iy,
                                               uint32_t* status) {
    const long long lx = (long long)__fadd_rn(__fadd_rn(__fmul_rn(ax, M.r00), M.r02), 1e-3f);
    const long long ly = (long long)__fadd_rn(__fadd_rn(__fmul_rn(ay, M.r11), M.r12), 1e-3f);
    bool ok = lx >= 0 && ly >= 0 && lx < M.nx && ly < M.ny;
    if (!ok) atomicOr(status, EVAD_STATUS_LUT_OOB);
    *ix = (int)min(max(lx, 0ll), (long long)M.nx - 1);
    *iy = (int)min(max(ly, 0ll), (long long)M.ny - 1);
    return ok;
}

// one dimension of spline_basis (degree 1, open): v = a*(ks-1); k0 = ((int64)v) % ks; k1 = ((int64)v + 1) % ks;
// frac = v - floor(v); w0 = 1 - frac - 0 + 2*frac*0 ; w1 = 1 - frac - 1 + 2*frac*1   (library expression, fp32)
__device__ __forceinline__ void evad_basis_1d(float a, int ks, int* k0, int* k1, float* w0, float* w1) {
    const float v = __fmul_rn(a, (float)(ks - 1));
    const long long vi = (long long)v;                      // truncation toward zero
    *k0 = (int)(vi % ks);                                   // C remainder (sign of the dividend)
    *k1 = (int)((vi + 1) % ks);
    const float f = __fsub_rn(v, floorf(v));
    *w0 = __fsub_rn(1.f, f);
    *w1 = __fadd_rn(__fsub_rn(__fsub_rn(1.f, f), 1.f), __fmul_rn(__fmul_rn(2.f, f), 1.f));
}

// the four (weight-matrix index, basis value) taps of an edge; s = (k_mod_x) + 2*(k_mod_y), dim 0 fastest
__device__ __forceinline__ void evad_taps(const EvadSplineMode& M, float ax, float ay, int* k, float* w,
                                          uint32_t* status) {
    int kx[2], ky[2];
    float wx[2], wy[2];
    if (M.use_lut) {
        int ix, iy;
        evad_lut_index(M, ax, ay, &ix, &iy, status);
        kx[0] = M.tab_k[2 * ix]; kx[1] = M.tab_k[2 * ix + 1];
        wx[0] = M.tab_w[2 * ix]; wx[1] = M.tab_w[2 * ix + 1];
        ky[0] = M.tab_k[2 * (M.nx + iy)]; ky[1] = M.tab_k[2 * (M.nx + iy) + 1];
        wy[0] = M.tab_w[2 * (M.nx + iy)]; wy[1] = M.tab_w[2 * (M.nx + iy) + 1];
    } else {
        evad_basis_1d(ax, M.ks, &kx[0], &kx[1], &wx[0], &wx[1]);
        evad_basis_1d(ay, M.ks, &ky[0], &ky[1], &wy[0], &wy[1]);
    }
#pragma unroll
    for (int s = 0; s < 4; s++) {
        const int mx = s & 1, my = s >> 1;
        int kk = kx[mx] + M.ks * ky[my];
        if (kk < 0) kk += M.ks * M.ks;                      // negative remainders cannot index; keep in range
        k[s] = kk;
        w[s] = __fmul_rn(wx[mx], wy[my]);
    }
}
