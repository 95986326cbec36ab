// eventad_b200/csrc/spline_small.cu -- fused LUT SplineConv for the full-resolution layer (cin <= 32, cout == 16).
//
// Reference path replaced per call (paths relative to /root/reference/src/dagr/model):
//   layers/spline_conv.py:39-47  message_lut: gather x_j, gather lut[ix,iy] ([E,cin,cout] materialised), einsum
//   PyG propagate sum aggregation (atomics on the GPU), layers/spline_conv.py:72-76 root weight
//   layers/conv.py:23-28,47-56   eval BatchNorm, activation, skip Linear + BatchNorm
// Here: the (2r+1)^2 LUT (+ root and skip matrices) sits in shared memory (103 KB at cin 19->20), one warp owns one
// destination node (CSR row), the two half-warps take alternate edges, lane&15 is the output channel; no atomics,
// no [E,cin,cout] tensor, BN/ReLU/skip fused into the epilogue.
#include "spline_common.cuh"

#define SS_COUT 16
#define SS_THREADS 512
#define SS_WARPS (SS_THREADS / 32)

__global__ void lut_build_kernel(const float* __restrict__ weight, const float* __restrict__ root_w,
                                 const float* __restrict__ skip_w, int cin, int cin_pad, int cout, EvadSplineMode M,
                                 float* __restrict__ lut) {
    const int nlut = M.nx * M.ny;
    const int64_t total = (int64_t)(nlut + 2) * cin_pad * cout;
    int64_t t = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (t >= total) return;
    const int o = (int)(t % cout);
    const int i = (int)((t / cout) % cin_pad);
    const int l = (int)(t / ((int64_t)cout * cin_pad));
    float v = 0.f;
    if (i < cin) {
        if (l < nlut) {
            const int ix = l / M.ny, iy = l - ix * M.ny;
            // (bil_w[..., None, None] * weight[indices]).sum(1)   (spline_conv.py:33)
#pragma unroll
            for (int s = 0; s < 4; s++) {
                const int mx = s & 1, my = s >> 1;
                int kk = M.tab_k[2 * ix + mx] + M.ks * M.tab_k[2 * (M.nx + iy) + my];
                if (kk < 0) kk += M.ks * M.ks;
                const float b = __fmul_rn(M.tab_w[2 * ix + mx], M.tab_w[2 * (M.nx + iy) + my]);
                v = __fadd_rn(v, __fmul_rn(b, weight[((int64_t)kk * cin + i) * cout + o]));
            }
        } else if (l == nlut) {
            v = root_w ? root_w[(int64_t)o * cin + i] : 0.f;
        } else {
            v = skip_w ? skip_w[(int64_t)o * cin + i] : 0.f;
        }
    }
    lut[t] = v;
}

extern "C" int evad_spline_lut_build(const float* weight, const float* root_w, const float* skip_w, int cin,
                                     int cin_pad, int cout, const evad_spline_mode_t* mode_h, float* lut,
                                     void* stream) {
    if (!weight || !mode_h || !lut || cin < 1 || cin_pad < cin || cout < 1) return EVAD_ERR_ARG;
    if (!mode_h->tab_k || !mode_h->tab_w) return EVAD_ERR_ARG;
    const int64_t total = (int64_t)(mode_h->nx * mode_h->ny + 2) * cin_pad * cout;
    lut_build_kernel<<<evad_cdiv(total, 256), 256, 0, (cudaStream_t)stream>>>(weight, root_w, skip_w, cin, cin_pad,
                                                                             cout, *mode_h, lut);
    EVAD_CHECK_LAUNCH();
    return EVAD_OK;
}

__device__ __forceinline__ float epi_apply(const evad_epilogue_t& E, float acc, int64_t row, int o) {
    float v = acc;
    if (E.scale) v = __fmul_rn(v, E.scale[o]);
    if (E.shift) v = __fadd_rn(v, E.shift[o]);
    if (E.addend) v = __fadd_rn(v, E.addend[row * E.ld_add + o]);
    if (E.relu) v = fmaxf(v, 0.f);
    return v;
}

// Thread mapping: one warp per destination node; lane = (edge slot es = lane>>2, output quad q = lane&3).  Eight in-edges
// are processed at once: the 4 lanes of a slot load the source row (5 x 128-bit, same address -> one transaction),
// then per input channel one 128-bit LUT read feeds 4 FFMAs.  Odd slots walk channel pairs swapped, so the two
// 64-byte LUT segments a quarter-warp reads in one LDS.128 wavefront always sit in different bank halves (the class
// stride Cin_pad*16 floats is a multiple of 32 banks).  The root weight (and the skip Linear) ride along as one more
// pseudo-edge whose source is the node itself.
template <int CIN_PAD>
__global__ void __launch_bounds__(SS_THREADS, 2) spline_small_kernel(
    const int32_t* __restrict__ rowptr, const int32_t* __restrict__ src, int64_t n, const float* __restrict__ x,
    int ldx, EvadPseudo P, EvadSplineMode M, const float* __restrict__ lut, evad_epilogue_t E, float* __restrict__ out,
    int ldo, int has_skip, evad_epilogue_t ES, float* __restrict__ skip_out, int ld_skip, uint32_t* __restrict__ status) {
    extern __shared__ __align__(16) float lut_s[];
    const int nlut = M.nx * M.ny;
    {   // stage the LUT (+root, +skip) in shared memory, 128-bit copies
        const int total4 = (nlut + 2) * CIN_PAD * SS_COUT / 4;
        const float4* g = reinterpret_cast<const float4*>(lut);
        float4* s = reinterpret_cast<float4*>(lut_s);
        for (int t = threadIdx.x; t < total4; t += SS_THREADS) s[t] = g[t];
    }
    __syncthreads();
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int es = lane >> 2, q = lane & 3, odd = es & 1;
    const int64_t stride = (int64_t)gridDim.x * SS_WARPS;
    for (int64_t i = (int64_t)blockIdx.x * SS_WARPS + warp; i < n; i += stride) {
        const int r0 = rowptr[i];
        const int deg = rowptr[i + 1] - r0;
        float acc0 = 0.f, acc1 = 0.f, acc2 = 0.f, acc3 = 0.f;       // conv (+root) partial sums of this edge slot
        float sk0 = 0.f, sk1 = 0.f, sk2 = 0.f, sk3 = 0.f;           // skip Linear (only the slot that owns the root)
        for (int e0 = 0; e0 <= deg; e0 += 8) {                      // pseudo-edge `deg` = root / skip on x_i
            const int e = e0 + es;
            if (e > deg) continue;
            const bool is_root = (e == deg);
            int j, cls;
            if (is_root) {
                j = (int)i;
                cls = nlut;
            } else {
                j = src[r0 + e];
                float ax, ay;
                evad_pseudo(P, r0 + e, j, (int)i, &ax, &ay);
                int ix, iy;
                evad_lut_index(M, ax, ay, &ix, &iy, status);
                cls = ix * M.ny + iy;
            }
            float xr[CIN_PAD];
            const float4* xp = reinterpret_cast<const float4*>(x + (int64_t)j * ldx);
#pragma unroll
            for (int k = 0; k < CIN_PAD / 4; k++) {
                const float4 v = __ldg(xp + k);
                // odd slots keep each channel pair swapped in registers
                xr[4 * k] = odd ? v.y : v.x; xr[4 * k + 1] = odd ? v.x : v.y;
                xr[4 * k + 2] = odd ? v.w : v.z; xr[4 * k + 3] = odd ? v.z : v.w;
            }
            const float* w = lut_s + (cls * CIN_PAD) * SS_COUT + 4 * q;
#pragma unroll
            for (int c = 0; c < CIN_PAD; c++) {
                const int ci = c ^ odd;                              // channel whose value sits in xr[c]
                const float4 w4 = *reinterpret_cast<const float4*>(w + ci * SS_COUT);
                acc0 = fmaf(w4.x, xr[c], acc0); acc1 = fmaf(w4.y, xr[c], acc1);
                acc2 = fmaf(w4.z, xr[c], acc2); acc3 = fmaf(w4.w, xr[c], acc3);
            }
            if (is_root && has_skip) {
                const float* wk = w + CIN_PAD * SS_COUT;
#pragma unroll
                for (int c = 0; c < CIN_PAD; c++) {
                    const int ci = c ^ odd;
                    const float4 w4 = *reinterpret_cast<const float4*>(wk + ci * SS_COUT);
                    sk0 = fmaf(w4.x, xr[c], sk0); sk1 = fmaf(w4.y, xr[c], sk1);
                    sk2 = fmaf(w4.z, xr[c], sk2); sk3 = fmaf(w4.w, xr[c], sk3);
                }
            }
        }
#pragma unroll
        for (int d = 4; d < 32; d <<= 1) {                          // reduce over the 8 edge slots
            acc0 += __shfl_xor_sync(0xffffffffu, acc0, d); acc1 += __shfl_xor_sync(0xffffffffu, acc1, d);
            acc2 += __shfl_xor_sync(0xffffffffu, acc2, d); acc3 += __shfl_xor_sync(0xffffffffu, acc3, d);
            sk0 += __shfl_xor_sync(0xffffffffu, sk0, d); sk1 += __shfl_xor_sync(0xffffffffu, sk1, d);
            sk2 += __shfl_xor_sync(0xffffffffu, sk2, d); sk3 += __shfl_xor_sync(0xffffffffu, sk3, d);
        }
        if (es == 0) {
            float4 o;
            o.x = epi_apply(E, acc0, i, 4 * q); o.y = epi_apply(E, acc1, i, 4 * q + 1);
            o.z = epi_apply(E, acc2, i, 4 * q + 2); o.w = epi_apply(E, acc3, i, 4 * q + 3);
            *reinterpret_cast<float4*>(out + i * ldo + 4 * q) = o;
        } else if (es == 1 && has_skip) {
            float4 o;
            o.x = epi_apply(ES, sk0, i, 4 * q); o.y = epi_apply(ES, sk1, i, 4 * q + 1);
            o.z = epi_apply(ES, sk2, i, 4 * q + 2); o.w = epi_apply(ES, sk3, i, 4 * q + 3);
            *reinterpret_cast<float4*>(skip_out + i * ld_skip + 4 * q) = o;
        }
    }
}

template <int CIN_PAD>
static int launch_small(const int32_t* rowptr, const int32_t* src, int64_t n, const float* x, int ldx,
                        const evad_pseudo_t& P, const evad_spline_mode_t& M, const float* lut, const evad_epilogue_t& E,
                        float* out, int ldo, int has_skip, const evad_epilogue_t& ES, float* skip_out, int ld_skip,
                        uint32_t* status, cudaStream_t st) {
    const size_t smem = (size_t)(M.nx * M.ny + 2) * CIN_PAD * SS_COUT * sizeof(float);
    if (smem > 227 * 1024) return EVAD_ERR_UNSUPPORTED;
    static bool attr_done = false;                          // per instantiation
    if (!attr_done) {
        if (cudaFuncSetAttribute(spline_small_kernel<CIN_PAD>, cudaFuncAttributeMaxDynamicSharedMemorySize, 227 * 1024) !=
            cudaSuccess)
            return EVAD_ERR_LAUNCH;
        attr_done = true;
    }
    const int ctas_per_sm = smem > 110 * 1024 ? 1 : 2;
    int64_t want = (n + SS_WARPS - 1) / SS_WARPS;
    int grid = (int)(want < (int64_t)148 * ctas_per_sm ? want : (int64_t)148 * ctas_per_sm);
    if (grid < 1) grid = 1;
    spline_small_kernel<CIN_PAD><<<grid, SS_THREADS, smem, st>>>(rowptr, src, n, x, ldx, P, M, lut, E, out, ldo, has_skip,
                                                               ES, skip_out, ld_skip, status);
    EVAD_CHECK_LAUNCH();
    return EVAD_OK;
}

extern "C" int evad_spline_conv_small(const int32_t* rowptr, const int32_t* src, int64_t n, const float* x, int ldx,
                                      int cin, int cin_pad, int cout, const evad_pseudo_t* pseudo_h,
                                      const evad_spline_mode_t* mode_h, const float* lut, const evad_epilogue_t* epi_h,
                                      float* out, int ldo, const evad_epilogue_t* skip_epi_h, float* skip_out,
                                      int ld_skip, uint32_t* status, void* stream) {
    if (!rowptr || !src || !x || !pseudo_h || !mode_h || !lut || !epi_h || !out || !status) return EVAD_ERR_ARG;
    if (cout != SS_COUT || cin_pad < cin || cin_pad % 4 != 0 || cin_pad > 32 || ldx % 4 != 0 || ldx < cin_pad)
        return EVAD_ERR_UNSUPPORTED;
    if (((uintptr_t)x & 15) || ((uintptr_t)lut & 15) || ((uintptr_t)out & 15) || ldo % 4 != 0) return EVAD_ERR_ARG;
    if (skip_out && (((uintptr_t)skip_out & 15) || ld_skip % 4 != 0)) return EVAD_ERR_ARG;
    if (!mode_h->use_lut) return EVAD_ERR_UNSUPPORTED;
    if (n <= 0) return EVAD_OK;
    const int has_skip = skip_epi_h != nullptr && skip_out != nullptr;
    evad_epilogue_t es = has_skip ? *skip_epi_h : *epi_h;
    cudaStream_t st = (cudaStream_t)stream;
#define EVAD_SS_CASE(C)                                                                                              \
    case C:                                                                                                          \
        return launch_small<C>(rowptr, src, n, x, ldx, *pseudo_h, *mode_h, lut, *epi_h, out, ldo, has_skip, es,     \
                               skip_out, ld_skip, status, st);
    switch (cin_pad) {
        EVAD_SS_CASE(4) EVAD_SS_CASE(8) EVAD_SS_CASE(12) EVAD_SS_CASE(16) EVAD_SS_CASE(20) EVAD_SS_CASE(24)
        EVAD_SS_CASE(28) EVAD_SS_CASE(32)
    default:
        return EVAD_ERR_UNSUPPORTED;
    }
#undef EVAD_SS_CASE
}
