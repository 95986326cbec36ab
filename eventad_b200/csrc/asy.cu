// eventad_b200/csrc/asy.cu -- masked row-update kernels of the asynchronous (incremental) DAGR layers.
//
// Reference replaced: the `asy_tools` CUDA extension, /root/reference/src/dagr/asynchronous/asy_tools/main.cu
//   masked_isdiff      :14-41  + launcher :97-125   (called from asynchronous/base/utils.py:42-43)
//   masked_inplace_BN  :43-67  + launcher :69-95    (asynchronous/batch_norm.py:40)
//   masked_lin(_no_bias) :128-188 + launchers :190-237 (asynchronous/conv.py:236-238, linear.py:46-48)
// Same results bit for bit (tests/test_asy_gpu.py runs the unmodified main.cu next to these on the B200): the serial
// read-modify-write accumulation order over Cin and the fused multiply-adds nvcc contracts in the reference are kept
// (explicit fmaf here so that the result does not depend on compiler flags).
// B200 shape of the work: the reference gives every (row, cout) its own thread, which walks Cin with a global
// read-modify-write per step and reads the weight row uncoalesced.  Here a warp owns a changed row: the row of x is read
// once with coalesced loads and broadcast by shuffles, each lane accumulates its output channels in registers (same
// ascending-Cin order, same fmaf), the weight is read through the read-only path (it is shared by all rows and stays in
// L1), and the output row is written once.
#include "common.cuh"

__global__ void __launch_bounds__(256) asy_masked_lin_kernel(const int64_t* __restrict__ indices, const float* __restrict__ x_in,
                                                             float* __restrict__ x_out, const float* __restrict__ weight,
                                                             const float* __restrict__ bias, int64_t K, int Cin, int Cout,
                                                             int add) {
    const int lane = threadIdx.x & 31;
    const int64_t row = (int64_t)blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
    if (row >= K) return;
    const int64_t i = indices[row];
    const float* xi = x_in + i * Cin;
    float* xo = x_out + i * Cout;
    for (int c0 = 0; c0 < Cout; c0 += 32) {
        const int co = c0 + lane;
        float acc = (add && co < Cout) ? xo[co] : 0.f;
        for (int k0 = 0; k0 < Cin; k0 += 32) {
            const float xv = (k0 + lane < Cin) ? xi[k0 + lane] : 0.f;
            const int kn = min(32, Cin - k0);
            for (int k = 0; k < kn; k++) {
                const float xk = __shfl_sync(0xffffffffu, xv, k);
                if (co < Cout) acc = fmaf(xk, __ldg(weight + (int64_t)co * Cin + k0 + k), acc);
            }
        }
        if (co < Cout) {
            if (bias) acc = __fadd_rn(acc, bias[co]);
            xo[co] = acc;
        }
    }
}

__global__ void __launch_bounds__(256) asy_masked_bn_kernel(const int64_t* __restrict__ indices, const float* __restrict__ x,
                                                            float* __restrict__ x_out, const float* __restrict__ mean,
                                                            const float* __restrict__ var, const float* __restrict__ weight,
                                                            const float* __restrict__ bias, int64_t K, int C, float eps) {
    const int64_t t = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (t >= K * C) return;
    const int64_t i = t / C;
    const int c = (int)(t - i * C);
    const int64_t j = (int64_t)C * indices[i] + c;
    // main.cu:66: (x - mean) / sqrt(var + eps) * weight + bias, the last two steps one FMA under nvcc's default -fmad=true
    x_out[j] = fmaf(__fdiv_rn(__fsub_rn(x[j], mean[c]), __fsqrt_rn(__fadd_rn(var[c], eps))), weight[c], bias[c]);
}

// one warp per candidate row: differs = any channel with |a - b| > atol + rtol * b (main.cu:36)
__global__ void __launch_bounds__(256) asy_masked_isdiff_kernel(int64_t* __restrict__ indices, const float* __restrict__ x_a,
                                                                const float* __restrict__ x_b, int64_t K, int C, float atol,
                                                                float rtol) {
    const int lane = threadIdx.x & 31;
    const int64_t row = (int64_t)blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
    if (row >= K) return;
    const int64_t i = indices[row];
    bool diff = false;
    for (int c = lane; c < C; c += 32) {
        const float a = x_a[i * C + c], b = x_b[i * C + c];
        diff |= fabsf(__fsub_rn(a, b)) > fmaf(rtol, b, atol);
    }
    diff = __any_sync(0xffffffffu, diff);
    if (lane == 0 && !diff) indices[row] = -1;
}

extern "C" int evad_asy_masked_lin(const int64_t* indices, int64_t k, const float* x_in, float* x_out, const float* weight,
                                   const float* bias, int cin, int cout, int add, void* stream) {
    if (k <= 0) return EVAD_OK;
    if (!indices || !x_in || !x_out || !weight || cin < 1 || cout < 1) return EVAD_ERR_ARG;
    asy_masked_lin_kernel<<<evad_cdiv(k, 8), 256, 0, (cudaStream_t)stream>>>(indices, x_in, x_out, weight, bias, k, cin, cout, add);
    EVAD_CHECK_LAUNCH();
    return EVAD_OK;
}

extern "C" int evad_asy_masked_bn(const int64_t* indices, int64_t k, const float* x, float* x_out, const float* running_mean,
                                  const float* running_var, const float* weight, const float* bias, int channels, float eps,
                                  void* stream) {
    if (k <= 0) return EVAD_OK;
    if (!indices || !x || !x_out || !running_mean || !running_var || !weight || !bias || channels < 1) return EVAD_ERR_ARG;
    asy_masked_bn_kernel<<<evad_cdiv(k * channels, 256), 256, 0, (cudaStream_t)stream>>>(indices, x, x_out, running_mean,
                                                                                          running_var, weight, bias, k,
                                                                                          channels, eps);
    EVAD_CHECK_LAUNCH();
    return EVAD_OK;
}

extern "C" int evad_asy_masked_isdiff(int64_t* indices, int64_t k, const float* x_a, const float* x_b, int channels, float atol,
                                      float rtol, void* stream) {
    if (k <= 0) return EVAD_OK;
    if (!indices || !x_a || !x_b || channels < 1) return EVAD_ERR_ARG;
    asy_masked_isdiff_kernel<<<evad_cdiv(k, 8), 256, 0, (cudaStream_t)stream>>>(indices, x_a, x_b, k, channels, atol, rtol);
    EVAD_CHECK_LAUNCH();
    return EVAD_OK;
}
