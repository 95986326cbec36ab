/*
 * oracle/asy_oracle.c -- CPU restatement of the reference's `asy_tools` CUDA extension
 * (/root/reference/src/dagr/asynchronous/asy_tools/main.cu), the masked row-update kernels of the asynchronous
 * (incremental) DAGR layers.
 *
 * TEST INFRASTRUCTURE ONLY: used by tests/ as the checker of eventad_b200/csrc/asy.cu and by oracle/ref_py.py as the
 * CPU stand-in of the `asy_tools` module under the reference's own Python.  Never linked or loaded by eventad_b200/.
 *
 * Each function is the kernel body with "one thread per item" turned into a loop:
 *   asyo_masked_isdiff      main.cu:14-41   (launcher :97-125; the host-side `indices.index({indices > -1})` stays in Python)
 *   asyo_masked_inplace_bn  main.cu:43-67   (launcher :69-95)
 *   asyo_masked_lin         main.cu:128-158 (launcher :215-237), bias != NULL
 *   asyo_masked_lin         main.cu:160-188 (launcher :190-213), bias == NULL
 * Parity status: pinned on the GPU box against the UNMODIFIED main.cu built by oracle/build_ref.py
 * (oracle/_ref/asy_tools.so; tests/test_asy_gpu.py).
 */
#include <math.h>
#include <stdint.h>

/* indices[k] <- indices[k] if row differs else -1;  |old - new| > atol + rtol * new, first arg named x_old is the
 * caller's NEW tensor at the call sites (base/utils.py:42-43) -- argument order kept as in main.cu */
void asyo_masked_isdiff(int64_t *indices, const float *x_old, const float *x_new, int64_t K, int C, float atol, float rtol) {
    for (int64_t k = 0; k < K; k++) {
        int64_t temp = indices[k];
        indices[k] = -1;
        int64_t off = temp * C;
        for (int i = 0; i < C; i++) {
            float input = x_old[off + i], other = x_new[off + i];
            if (fabsf(input - other) > fmaf(rtol, other, atol)) { indices[k] = temp; break; }   /* nvcc -fmad=true */
        }
    }
}

void asyo_masked_inplace_bn(const int64_t *indices, const float *x, float *x_out, const float *mean, const float *var,
                            const float *weight, const float *bias, int64_t K, int C, float eps) {
    for (int64_t k = 0; k < K; k++)
        for (int c = 0; c < C; c++) {
            int64_t j = (int64_t)C * indices[k] + c;
            /* main.cu:66: `sqrt(float + float)` in device code resolves to the float overload */
            x_out[j] = fmaf((x[j] - mean[c]) / (sqrtf(var[c] + eps)), weight[c], bias[c]);   /* a*w+b contracts under nvcc -fmad=true */
        }
}

void asyo_masked_lin(const int64_t *indices, const float *x_in, float *x_out, const float *weight, const float *bias,
                     int64_t K, int Cin, int Cout, int add) {
    for (int64_t k = 0; k < K; k++)
        for (int co = 0; co < Cout; co++) {
            int64_t o = (int64_t)Cout * indices[k] + co;
            int64_t in0 = (int64_t)Cin * indices[k];
            float acc = add ? x_out[o] : 0.f;                  /* serial read-modify-write of x_out in the kernel */
            for (int ci = 0; ci < Cin; ci++) {
                /* nvcc contracts `x_out[o] += a * b` into an FMA; fmaf keeps this restatement bit-identical to it */
                acc = fmaf(x_in[in0 + ci], weight[(int64_t)co * Cin + ci], acc);
            }
            if (bias) acc += bias[co];
            x_out[o] = acc;
        }
}
