// eventad_b200/csrc/common.cuh -- shared helpers for the sm_100a kernels (no torch headers anywhere in csrc/).
#pragma once
#include <cuda_fp16.h>
#include <cuda_runtime.h>
#include <stdint.h>
#include <stddef.h>

#include "../../include/eventad_b200.h"

// every kernel launch of the library is followed by this macro: it also feeds evad_launch_count()
extern long long g_evad_launches;
#define EVAD_CHECK_LAUNCH()                                   \
    do {                                                      \
        cudaError_t e__ = cudaGetLastError();                 \
        if (e__ != cudaSuccess) return EVAD_ERR_LAUNCH;       \
        ++g_evad_launches;                                    \
    } while (0)

#define EVAD_TRY(expr)                                        \
    do {                                                      \
        int r__ = (expr);                                     \
        if (r__ != EVAD_OK) return r__;                       \
    } while (0)

static inline int evad_cdiv(long long a, long long b) { return (int)((a + b - 1) / b); }
static inline size_t evad_align(size_t x, size_t a = 256) { return (x + a - 1) / a * a; }

// bump allocator over a caller-provided workspace; never allocates device memory itself
struct EvadArena {
    char* base;
    size_t cap;
    size_t off;
    bool ok;
    EvadArena(void* p, size_t bytes) : base((char*)p), cap(bytes), off(0), ok(true) {}
    template <typename T>
    T* take(size_t n) {
        size_t bytes = evad_align(n * sizeof(T));
        if (base != nullptr && off + bytes > cap) ok = false;
        T* r = base ? (T*)(base + off) : nullptr;
        off += bytes;
        return r;
    }
};

// ---- primitives implemented in primitives.cu ----
// exclusive scan of n uint32 (in may alias out); if total != nullptr the grand total is written to *total.
// tmp must hold evad_scan_tmp_elems(n) uint32.
size_t evad_scan_tmp_elems(long long n);
int evad_exclusive_scan_u32(const uint32_t* in, uint32_t* out, long long n, uint32_t* total, uint32_t* tmp,
                            cudaStream_t st);
// stable LSD radix sort of (key,val) uint32 pairs on the low key_bits bits, ping-ponging between the buffer
// pairs (k0,v0) and (k1,v1); the input is in k0 (and v0 unless iota_vals, which means val = position).
// The buffer pair that holds the result is returned through res_k/res_v (host pointers, known at enqueue time).
size_t evad_sort_tmp_elems(long long n);
int evad_sort_pairs_u32(uint32_t* k0, uint32_t* v0, uint32_t* k1, uint32_t* v1, int iota_vals, long long n,
                        int key_bits, uint32_t* tmp, uint32_t** res_k, uint32_t** res_v, cudaStream_t st);

// deterministic split-K reduce + epilogue for [m, 64] outputs (gemm_tc.cu): out = epi(sum_y partial[y][m][64])
int evad_gemm64_reduce(const float* partial, int splits, int64_t m, const int32_t* m_dev, const evad_epilogue_t* E,
                       float* out, int ldo, cudaStream_t st);

// "h2 word" of an fp32 value (operand format of csrc/spline_l1_h16.cu): low half = rn16(v), high half = rn16((v - hi) * 2048)
#define EVAD_H2_RANGE 2048.f
__device__ __forceinline__ uint32_t evad_h2_split(float v) {
    const __half h = __float2half_rn(v);
    const __half l = __float2half_rn(__fmul_rn(__fsub_rn(v, __half2float(h)), 2048.f));
    return (uint32_t)__half_as_ushort(h) | ((uint32_t)__half_as_ushort(l) << 16);
}

__device__ __forceinline__ unsigned evad_lanemask_lt() {
    unsigned m;
    asm("mov.u32 %0, %%lanemask_lt;" : "=r"(m));
    return m;
}
