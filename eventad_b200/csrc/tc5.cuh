// eventad_b200/csrc/tc5.cuh -- tcgen05 / TMEM / mbarrier / cp.async PTX wrappers shared by the sm_100a tensor-core kernels
// (gemm_tc5.cu, spline_tc5.cu).  Tile shape: 128 x 64 output per CTA, K walked in 32-float (128-byte) swizzle rows.
#pragma once
#include "common.cuh"

#define T5_BM 128
#define T5_BN 64
#define T5_BK 32                                             // floats per slab = 128 bytes = one swizzle row
#define T5_A_BYTES (T5_BM * 128)                             // 16 KB
#define T5_B_BYTES (T5_BN * 128)                             // 8 KB
#define T5_OB_BYTES (2 * T5_A_BYTES + 2 * T5_B_BYTES)        // operand buffer: A_hi, A_lo, W_hi, W_lo = 48 KB
#define T5_TMEM_COLS 128                                     // two 64-column accumulator buffers

// ---- PTX wrappers (sm_100a) ----
__device__ __forceinline__ uint32_t t5_smem(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }

__device__ __forceinline__ void t5_cp16(uint32_t dst, const void* gmem, int src_bytes) {
    asm volatile("cp.async.cg.shared.global [%0], [%1], 16, %2;" ::"r"(dst), "l"(gmem), "r"(src_bytes));
}
__device__ __forceinline__ void t5_commit_group() { asm volatile("cp.async.commit_group;"); }
template <int N>
__device__ __forceinline__ void t5_wait_group() { asm volatile("cp.async.wait_group %0;" ::"n"(N) : "memory"); }

__device__ __forceinline__ void t5_mbar_init(uint32_t bar, int count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(bar), "r"(count));
}
__device__ __forceinline__ void t5_mbar_wait(uint32_t bar, uint32_t parity) {
    asm volatile(
        "{\n\t"
        ".reg .pred p;\n\t"
        "T5_WAIT:\n\t"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n\t"
        "@p bra T5_DONE;\n\t"
        "bra T5_WAIT;\n\t"
        "T5_DONE:\n\t"
        "}\n" ::"r"(bar), "r"(parity)
        : "memory");
}

__device__ __forceinline__ void t5_mbar_arrive(uint32_t bar) {
    asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(bar) : "memory");
}

// 64-bit shared-memory matrix descriptor: K-major operand tile, 128-byte swizzle, rows of 128 bytes, 8-row groups
// 1024 bytes apart (cute::UMMA::SmemDescriptor: start >> 4 | LBO(16 B) << 16 | SBO(1024 B) << 32 | version 1 << 46 |
// SWIZZLE_128B (2) << 61).  The tile base is 1024-byte aligned; a K step of 8 floats advances the start by 32 bytes.
__device__ __forceinline__ uint64_t t5_desc(uint32_t smem_addr) {
    return (uint64_t)((smem_addr & 0x3FFFFu) >> 4) | (1ull << 16) | (64ull << 32) | (1ull << 46) | (2ull << 61);
}

// instruction descriptor (cute::UMMA::InstrDescriptor): D = F32 (1 << 4), A = B = TF32 (2 << 7, 2 << 10), both K-major,
// N = 64 (N >> 3 at bit 17), M = 128 (M >> 4 at bit 24)
#define T5_IDESC ((1u << 4) | (2u << 7) | (2u << 10) | ((T5_BN >> 3) << 17) | ((T5_BM >> 4) << 24))

__device__ __forceinline__ void t5_mma(uint32_t tmem_d, uint64_t adesc, uint64_t bdesc, uint32_t accumulate) {
    asm volatile(
        "{\n\t"
        ".reg .pred p;\n\t"
        "setp.ne.b32 p, %4, 0;\n\t"
        "tcgen05.mma.cta_group::1.kind::tf32 [%0], %1, %2, %3, p;\n\t"
        "}\n" ::"r"(tmem_d), "l"(adesc), "l"(bdesc), "r"(T5_IDESC), "r"(accumulate)
        : "memory");
}
__device__ __forceinline__ void t5_mma_commit(uint32_t bar) {
    asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(bar) : "memory");
}


// split an fp32 quad into TF32 parts: hi = upper 19 bits (what the tensor core reads), lo = exact remainder
__device__ __forceinline__ void t5_split4(const float4 v, float4* h, float4* l) {
    h->x = __uint_as_float(__float_as_uint(v.x) & 0xffffe000u); l->x = __fsub_rn(v.x, h->x);
    h->y = __uint_as_float(__float_as_uint(v.y) & 0xffffe000u); l->y = __fsub_rn(v.y, h->y);
    h->z = __uint_as_float(__float_as_uint(v.z) & 0xffffe000u); l->z = __fsub_rn(v.z, h->z);
    h->w = __uint_as_float(__float_as_uint(v.w) & 0xffffe000u); l->w = __fsub_rn(v.w, h->w);
}

// 32 lanes x 32 columns of a TMEM accumulator into registers (one row per thread)
__device__ __forceinline__ void t5_ld32(uint32_t taddr, uint32_t (&r)[32]) {
    asm volatile(
        "tcgen05.ld.sync.aligned.32x32b.x32.b32 "
        "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, "
        "%16, %17, %18, %19, %20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31}, [%32];"
        : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]),
          "=r"(r[8]), "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15]),
          "=r"(r[16]), "=r"(r[17]), "=r"(r[18]), "=r"(r[19]), "=r"(r[20]), "=r"(r[21]), "=r"(r[22]), "=r"(r[23]),
          "=r"(r[24]), "=r"(r[25]), "=r"(r[26]), "=r"(r[27]), "=r"(r[28]), "=r"(r[29]), "=r"(r[30]), "=r"(r[31])
        : "r"(taddr)
        : "memory");
    asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
}

// 32 lanes x 16 columns
__device__ __forceinline__ void t5_ld16(uint32_t taddr, uint32_t (&r)[16]) {
    asm volatile(
        "tcgen05.ld.sync.aligned.32x32b.x16.b32 "
        "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15}, [%16];"
        : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]),
          "=r"(r[8]), "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15])
        : "r"(taddr)
        : "memory");
    asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
}
