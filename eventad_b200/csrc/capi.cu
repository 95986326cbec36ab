// eventad_b200/csrc/capi.cu -- small C-ABI entry points that wrap the primitives (see include/eventad_b200.h).
#include "common.cuh"

long long g_evad_launches = 0;

extern "C" int evad_version(void) { return 1; }

extern "C" long long evad_launch_count(void) { return g_evad_launches; }

extern "C" int evad_compiled_arch(void) {
#ifdef EVAD_ARCH
    return EVAD_ARCH;
#else
    return 100;
#endif
}

extern "C" size_t evad_scan_ws_bytes(int64_t n) { return evad_scan_tmp_elems(n) * sizeof(uint32_t) + 256; }

extern "C" int evad_exclusive_scan(const uint32_t* in, uint32_t* out, int64_t n, uint32_t* total, void* ws,
                                   size_t ws_bytes, void* stream) {
    if (!in || !out || !ws) return EVAD_ERR_ARG;
    if (ws_bytes < evad_scan_ws_bytes(n)) return EVAD_ERR_WORKSPACE;
    return evad_exclusive_scan_u32(in, out, n, total, (uint32_t*)ws, (cudaStream_t)stream);
}

extern "C" size_t evad_sort_ws_bytes(int64_t n) { return evad_sort_tmp_elems(n) * sizeof(uint32_t) + 256; }

__global__ void copy_pairs_kernel(const uint32_t* __restrict__ k, const uint32_t* __restrict__ v,
                                  uint32_t* __restrict__ ko, uint32_t* __restrict__ vo, int64_t n) {
    int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) { ko[i] = k[i]; vo[i] = v[i]; }
}

extern "C" int evad_sort_pairs(uint32_t* keys_in, uint32_t* vals_in, uint32_t* keys_out, uint32_t* vals_out,
                               int64_t n, int key_bits, void* ws, size_t ws_bytes, void* stream) {
    if (!keys_in || !keys_out || !vals_out || !ws) return EVAD_ERR_ARG;
    if (ws_bytes < evad_sort_ws_bytes(n) + (vals_in ? 0 : evad_align((size_t)(n > 0 ? n : 1) * 4))) return EVAD_ERR_WORKSPACE;
    cudaStream_t st = (cudaStream_t)stream;
    uint32_t* tmp = (uint32_t*)ws;
    uint32_t* v0 = vals_in;
    if (!v0) {                                               // iota: still need a ping-pong buffer for the values
        v0 = (uint32_t*)((char*)ws + evad_sort_ws_bytes(n) - 256);
    }
    uint32_t *rk, *rv;
    EVAD_TRY(evad_sort_pairs_u32(keys_in, v0, keys_out, vals_out, vals_in == nullptr, n, key_bits, tmp, &rk, &rv, st));
    if (rk != keys_out && n > 0) {
        copy_pairs_kernel<<<evad_cdiv(n, 256), 256, 0, st>>>(rk, rv, keys_out, vals_out, n);
        EVAD_CHECK_LAUNCH();
    }
    return EVAD_OK;
}
