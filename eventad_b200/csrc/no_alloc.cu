// eventad_b200/csrc/no_alloc.cu -- an exclusive scan that allocates nothing: its scratch is a file-scope device array.
//
// No counterpart in the reference; added for callers that cannot call
// cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses that forbid it).  It is the scan of
// primitives.cu (evad_exclusive_scan_u32: tile reduce -> scan of the tile sums -> apply) with the tile-sum scratch taken
// from `g_scan_scratch` below instead of a caller workspace.
//
//   * limit: n <= EVAD_SCAN_STATIC_MAX_N (16,777,216 values); the array is 8,256 words = 33,024 bytes, and every GPU
//     that loads the library holds it until the process exits.  Larger n is rejected with EVAD_ERR_STATIC_LIMIT.
//   * `g_scan_scratch` is a device symbol: in HOST code its name is a host-side stand-in, not a device address.  The
//     device address is taken once per device, with that device current, by cudaGetSymbolAddress, and that pointer is
//     what the kernels get.
//   * every call on a device shares the array and finds the previous call's contents in it.  Nothing has to start at
//     zero: the reduce kernel writes every tile sum the later kernels read.  Two calls that OVERLAP on one GPU (different
//     streams, or two host threads) race on it; calls on one stream are ordered and safe.
//   * no cudaMalloc / cudaFree / allocating library routine is called: cudaGetDevice, cudaGetSymbolAddress and three
//     kernel launches only.
#include <mutex>

#include "common.cuh"

#define EVAD_SCAN_STATIC_MAX_N (1ll << 24)
#define EVAD_SCAN_STATIC_WORDS 8256           // evad_scan_tmp_elems(2^24) = 2^24 / 2048 + 64
#define EVAD_MAX_DEVICES 64

static __device__ __align__(256) uint32_t g_scan_scratch[EVAD_SCAN_STATIC_WORDS];

static std::mutex g_scratch_mutex;
static uint32_t* g_scratch_addr[EVAD_MAX_DEVICES];   // device address of g_scan_scratch, per device ordinal
static thread_local const char* g_no_alloc_error = "";

extern "C" int64_t evad_exclusive_scan_static_max_n(void) { return EVAD_SCAN_STATIC_MAX_N; }
extern "C" size_t evad_exclusive_scan_static_bytes(void) { return sizeof(uint32_t) * EVAD_SCAN_STATIC_WORDS; }
extern "C" const char* evad_no_alloc_last_error(void) { return g_no_alloc_error; }

// device address of the scratch on the CURRENT device (resolved on first use per device)
static int scan_scratch_ptr(uint32_t** out) {
    int dev = -1;
    if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= EVAD_MAX_DEVICES) {
        g_no_alloc_error = "evad_exclusive_scan_static: no current CUDA device (or ordinal >= 64)";
        return EVAD_ERR_ARG;
    }
    std::lock_guard<std::mutex> lock(g_scratch_mutex);
    if (g_scratch_addr[dev] == nullptr) {
        uint32_t* p = nullptr;
        cudaError_t e = cudaGetSymbolAddress((void**)&p, g_scan_scratch);
        if (e != cudaSuccess || p == nullptr) {
            g_no_alloc_error = "evad_exclusive_scan_static: cudaGetSymbolAddress(g_scan_scratch) failed";
            return EVAD_ERR_LAUNCH;
        }
        g_scratch_addr[dev] = p;
    }
    *out = g_scratch_addr[dev];
    return EVAD_OK;
}

extern "C" int evad_exclusive_scan_static(const uint32_t* in, uint32_t* out, int64_t n, uint32_t* total, void* stream) {
    g_no_alloc_error = "";
    if (!in || !out || n < 0) {
        g_no_alloc_error = "evad_exclusive_scan_static: null pointer or negative n";
        return EVAD_ERR_ARG;
    }
    if (n > EVAD_SCAN_STATIC_MAX_N) {
        g_no_alloc_error = "evad_exclusive_scan_static: n exceeds the static scratch's maximum of 16777216 values "
                           "(use evad_exclusive_scan with a caller workspace)";
        return EVAD_ERR_STATIC_LIMIT;
    }
    if (evad_scan_tmp_elems(n) > EVAD_SCAN_STATIC_WORDS) {       // guards the constant against a change of the tile size
        g_no_alloc_error = "evad_exclusive_scan_static: static scratch smaller than the scan needs";
        return EVAD_ERR_STATIC_LIMIT;
    }
    uint32_t* scratch = nullptr;
    EVAD_TRY(scan_scratch_ptr(&scratch));
    int rc = evad_exclusive_scan_u32(in, out, n, total, scratch, (cudaStream_t)stream);
    if (rc != EVAD_OK) g_no_alloc_error = "evad_exclusive_scan_static: kernel launch failed";
    return rc;
}
