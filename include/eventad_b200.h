/*
 * include/eventad_b200.h -- C ABI of libeventad_b200.so (sm_100a kernels for EventAD's DAGR event branch).
 *
 * Conventions
 *   - every pointer is a DEVICE pointer unless its name ends in _h (host);
 *   - `stream` is a cudaStream_t passed as void*; every call only ENQUEUES work on it: no allocation, no
 *     synchronisation, no host read-back.  Data-dependent sizes (edge counts, cluster counts) stay on the device
 *     in small int32 "count" cells; output arrays are sized by the host-known upper bounds documented per call;
 *   - `ws`/`ws_bytes` is caller-owned scratch (query the size with the matching *_ws_bytes function);
 *   - return value: EVAD_OK (0) or a negative EVAD_ERR_* code.  Device-side argument problems (an event outside
 *     the sensor, a LUT offset outside its table) raise bits in the caller-provided `status` cell instead, because
 *     nothing may synchronise; the Python layer turns both into RuntimeError like the reference's AT_ASSERTM
 *     (/root/reference/src/dagr/graph/ev_graph.cu:9-12).
 *
 * Each entry point names the reference interface it replaces (paths relative to /root/reference).
 */
#ifndef EVENTAD_B200_H
#define EVENTAD_B200_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define EVAD_OK 0
#define EVAD_ERR_ARG (-1)       /* bad argument (null pointer, negative size, unsupported width) */
#define EVAD_ERR_WORKSPACE (-2) /* workspace too small */
#define EVAD_ERR_LAUNCH (-3)    /* cudaGetLastError() after a launch */
#define EVAD_ERR_UNSUPPORTED (-4)

/* bits of the device status word */
#define EVAD_STATUS_EVENT_OOB 1u     /* event outside [0,W)x[0,H) or batch outside [0,B)           */
#define EVAD_STATUS_LUT_OOB 2u       /* LUT offset outside [-r, r] (reference would index out of range) */
#define EVAD_STATUS_CLUSTER_OOB 4u   /* voxel id outside the cluster table (t >= 2.0 or pos outside [0,1)) */
#define EVAD_STATUS_EDGE_OVERFLOW 8u /* coarse edge buffer too small */

int evad_version(void);
/* returns the SM architecture the library was compiled for (100) */
int evad_compiled_arch(void);

/* ------------------------------------------------------------------------------------------------------------
 * Primitives (exported for the unit tests).
 * ---------------------------------------------------------------------------------------------------------- */
size_t evad_scan_ws_bytes(int64_t n);
int evad_exclusive_scan(const uint32_t* in, uint32_t* out, int64_t n, uint32_t* total, void* ws, size_t ws_bytes,
                        void* stream);
size_t evad_sort_ws_bytes(int64_t n);
/* stable sort of (key,val) pairs on the low key_bits bits; keys_in/vals_in are clobbered; vals_in may be NULL
 * (values = 0..n-1).  Result in keys_out/vals_out. */
int evad_sort_pairs(uint32_t* keys_in, uint32_t* vals_in, uint32_t* keys_out, uint32_t* vals_out, int64_t n,
                    int key_bits, void* ws, size_t ws_bytes, void* stream);

/* ------------------------------------------------------------------------------------------------------------
 * Event-graph construction, synchronous ("reset") mode.
 * Replaces, for one call of AsyncGraph.forward after reset() (src/dagr/graph/ev_graph.py:63-103):
 *   _insert_events_into_queue + insert_in_queue_cuda / insert_in_queue_single_cuda
 *       (src/dagr/graph/utils.py:6-18, src/dagr/graph/ev_graph.cu:130-212,215-276)
 *   _search_for_edges + fill_edges_cuda + the boolean-mask compaction
 *       (src/dagr/graph/utils.py:20-23, src/dagr/graph/ev_graph.cu:15-128)
 * without the [B,Q,H,W] FIFO tensor: events are bucketed per pixel by a stable sort, which holds exactly the FIFO
 * content (newest Q indices per pixel).
 *
 *   batch[N] int32, pos[N*3] int32 (x px, y px, t us)        -- same tensors the reference kernel takes
 *   radius (px, Chebyshev), delta_t_us, Q = max_queue_size, K = max_num_neighbors (self loop included, K <= 32)
 * Outputs (dst-major CSR in event order = the reference's edge order):
 *   rowptr[N+1] int32, src[K*N] int32 (first E entries valid), num_edges[1] int32,
 *   edge_index (optional, may be NULL) int64 [2][edge_index_stride]: row 0 = src, row 1 = dst, first E columns.
 *   cell_perm (optional) int32[N]: event ids ordered by (b,y,x) pixel then id; cell_start (optional) int32[B*H*W+1].
 *   single_event_quirk: when N == 1 the reference ignores `batch` and inserts into sample 0
 *   (ev_graph.cu:152); the flag reproduces that.
 * ---------------------------------------------------------------------------------------------------------- */
size_t evad_graph_ws_bytes(int64_t n_events, int batch_size, int height, int width, int max_neighbors);
int evad_graph_build(const int32_t* batch, const int32_t* pos, int64_t n_events, int batch_size, int height,
                     int width, int radius, int delta_t_us, int max_queue_size, int max_neighbors,
                     int32_t* rowptr, int32_t* src, int32_t* num_edges, int64_t* edge_index,
                     int64_t edge_index_stride, int32_t* cell_perm, int32_t* cell_start, uint32_t* status,
                     void* ws, size_t ws_bytes, void* stream);

#ifdef __cplusplus
}
#endif
#endif /* EVENTAD_B200_H */
