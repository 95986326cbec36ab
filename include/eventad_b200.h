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
#define EVAD_ERR_STATIC_LIMIT (-5) /* size beyond the maximum of an entry point that works in static device scratch */

/* bits of the device status word */
#define EVAD_STATUS_EVENT_OOB 1u     /* event outside [0,W)x[0,H) or batch outside [0,B)           */
#define EVAD_STATUS_LUT_OOB 2u       /* LUT offset outside [-r, r] (reference would index out of range) */
#define EVAD_STATUS_CLUSTER_OOB 4u   /* voxel id outside the cluster table (t >= 2.0 or pos outside [0,1)) */
#define EVAD_STATUS_EDGE_OVERFLOW 8u /* coarse edge buffer too small */
#define EVAD_STATUS_H16_RANGE 16u    /* a value with |v| >= 2048 met the fp16-pair ("h2") operand format of evad_spline_conv_l1_h16 */

int evad_version(void);
/* returns the SM architecture the library was compiled for (100) */
int evad_compiled_arch(void);
/* number of kernels this library has launched in this process (for bench.py's gpu_launches) */
long long evad_launch_count(void);

/* ------------------------------------------------------------------------------------------------------------
 * Primitives (exported for the unit tests).
 * ---------------------------------------------------------------------------------------------------------- */
size_t evad_scan_ws_bytes(int64_t n);
int evad_exclusive_scan(const uint32_t* in, uint32_t* out, int64_t n, uint32_t* total, void* ws, size_t ws_bytes,
                        void* stream);
/* No counterpart in the reference: added for callers that cannot call
 * cudaMalloc (CUDA graph capture, tight serving loops, harnesses that forbid it).  evad_exclusive_scan without a
 * workspace argument: the tile-sum scratch is a file-scope `static __device__` array of 33,024 bytes
 * (csrc/no_alloc.cu; every GPU that loads the library holds it until exit), so n is limited to
 * evad_exclusive_scan_static_max_n() = 16,777,216 values; larger n returns EVAD_ERR_STATIC_LIMIT and
 * evad_no_alloc_last_error() says why.  Calls no cudaMalloc / cudaFree / allocating routine.  All calls on a device
 * share the scratch: calls ordered on one stream are safe, calls that overlap on the GPU race.  in may alias out;
 * total may be NULL. */
int evad_exclusive_scan_static(const uint32_t* in, uint32_t* out, int64_t n, uint32_t* total, void* stream);
int64_t evad_exclusive_scan_static_max_n(void);
size_t evad_exclusive_scan_static_bytes(void);
const char* evad_no_alloc_last_error(void);
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
 *   (With N == 1 the reference inserts into sample 0 whatever `batch` says, ev_graph.cu:152; after a reset the
 *   result is the self loop either way.  The streaming entry point below reproduces the quirk.)
 * ---------------------------------------------------------------------------------------------------------- */
size_t evad_graph_ws_bytes(int64_t n_events, int batch_size, int height, int width, int max_neighbors);
int evad_graph_build(const int32_t* batch, const int32_t* pos, int64_t n_events, int batch_size, int height,
                     int width, int radius, int delta_t_us, int max_queue_size, int max_neighbors,
                     int32_t* rowptr, int32_t* src, int32_t* num_edges, int64_t* edge_index,
                     int64_t edge_index_stride, int32_t* cell_perm, int32_t* cell_start,
                     uint8_t* edge_cls /* optional [K*N]: (dx + r) * (2r + 1) + (dy + r) of every CSR slot, the pixel
                                          offset source - destination the spiral walk found the edge at */,
                     uint32_t* status, void* ws, size_t ws_bytes, void* stream);
/* The same graph as FIXED-STRIDE ROWS, without the prefix scan and the compaction into a CSR: row i =
 * nbr[i*K .. i*K + count[i]) in the reference's edge order for destination i (self loop first), nbr_cls the matching
 * offset-class bytes (optional).  evad_spline_conv_l1 and evad_voxel_pool read this layout directly (ell_stride = K);
 * the reference's mask compaction `edges[:, edges[1] >= 0]` (graph/utils.py:23) is what evad_graph_build adds on top.
 * Workspace: evad_graph_ws_bytes. */
int evad_graph_build_rows(const int32_t* batch, const int32_t* pos, int64_t n_events, int batch_size, int height,
                          int width, int radius, int delta_t_us, int max_queue_size, int max_neighbors, int32_t* nbr,
                          int32_t* count, uint8_t* nbr_cls, int32_t* cell_perm, int32_t* cell_start, uint32_t* status,
                          void* ws, size_t ws_bytes, void* stream);

/* Tuning knob of the graph build (no reference counterpart).  The spiral search of an event runs over the 2r+1 row spans
 * of the pixel-sorted events when each of them holds at most min(limit, max_queue_size, 32) records, and as the per-cell
 * walk (early exit at K neighbours, FIFO cap) otherwise -- decided per event on the device; both give the same edges.
 * Default 32; 0 = always walk.  Returns the previous value; a negative argument only reads it. */
int evad_graph_set_span_limit(int limit);

/* ------------------------------------------------------------------------------------------------------------
 * Event-graph construction, streaming ("append") mode: EV_TGN.forward(reset=False)
 * (src/dagr/model/layers/ev_tgn.py:45-56), AsyncGraph.forward growth path (src/dagr/graph/ev_graph.py:75-103),
 * insert_in_queue_cuda / insert_in_queue_single_cuda / fill_edges_cuda with min_index > 0
 * (src/dagr/graph/ev_graph.cu:15-80,130-212).  `state` is caller-owned device memory of evad_stream_state_bytes()
 * holding a per-pixel ring [B*H*W][Q] of (event index, timestamp) plus insert counters; it replaces the reference's
 * [B,Q,H,W] FIFO tensor.  New events get indices base_index .. base_index+n-1; edges are reported relative to
 * min_index like the reference (SlidingWindowGraph.delete_nodes only bumps min_index, ev_graph.py:121-136).
 * do_search = 0 only inserts (used to replay the events of a preceding reset-mode call into the ring).
 * ---------------------------------------------------------------------------------------------------------- */
size_t evad_stream_state_bytes(int batch_size, int height, int width, int max_queue_size);
int evad_stream_reset(void* state, int batch_size, int height, int width, int max_queue_size, void* stream);
size_t evad_stream_ws_bytes(int64_t n_events, int max_neighbors);
int evad_stream_append(void* state, int batch_size, int height, int width, int max_queue_size, const int32_t* batch,
                       const int32_t* pos, int64_t n_events, int base_index, int min_index, int radius,
                       int delta_t_us, int max_neighbors, int do_search, int32_t* rowptr, int32_t* src,
                       int32_t* num_edges, int64_t* edge_index, int64_t edge_index_stride, uint32_t* status, void* ws,
                       size_t ws_bytes, void* stream);

/* Incremental layer updates (src/dagr/asynchronous/conv.py:110-127,183-193: new nodes are recomputed from
 * their in-edges): the CSR rows evad_stream_append returned for n_new events (rowptr[n_new+1], src relative to
 * min_index) are written into GLOBAL fixed-stride rows nbr[N][K] / count[N] / nbr_cls[N][K] at node ids
 * base_index + p, the layout evad_spline_conv_l1_h16 and evad_voxel_pool read; pos_int = [N][3] pixel / microsecond
 * coordinates of all resident events.  Class byte = (dx + r)(2r + 1) + (dy + r) like evad_graph_build_rows. */
int evad_stream_rows(const int32_t* rowptr, const int32_t* src, int64_t n_new, int base_index, int min_index,
                     const int32_t* pos_int, int radius, int max_neighbors, int32_t* nbr, int32_t* count,
                     uint8_t* nbr_cls, uint32_t* status, void* stream);

/* ------------------------------------------------------------------------------------------------------------
 * SplineConv message passing.
 * Replaces MySplineConv.forward/_forward/message_lut/init_lut (src/dagr/model/layers/spline_conv.py:16-78), PyG
 * MessagePassing.propagate (gather x_j, sum aggregation), torch_spline_conv.spline_basis/spline_weighting, and the
 * eval-mode BatchNormData + activation (+ skip branch) that ConvBlock / ConvBlockWithSkip apply right after
 * (src/dagr/model/layers/conv.py:23-28,47-56).
 *
 * Graphs are passed as dst-major CSR: rowptr[n+1], src[E]; slot order inside a row = the reference's edge order
 * for that destination, so sums are accumulated in the reference's (CPU scatter_add) order.
 * ---------------------------------------------------------------------------------------------------------- */
typedef struct {
    const float* edge_attr; /* [E, ld_attr] pseudo-coordinates in ORIGINAL edge order, or NULL -> computed from pos */
    const int32_t* eid;     /* original edge id of CSR slot e (NULL = identity); only used with edge_attr */
    int ld_attr;
    const float* pos;       /* [n, ld_pos] node positions (x, y, ...) */
    int ld_pos;
    float two_m;            /* fp32(2 * max_value) of T.Cartesian (components.py:25-35) */
    int clamp01;            /* clamp attr to [0,1] (net.py:121, first layer only) */
} evad_pseudo_t;

typedef struct {
    int use_lut;            /* 1: message_lut semantics (snap to the integer LUT lattice); 0: spline_basis on attr */
    float r00, r02, r11, r12; /* attr_remapping_matrix (spline_conv.py:21-24) */
    int nx, ny;             /* LUT extent 2*rx+1, 2*ry+1 */
    const int32_t* tab_k;   /* [nx+ny][2] kernel tap index per lattice coordinate (x table, then y table) */
    const float* tab_w;     /* [nx+ny][2] tap weights */
    int ks;                 /* kernel size per dimension */
} evad_spline_mode_t;

/* per-output-channel affine (eval BatchNorm folded: scale = gamma/sqrt(var+eps), shift = beta - mean*scale; or the
 * conv bias), optional residual addend [n, ld_add] and ReLU:  y = act(acc*scale + shift + addend) */
typedef struct {
    const float* scale;     /* [cout] or NULL (=1) */
    const float* shift;     /* [cout] or NULL (=0) */
    const float* addend;    /* [n, ld_add] or NULL */
    int ld_add;
    int relu;
} evad_epilogue_t;

/* Dense LUT of init_lut: lut[(ix*ny+iy)][cin_pad][cout] = sum_s basis_s * weight[idx_s]; rows cin..cin_pad-1 zero.
 * Two extra matrices are appended: [nx*ny] = root weight lin.weight^T, [nx*ny+1] = skip weight^T (or zeros). */
int evad_spline_lut_build(const float* weight /*[ks*ks,cin,cout]*/, const float* root_w /*[cout,cin]*/,
                          const float* skip_w /*[cout,cin] or NULL*/, int cin, int cin_pad, int cout,
                          const evad_spline_mode_t* mode_h, float* lut, void* stream);

/* Fused small-channel SplineConv (cin_pad <= 32, cout == 16), LUT resident in shared memory, one warp per
 * destination node:  out = epi(sum_e lut[ix_e,iy_e]^T x[src_e] + root^T x[i]);
 * optionally also skip_out = skip_epi(skip^T x[i])  (the Linear+BN skip branch of ConvBlockWithSkip). */
int evad_spline_conv_small(const int32_t* rowptr, const int32_t* src, int64_t n, const float* x, int ldx, int cin,
                           int cin_pad, int cout, const evad_pseudo_t* pseudo_h, const evad_spline_mode_t* mode_h,
                           const float* lut, const evad_epilogue_t* epi_h, float* out, int ldo,
                           const evad_epilogue_t* skip_epi_h, float* skip_out, int ld_skip, uint32_t* status,
                           void* stream);

/* General SplineConv, step 1: per destination node aggregate the sources per kernel tap:
 *   A[i, k*cin + c] = sum_{e: tap k} basis_e,k * x[src_e, c]   (k < ks*ks),   A[i, ks*ks*cin + c] = x[i, c]
 * n_dev (device int32, may be NULL) = actual node count when only an upper bound n is known on the host. */
int evad_spline_aggregate(const int32_t* rowptr, const int32_t* src, int64_t n, const int32_t* n_dev, const float* x,
                          int ldx, int cin, const evad_pseudo_t* pseudo_h, const evad_spline_mode_t* mode_h, float* A,
                          int64_t lda, uint32_t* status, void* stream);

/* General SplineConv, step 2 (also Linear): out[m, :] = epi(A[m, :K] @ W[K, N]) in fp32 on the CUDA cores. */
size_t evad_gemm_ws_bytes(int64_t m, int k, int n);
/* ws (optional, evad_gemm_ws_bytes): split-K partial sums for small m; without it the kernel runs unsplit. */
int evad_gemm_epilogue(const float* A, int64_t lda, const float* W /*[K,N] row-major*/, int64_t m,
                       const int32_t* m_dev, int k, int n, const evad_epilogue_t* epi_h, float* out, int ldo,
                       void* ws, size_t ws_bytes, void* stream);

/* Full-resolution SplineConv with cout = 16 (layer 1) as tap aggregation + tensor-core contraction (csrc/spline_l1.cu):
 * replaces message_lut + propagate + root weight + BatchNorm/ReLU/skip (spline_conv.py:39-47,72-76; conv.py:23-28,47-56)
 * like evad_spline_conv_small, without per-edge weight matrices.
 *   evad_edge_class: cls[e] = ix*ny + iy, the reference's LUT index of edge e (spline_conv.py:39-42, fp32 arithmetic of
 *     T.Cartesian + attr_remapping_matrix), one byte per CSR slot; shared by both convolutions of a Layer.
 *   tapw [ncls][16]: basis value of lattice class c on compact tap slot s (0 for the taps the class does not touch;
 *     slot 15 unused); Whi/Wlo [16*cin_pad][16]: TF32 hi/lo parts of the weight rows [slot][channel] (slot 15 = root
 *     weight `lin`), Shi/Slo [roundup(cin_pad,8)][16]: skip Linear (optional).  cin_pad in {16, 20}; x rows zero padded. */
int evad_edge_class(const int32_t* rowptr, const int32_t* src, int64_t n, const evad_pseudo_t* pseudo_h,
                    const evad_spline_mode_t* mode_h, uint8_t* cls, uint32_t* status, void* stream);
int evad_spline_conv_l1(const int32_t* rowptr, const int32_t* src, const uint8_t* cls,
                        const int32_t* perm /* optional processing order, e.g. cell_perm of evad_graph_build */,
                        int ell_stride /* 0: CSR, rowptr[n+1]; K > 0: fixed-stride rows of evad_graph_build_rows,
                                          rowptr[i] = length of row i, src / cls row i at i*K */,
                        int64_t n, const float* x, int ldx,
                        int cin_pad, const float* tapw, int ncls, const float* Whi, const float* Wlo, const float* Shi,
                        const float* Slo, const evad_epilogue_t* epi_h, float* out, int ldo,
                        const evad_epilogue_t* skip_epi_h, float* skip_out, int ld_skip, void* stream);
/* The same layer with the tap aggregation on the tensor cores too (one mma.sync product [16 slots x 16 row positions] @
 * [16 row positions x cin_pad] per node, csrc/spline_l1.cu second half).  Reads only the fixed-stride rows of
 * evad_graph_build_rows (K = 16: count[i], nbr / nbr_cls row i at 16*i, 16-byte aligned).  Tap weights and features
 * are both split into TF32 hi / lo parts (3xTF32).  Same operands, epilogues and result (rtol 1e-4 / atol 1e-5) as
 * evad_spline_conv_l1. */
int evad_spline_conv_l1_mma(const int32_t* count, const int32_t* nbr, const uint8_t* nbr_cls, const int32_t* perm,
                            int64_t n, const float* x, int ldx, int cin_pad, const float* tapw, int ncls,
                            const float* Whi, const float* Wlo, const float* Shi, const float* Slo,
                            const evad_epilogue_t* epi_h, float* out, int ldo, const evad_epilogue_t* skip_epi_h,
                            float* skip_out, int ld_skip, void* stream);

/* The same layer with the weight contraction on the 5th-generation tensor cores (csrc/spline_l1_tc5.cu): x_h2 are h2
 * words (see evad_spline_conv_l1_h16 below), the per-node tap aggregation is that kernel's (mma.sync m16n8k16 on fp16
 * pairs); the [128 nodes x 16*cin_pad] P tile is written as two fp16 parts (22 significant bits) in the K-major
 * SWIZZLE_128B operand layout and contracted by tcgen05.mma kind::f16 (M 128, N 16) into tensor memory; epilogue
 * (BatchNorm, skip add, ReLU; h2 words when out_h2) from tcgen05.ld.  w16 = [2 parts][16 output rows][K] fp16 (K = 256,
 * or 320 for cin_pad 20, in the kernel's operand order: MySplineConv.l1_pack), W * 2^kw = w1 + w2 * 2^-11 with
 * inv_sw = 2^-kw; s16 = [2][16][128] fp16, the skip Linear weight at the K positions of the root slot (or NULL), inv_ss
 * its un-scale.  Replaces spline_conv.py:39-47,72-76 + conv.py:23-28,47-56 like evad_spline_conv_l1; result within
 * rtol 1e-4 / atol 1e-5. */
int evad_spline_conv_l1_tc5(const int32_t* count, const int32_t* nbr, const uint8_t* nbr_cls, const int32_t* perm,
                            int64_t n, const uint32_t* x_h2, int ldx, int cin_pad, const float* tapw, int ncls,
                            const uint16_t* w16, float inv_sw, const uint16_t* s16, float inv_ss,
                            const evad_epilogue_t* epi_h, float* out, int ldo, int out_h2,
                            const evad_epilogue_t* skip_epi_h, float* skip_out, int ld_skip, uint32_t* status,
                            void* stream);

/* Layer-1 block as evad_spline_conv_l1_mma (same reference lines: spline_conv.py:39-47,72-76, conv.py:23-28,47-56), with
 * both tensor-core products on fp16 operand pairs (mma.sync m16n8k16, csrc/spline_l1_h16.cu): half the HMMA per node.
 * x_h2: node features as "h2 words" (low half = rn16(v), high half = rn16((v - hi) * 2048); |v| < 2048), produced by
 * evad_split_h2 from fp32 rows or by this entry point itself when out_h2 != 0 (the layer's second convolution gathers the
 * first one's output).  Weights as for evad_spline_conv_l1_mma (the TF32 pair; their sum is re-split in the kernel).
 * status: device word, EVAD_STATUS_H16_RANGE is raised when an h2 output leaves the range (may be NULL).
 * n nodes are computed (ids perm[0..n), or 0..n-1 without perm); n_rows >= n is the row count of x_h2 / the row arrays
 * (incremental updates compute a few new nodes of a larger resident graph: n << n_rows). */
int evad_split_h2(const float* in, uint32_t* out, int64_t n_words, uint32_t* status, void* stream);
int evad_spline_conv_l1_h16(const int32_t* count, const int32_t* nbr, const uint8_t* nbr_cls, const int32_t* perm,
                            int64_t n, int64_t n_rows, const uint32_t* x_h2, int ldx, int cin_pad, const float* tapw, int ncls,
                            const float* Whi, const float* Wlo, const float* Shi, const float* Slo,
                            const evad_epilogue_t* epi_h, float* out, int ldo, int out_h2,
                            const evad_epilogue_t* skip_epi_h, float* skip_out, int ld_skip, uint32_t* status,
                            void* stream);

/* The same contraction on the tensor cores for N = 64 (csrc/gemm_tc.cu): every product is split into TF32 parts
 * (a_hi + a_lo)(w_hi + w_lo) and accumulated in fp32 as a_lo*w_hi + a_hi*w_lo + a_hi*w_hi, so the result keeps fp32
 * accuracy (rtol 1e-4 / atol 1e-5 parity bar); the split happens in registers.  W = spline_conv.py:64-78 `weight`
 * [K,Cin,Cout] stacked with `lin.weight`^T (MySplineConv.wcat).  evad_split_tf32 writes round-to-nearest hi / lo parts
 * of a weight for kernels that keep it resident in shared memory (evad_spline_conv_l1).
 * Requirements: lda % 4 == 0, A / W 16-byte aligned, out 16-byte aligned with ldo % 4 == 0. */
int evad_split_tf32(const float* w, int64_t n, float* hi, float* lo, void* stream);
size_t evad_gemm64_tc_ws_bytes(int64_t m, int k);
int evad_gemm64_tc(const float* A, int64_t lda, const float* W /*[K,64] row-major fp32*/, int64_t m, const int32_t* m_dev,
                   int k, const evad_epilogue_t* epi_h, float* out, int ldo, void* ws, size_t ws_bytes, void* stream);

/* The same contraction on the 5th-generation tensor cores (csrc/gemm_tc5.cu): tcgen05.mma kind::tf32 with the
 * accumulator in tensor memory, 128 x 64 tile per CTA, operands staged in the SWIZZLE_128B K-major shared-memory layout,
 * 3xTF32 split as above.  Wt is W TRANSPOSED: [64, ldw] row-major (row n = the K weights of output n), ldw % 4 == 0,
 * zero padded beyond K.  Requirements: lda % 4 == 0, ldo % 4 == 0, A / Wt / out 16-byte aligned. */
size_t evad_gemm64_tc5_ws_bytes(int64_t m, int k);
int evad_gemm64_tc5(const float* A, int64_t lda, const float* Wt, int64_t ldw, int64_t m, const int32_t* m_dev, int k,
                    const evad_epilogue_t* epi_h, float* out, int ldo, void* ws, size_t ws_bytes, void* stream);

/* Pooled-layer contraction on fp16 operand pairs fed by bulk copies (csrc/gemm_h2.cu; replaces the same reference lines
 * as evad_gemm64_tc5: spline_conv.py:39-47,64-78 + conv.py:23-28,47-56).  evad_spline_aggregate_h2 = evad_spline_aggregate
 * with A written as the operand image: per (128-row tile, 64-column K chunk) one 32 KB block [a1 | a2], v = a1 + a2 *
 * 2^-11 in fp16, each a 128 x 128-byte tile in the SWIZZLE_128B K-major layout (16-byte unit u of row r stored at
 * u ^ (r & 7)); |v| < 32768 or EVAD_STATUS_H16_RANGE.  W_h2: per K chunk 16 KB [w1 | w2], 64 rows (output channels) x
 * 128 bytes, same swizzle (eventad_b200/ops.py: pack_weight_h2).  Both buffers 1024-byte aligned.  evad_gemm64_h2:
 * out[m, 64] = epi(A @ W), one cp.async.bulk pair per stage -> tcgen05.mma kind::f16 -> TMEM drain, split-K over the SMs
 * with the deterministic reduce of evad_gemm64_tc (ws: evad_gemm64_h2_ws_bytes). */
size_t evad_gemm64_h2_a_bytes(int64_t m, int k);
size_t evad_gemm64_h2_w_bytes(int k);
size_t evad_gemm64_h2_ws_bytes(int64_t m, int k);
int evad_spline_aggregate_h2(const int32_t* rowptr, const int32_t* src, int64_t n, const int32_t* n_dev, const float* x,
                             int ldx, int cin, const evad_pseudo_t* pseudo_h, const evad_spline_mode_t* mode_h,
                             void* A_h2, uint32_t* status, void* stream);
int evad_gemm64_h2(const void* A_h2, const void* W_h2, int64_t m, const int32_t* m_dev, int k,
                   const evad_epilogue_t* epi_h, float* out, int ldo, void* ws, size_t ws_bytes, void* stream);

/* ------------------------------------------------------------------------------------------------------------
 * SplineConv with Cout = 64 as one kernel (csrc/spline_tc5.cu): per-tap aggregation in registers feeding tcgen05.mma
 * (kind::tf32, 3xTF32, TMEM accumulator) -- evad_spline_aggregate + evad_gemm64_tc5 without the [n, 26*Cin] operand in HBM.
 *   evad_spline_tap_lists: per node the (source, basis value) entries of its in-edges grouped by kernel tap
 *     (spline_conv.py:39-42 LUT index or torch_spline_conv.spline_basis, as evad_spline_aggregate evaluates them);
 *     tptr [n, ntap+1] int32 entry offsets, entries [4 * rowptr[n]] (int32 source, fp32 basis value) pairs; one pass per
 *     graph, shared by the two convolutions of a Layer.
 *   evad_spline_conv64_fused: out = epi(sum_k W_k^T A_k + W_root^T x); x rows zero padded to ldx; Wt_hi / Wt_lo
 *     [64, ldw]: TF32 parts (evad_split_tf32) of the transposed stacked weight, `cp` (multiple of 8, >= ldx) columns per
 *     tap in the order taps 0..ntap-1, root; zero in the padding columns. */
int evad_spline_tap_lists(const int32_t* rowptr, const int32_t* src, int64_t n, const int32_t* n_dev,
                          const evad_pseudo_t* pseudo_h, const evad_spline_mode_t* mode_h, int32_t* tptr, void* entries,
                          uint32_t* status, void* stream);
size_t evad_spline_conv64_fused_ws_bytes(int64_t n, int ntap);
int evad_spline_conv64_fused(const int32_t* tptr, const void* entries, const float* x, int ldx, int64_t n,
                             const int32_t* n_dev, int ntap, int cp, const float* Wt_hi, const float* Wt_lo, int64_t ldw,
                             const evad_epilogue_t* epi_h, float* out, int ldo, void* ws, size_t ws_bytes, void* stream);

/* ------------------------------------------------------------------------------------------------------------
 * Voxel-grid pooling.
 * Replaces Pooling.forward (src/dagr/model/layers/pooling.py:51-97): torch_cluster.grid_cluster on [x,y,t,batch],
 * consecutive_cluster, cluster[edge_index] + self-loop drop + unique(dim=-1), pool_pos, scatter_max / _avg_pool_x,
 * round_to_pixel; and the sampling_skip that precedes every pooling call (src/dagr/model/networks/net.py:15-17,200-228)
 * when `feat` is given (the [n, C+64] concat is never materialised).
 * The input graph is dst-major CSR; the coarse graph comes out as dst-major CSR with ascending sources per row
 * (the transposed order of the reference's lexicographic unique; see evad_csr_to_coo).
 * ---------------------------------------------------------------------------------------------------------- */
typedef struct {
    float size_x, size_y, size_t; /* voxel_size[0..2] (pooling.py:24) */
    int nvx, nvy, nvt;            /* ((end - start) / size).long() + 1 per dim, as torch_cluster computes it */
    int batch_size;               /* run-time number of samples; ids of the (t == 1.0) quirk use slot batch_size */
    float inv_w, inv_h;           /* wh_inv = 1 / [W, H] in fp32 (pooling.py:32) */
    int aggr_mean;                /* 0: max (scatter_max), 1: mean (_avg_pool_x) */
    int append_pos;               /* also write the coarse (x, y) after the pooled channels (net.py:135) */
    const float* feat;            /* optional CHANNEL-LAST CNN map [feat_b, feat_h, feat_w, feat_c], or NULL */
    int feat_b, feat_h, feat_w, feat_c;
    float width, height;          /* data.width[0], data.height[0] */
    int ell_stride;               /* 0: rowptr_in is a CSR row pointer [n+1]; K > 0: the fine graph comes as fixed-stride
                                     rows (evad_graph_build_rows): row i = src_in[i*K .. i*K + rowptr_in[i]) */
    int self_loop;                /* Pooling(self_loop=True): coarse self loops are kept (pooling.py:59-60) */
} evad_pool_params_t;

size_t evad_voxel_pool_ws_bytes(int64_t n_cap, int64_t c_cap);
/* n_cap nodes at most (n_dev: device int32 actual count or NULL), c_cap >= number of output clusters (host bound),
 * e_cap = capacity of src_out.  Outputs: cluster[n_cap] dense ids in node order, n_out[1], x_out[c_cap, ldo],
 * pos_out[c_cap,3], batch_out[c_cap], rowptr_out[c_cap+1], src_out[e_cap], num_edges_out[1] (edge outputs optional:
 * pass rowptr_in = NULL to skip). */
int evad_voxel_pool(const evad_pool_params_t* params_h, int64_t n_cap, const int32_t* n_dev, const float* pos,
                    int ld_pos, const int32_t* batch, const float* x, int ldx, int cx, const int32_t* rowptr_in,
                    const int32_t* src_in, int64_t c_cap, int64_t e_cap, int32_t* cluster, int32_t* n_out,
                    float* x_out, int ldo, float* pos_out, int32_t* batch_out, int32_t* rowptr_out, int32_t* src_out,
                    int32_t* num_edges_out, uint32_t* status, void* ws, size_t ws_bytes, void* stream);

/* ------------------------------------------------------------------------------------------------------------
 * Format glue.
 * ---------------------------------------------------------------------------------------------------------- */
/* Incremental voxel pooling (src/dagr/asynchronous/max_pool.py:123-243 keeps per-voxel maxima / position
 * sums and updates the voxels that receive nodes).  `state` (evad_pool_inc_state_bytes, 256-byte aligned, caller-owned)
 * holds per voxel key the running maximum of [x | sampled CNN map], serial position sums, member count and one bitmap
 * row of coarse in-edges.  evad_pool_inc_update folds the appended nodes n0 .. n0 + n_new - 1 (ids ascending with time)
 * and their fixed-stride rows count / nbr (K = params->ell_stride) in; node_key[>= n0 + n_new] is the caller's per-node
 * key array.  evad_pool_inc_emit writes the pooled graph of all folded nodes in evad_voxel_pool's output layout:
 * bit-identical to evad_voxel_pool over the same nodes.  Max aggregation, no coarse self loops. */
size_t evad_pool_inc_state_bytes(const evad_pool_params_t* params_h, int cx);
int evad_pool_inc_reset(void* state, const evad_pool_params_t* params_h, int cx, void* stream);
size_t evad_pool_inc_update_ws_bytes(int64_t n_new);
int evad_pool_inc_update(void* state, const evad_pool_params_t* params_h, int64_t n0, int64_t n_new, const float* pos,
                         int ld_pos, const int32_t* batch, const float* x, int ldx, int cx, const int32_t* count,
                         const int32_t* nbr, uint32_t* node_key, uint32_t* status, void* ws, size_t ws_bytes, void* stream);
size_t evad_pool_inc_emit_ws_bytes(const evad_pool_params_t* params_h, int64_t c_cap);
int evad_pool_inc_emit(void* state, const evad_pool_params_t* params_h, int cx, int64_t c_cap, int64_t e_cap, int32_t* n_out,
                       float* x_out, int ldo, float* pos_out, int32_t* batch_out, int32_t* rowptr_out, int32_t* src_out,
                       int32_t* num_edges_out, uint32_t* status, void* ws, size_t ws_bytes, void* stream);

/* edge_index int64 [2][stride] in any order -> dst-major CSR (stable: slots of a row keep the original edge order,
 * i.e. the order in which PyG's scatter-add visits them); eid[slot] = original edge id. */
size_t evad_coo_to_csr_ws_bytes(int64_t num_edges);
int evad_coo_to_csr(const int64_t* edge_index, int64_t stride, int64_t num_edges, int64_t n_nodes, int32_t* rowptr,
                    int32_t* src, int32_t* eid, uint32_t* status, void* ws, size_t ws_bytes, void* stream);
/* dst-major CSR -> edge_index int64 [2][stride]; src_major = 0 keeps CSR order (the graph builder's reference order),
 * src_major = 1 gives the (src, dst)-lexicographic order of `edge_index.unique(dim=-1)` (pooling.py:64).
 * slot_of (optional) [e_cap]: CSR slot of each output column. */
size_t evad_csr_to_coo_ws_bytes(int64_t e_cap);
int evad_csr_to_coo(const int32_t* rowptr, const int32_t* src, int64_t n_cap, const int32_t* n_dev, int64_t e_cap,
                    const int32_t* num_edges, int src_major, int64_t* edge_index, int64_t stride, int32_t* slot_of,
                    void* ws, size_t ws_bytes, void* stream);
/* PyG T.Cartesian(norm=True, cat=False, max_value=m) (+ optional clamp to [0,1], net.py:121):
 * out[e, d] = (pos[row[e], d] - pos[col[e], d]) / two_m + 0.5 for d < dims. */
int evad_cartesian(const float* pos, int ld_pos, int dims, const int64_t* edge_index, int64_t stride, int64_t e_cap,
                   const int32_t* num_edges, float two_m, int clamp01, float* out, void* stream);
/* denormalize_pos (src/dagr/model/layers/ev_tgn.py:11-16): pos_int[i,d] = (int)(pos[i,d] * {W,H,T}[d] + 1e-3), fp32. */
/* format_data (src/dagr/utils/buffers.py:33-44) on the raw event layout of src/dagr/data/utils.py:6-20:
 * pos_xy int16 [n,2], t int32 [n], polarity int8 [n], batch64 int64 [n] (NULL = all zero)  ->  x float32 [n,1],
 * pos float32 [n,3] = cat([pos, t]) / [width, height, time_window], batch int32 [n]. */
int evad_format_events(const int16_t* pos_xy, const int32_t* t, const int8_t* polarity, const int64_t* batch64, int64_t n,
                       int width, int height, int time_window, float* x, float* pos, int32_t* batch, void* stream);
int evad_denormalize_pos(const float* pos, int ld_pos, int64_t n, int width, int height, int time_window,
                         int32_t* pos_int, void* stream);
/* sampling_skip + cat(x, pos[:, :2]) (net.py:15-17,116,122-123): out[i] = [x[i,:cx] | grid_sample(feat)(feat_c) |
 * pos_x, pos_y | zero padding]; feat is CHANNEL-LAST [feat_b, feat_h, feat_w, feat_c] or NULL. */
int evad_sample_concat(const float* x, int ldx, int cx, const float* pos, int ld_pos, const int32_t* batch,
                       int64_t n_cap, const int32_t* n_dev, const float* feat, int feat_b, int feat_h, int feat_w,
                       int feat_c, float width, float height, int append_pos, float* out, int ldo, void* stream);
/* the same rows for the layer-1 shape (cx 1, feat_c 16, ldo 20) written as h2 words, the input of
 * evad_spline_conv_l1_h16 (no separate evad_split_h2 pass); EVAD_ERR_UNSUPPORTED for other shapes. */
int evad_sample_concat_h2(const float* x, int ldx, int cx, const float* pos, int ld_pos, const int32_t* batch,
                          int64_t n_cap, const int32_t* n_dev, const float* feat, int feat_b, int feat_h, int feat_w,
                          int feat_c, float width, float height, int append_pos, uint32_t* out, int ldo,
                          const int32_t* visit /* optional permutation: thread order over the nodes, output unchanged */,
                          uint32_t* status, void* stream);

/* ---- Event ingest on the device (csrc/ingest.cu) ----
 * A recording's arrays (int32 x, y, t in microseconds relative to t_offset and ascending; int8 p) stay in HBM.
 * evad_ingest_ms_to_idx: create_ms_to_idx (/root/reference/scripts/downsample_events.py:36-42): ms_to_idx[m] = number of
 *   events with t / 1000 < m, m < n_ms (= t[n-1] / 1000 + 1).
 * evad_ingest_slice: the array half of _load_events (/root/reference/src/dagr/data/dsec_utils.py:80-120) for events
 *   [idx0, idx0 + cnt): xy int16 [.,2], t = time_window - tq + t_ev (int32 [.,1]), p = 2 p - 1 (int8 [.,1]), kept where
 *   t > 0 and y < height (height < 0: no crop), order preserved; *n_out = survivors (device).
 * evad_ingest_downsample: downsample_events + _filter_events_resize (downsample_events.py:91-122): the per-cell
 *   integrate-and-fire filter (float32 cells, float64 expression) over a stable sort by coarse cell; change_map
 *   [out_h, out_w] is read and updated; survivors in original order with x / fx, y / fy. */
int evad_ingest_ms_to_idx(const int32_t* t_us, int64_t n, int64_t n_ms, int64_t* ms_to_idx, void* stream);
size_t evad_ingest_slice_ws_bytes(int64_t cnt);
int evad_ingest_slice(const int32_t* x, const int32_t* y, const int32_t* t, const int8_t* p, int64_t idx0, int64_t cnt,
                      int64_t tq, int64_t time_window, int height, int16_t* xy_out, int32_t* t_out, int8_t* p_out,
                      int32_t* n_out, void* ws, size_t ws_bytes, void* stream);
size_t evad_ingest_downsample_ws_bytes(int64_t n);
int evad_ingest_downsample(const int32_t* x, const int32_t* y, const int32_t* t, const int8_t* p, int64_t n, int fx, int fy,
                           int out_w, int out_h, float* change_map, int32_t* x_out, int32_t* y_out, int32_t* t_out,
                           int8_t* p_out, int32_t* n_out, uint32_t* status, void* ws, size_t ws_bytes, void* stream);

/* ---- EventAD head input: per-bounding-box pooling of the out4 graph (csrc/bbox_pool.cu) ----
 * Replaces the python loops of /root/reference/models/EventAD.py:382-397 and _process_bbox (:406-498): for every box
 * [x, y, w, h, class, track_id] (pixels) of `bbox` (sample ids bbox_batch int64) the feature of the nodes of its sample
 * that lie inside the box, weighted by the inverse distance to the box centre; without such nodes the mean of the 5
 * closest nodes; written to batch_features[sample][frame][track_id][channels] (caller zero-fills; of several boxes with
 * one (sample, track) the last wins, as in the reference's sequential loop).  width / height: fp32 [B].  node_batch int32
 * [n_cap] with *n_dev valid rows (NULL: n_cap).  One launch per frame list (frame 1 = bbox, frame 0 = bbox0). */
int evad_bbox_pool(const float* feat, int ldf, int channels, const float* pos, int ldp, const int32_t* node_batch,
                   int64_t n_cap, const int32_t* n_dev, const float* bbox, int ldb, const int64_t* bbox_batch,
                   int num_boxes, int frame, const float* width, const float* height, int n_frames, int max_boxes,
                   float* batch_features, void* stream);

/* ---- asynchronous (incremental) layers: masked row updates (csrc/asy.cu) ----
 * C-ABI counterparts of the reference's `asy_tools` pybind module (/root/reference/src/dagr/asynchronous/asy_tools/main.cu
 * :239-244), bit-identical results:
 *   evad_asy_masked_lin     masked_lin (:215-237) when bias != NULL, masked_lin_no_bias (:190-213) when NULL:
 *                           x_out[idx] (+)= x_in[idx] @ weight^T (+ bias), weight [cout, cin] row-major, `add` keeps x_out
 *   evad_asy_masked_bn      masked_inplace_BN (:69-95): eval BatchNorm of the selected rows of x into x_out
 *   evad_asy_masked_isdiff  masked_isdiff (:97-125) without the host-side compaction: indices[k] stays where row
 *                           indices[k] of x_a and x_b differ (|a - b| > atol + rtol * b), else becomes -1
 * indices int64 [k] row numbers; x rows are dense [*, channels] fp32. */
int evad_asy_masked_lin(const int64_t* indices, int64_t k, const float* x_in, float* x_out, const float* weight,
                        const float* bias, int cin, int cout, int add, void* stream);
int evad_asy_masked_bn(const int64_t* indices, int64_t k, const float* x, float* x_out, const float* running_mean,
                       const float* running_var, const float* weight, const float* bias, int channels, float eps,
                       void* stream);
int evad_asy_masked_isdiff(int64_t* indices, int64_t k, const float* x_a, const float* x_b, int channels, float atol,
                           float rtol, void* stream);

#ifdef __cplusplus
}
#endif
#endif /* EVENTAD_B200_H */
