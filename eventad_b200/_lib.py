"""ctypes loader of libeventad_b200.so (the C ABI in include/eventad_b200.h).

There is no CPU fallback: if the library is missing or a call fails, a RuntimeError is raised (the reference raises
RuntimeError through AT_ASSERTM for non-CUDA / non-contiguous inputs, /root/reference/src/dagr/graph/ev_graph.cu:9-12).
"""
import ctypes
import os

import torch

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "libeventad_b200.so")
_lib = None

ERRORS = {-1: "EVAD_ERR_ARG (bad argument)", -2: "EVAD_ERR_WORKSPACE (workspace too small)",
          -3: "EVAD_ERR_LAUNCH (CUDA launch failed)", -4: "EVAD_ERR_UNSUPPORTED",
          -5: "EVAD_ERR_STATIC_LIMIT (size beyond the static-scratch entry point's maximum)"}
STATUS_BITS = {1: "event outside the sensor / batch range", 2: "LUT offset outside its table",
               4: "voxel id outside the cluster table", 8: "coarse edge buffer overflow",
               16: "feature value with |v| >= 2048 in the fp16-pair layer-1 kernel (set EVAD_L1_H16=0)"}

c_p = ctypes.c_void_p
c_i = ctypes.c_int
c_i64 = ctypes.c_int64
c_sz = ctypes.c_size_t
c_f = ctypes.c_float


class PseudoT(ctypes.Structure):
    _fields_ = [("edge_attr", c_p), ("eid", c_p), ("ld_attr", c_i), ("pos", c_p), ("ld_pos", c_i),
                ("two_m", c_f), ("clamp01", c_i)]


class SplineModeT(ctypes.Structure):
    _fields_ = [("use_lut", c_i), ("r00", c_f), ("r02", c_f), ("r11", c_f), ("r12", c_f), ("nx", c_i), ("ny", c_i),
                ("tab_k", c_p), ("tab_w", c_p), ("ks", c_i)]


class EpilogueT(ctypes.Structure):
    _fields_ = [("scale", c_p), ("shift", c_p), ("addend", c_p), ("ld_add", c_i), ("relu", c_i)]


class PoolParamsT(ctypes.Structure):
    _fields_ = [("size_x", c_f), ("size_y", c_f), ("size_t", c_f), ("nvx", c_i), ("nvy", c_i), ("nvt", c_i),
                ("batch_size", c_i), ("inv_w", c_f), ("inv_h", c_f), ("aggr_mean", c_i), ("append_pos", c_i),
                ("feat", c_p), ("feat_b", c_i), ("feat_h", c_i), ("feat_w", c_i), ("feat_c", c_i),
                ("width", c_f), ("height", c_f), ("ell_stride", c_i), ("self_loop", c_i)]


def lib():
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise RuntimeError(f"{LIB_PATH} is missing: run `python -m eventad_b200.build` (there is no CPU fallback)")
        L = ctypes.CDLL(LIB_PATH)
        _declare(L)
        _lib = L
    return _lib


def _sig(L, name, restype, argtypes):
    f = getattr(L, name)
    f.restype = restype
    f.argtypes = argtypes


def _declare(L):
    _sig(L, "evad_version", c_i, [])
    _sig(L, "evad_compiled_arch", c_i, [])
    _sig(L, "evad_launch_count", ctypes.c_longlong, [])
    _sig(L, "evad_scan_ws_bytes", c_sz, [c_i64])
    _sig(L, "evad_exclusive_scan", c_i, [c_p, c_p, c_i64, c_p, c_p, c_sz, c_p])
    _sig(L, "evad_exclusive_scan_static", c_i, [c_p, c_p, c_i64, c_p, c_p])
    _sig(L, "evad_exclusive_scan_static_max_n", c_i64, [])
    _sig(L, "evad_exclusive_scan_static_bytes", c_sz, [])
    _sig(L, "evad_no_alloc_last_error", ctypes.c_char_p, [])
    _sig(L, "evad_sort_ws_bytes", c_sz, [c_i64])
    _sig(L, "evad_sort_pairs", c_i, [c_p, c_p, c_p, c_p, c_i64, c_i, c_p, c_sz, c_p])
    _sig(L, "evad_graph_ws_bytes", c_sz, [c_i64, c_i, c_i, c_i, c_i])
    _sig(L, "evad_graph_build", c_i, [c_p, c_p, c_i64, c_i, c_i, c_i, c_i, c_i, c_i, c_i,
                                      c_p, c_p, c_p, c_p, c_i64, c_p, c_p, c_p, c_p, c_p, c_sz, c_p])
    _sig(L, "evad_graph_build_rows", c_i, [c_p, c_p, c_i64, c_i, c_i, c_i, c_i, c_i, c_i, c_i,
                                           c_p, c_p, c_p, c_p, c_p, c_p, c_p, c_sz, c_p])
    _sig(L, "evad_graph_set_span_limit", c_i, [c_i])
    _sig(L, "evad_stream_state_bytes", c_sz, [c_i, c_i, c_i, c_i])
    _sig(L, "evad_stream_reset", c_i, [c_p, c_i, c_i, c_i, c_i, c_p])
    _sig(L, "evad_stream_ws_bytes", c_sz, [c_i64, c_i])
    _sig(L, "evad_stream_append", c_i, [c_p, c_i, c_i, c_i, c_i, c_p, c_p, c_i64, c_i, c_i, c_i, c_i, c_i, c_i,
                                        c_p, c_p, c_p, c_p, c_i64, c_p, c_p, c_sz, c_p])
    P = ctypes.POINTER
    _sig(L, "evad_spline_lut_build", c_i, [c_p, c_p, c_p, c_i, c_i, c_i, P(SplineModeT), c_p, c_p])
    _sig(L, "evad_spline_conv_small", c_i, [c_p, c_p, c_i64, c_p, c_i, c_i, c_i, c_i, P(PseudoT), P(SplineModeT), c_p,
                                            P(EpilogueT), c_p, c_i, P(EpilogueT), c_p, c_i, c_p, c_p])
    _sig(L, "evad_spline_aggregate", c_i, [c_p, c_p, c_i64, c_p, c_p, c_i, c_i, P(PseudoT), P(SplineModeT), c_p, c_i64,
                                           c_p, c_p])
    _sig(L, "evad_gemm_ws_bytes", c_sz, [c_i64, c_i, c_i])
    _sig(L, "evad_gemm_epilogue", c_i, [c_p, c_i64, c_p, c_i64, c_p, c_i, c_i, P(EpilogueT), c_p, c_i, c_p, c_sz, c_p])
    _sig(L, "evad_edge_class", c_i, [c_p, c_p, c_i64, P(PseudoT), P(SplineModeT), c_p, c_p, c_p])
    _sig(L, "evad_spline_conv_l1", c_i, [c_p, c_p, c_p, c_p, c_i, c_i64, c_p, c_i, c_i, c_p, c_i, c_p, c_p, c_p, c_p, P(EpilogueT),
                                         c_p, c_i, P(EpilogueT), c_p, c_i, c_p])
    _sig(L, "evad_spline_conv_l1_mma", c_i, [c_p, c_p, c_p, c_p, c_i64, c_p, c_i, c_i, c_p, c_i, c_p, c_p, c_p, c_p, P(EpilogueT),
                                             c_p, c_i, P(EpilogueT), c_p, c_i, c_p])
    _sig(L, "evad_spline_conv_l1_tc5", c_i, [c_p, c_p, c_p, c_p, c_i64, c_p, c_i, c_i, c_p, c_i, c_p, c_f, c_p, c_f, P(EpilogueT),
                                             c_p, c_i, c_i, P(EpilogueT), c_p, c_i, c_p, c_p])
    _sig(L, "evad_split_h2", c_i, [c_p, c_p, c_i64, c_p, c_p])
    _sig(L, "evad_stream_rows", c_i, [c_p, c_p, c_i64, c_i, c_i, c_p, c_i, c_i, c_p, c_p, c_p, c_p, c_p])
    _sig(L, "evad_spline_conv_l1_h16", c_i, [c_p, c_p, c_p, c_p, c_i64, c_i64, c_p, c_i, c_i, c_p, c_i, c_p, c_p, c_p, c_p, P(EpilogueT),
                                             c_p, c_i, c_i, P(EpilogueT), c_p, c_i, c_p, c_p])
    _sig(L, "evad_spline_tap_lists", c_i, [c_p, c_p, c_i64, c_p, P(PseudoT), P(SplineModeT), c_p, c_p, c_p, c_p])
    _sig(L, "evad_spline_conv64_fused_ws_bytes", c_sz, [c_i64, c_i])
    _sig(L, "evad_spline_conv64_fused", c_i, [c_p, c_p, c_p, c_i, c_i64, c_p, c_i, c_i, c_p, c_p, c_i64, P(EpilogueT),
                                              c_p, c_i, c_p, c_sz, c_p])
    _sig(L, "evad_split_tf32", c_i, [c_p, c_i64, c_p, c_p, c_p])
    _sig(L, "evad_pool_inc_state_bytes", c_sz, [P(PoolParamsT), c_i])
    _sig(L, "evad_pool_inc_reset", c_i, [c_p, P(PoolParamsT), c_i, c_p])
    _sig(L, "evad_pool_inc_update_ws_bytes", c_sz, [c_i64])
    _sig(L, "evad_pool_inc_update", c_i, [c_p, P(PoolParamsT), c_i64, c_i64, c_p, c_i, c_p, c_p, c_i, c_i, c_p, c_p, c_p, c_p,
                                          c_p, c_sz, c_p])
    _sig(L, "evad_pool_inc_emit_ws_bytes", c_sz, [P(PoolParamsT), c_i64])
    _sig(L, "evad_pool_inc_emit", c_i, [c_p, P(PoolParamsT), c_i, c_i64, c_i64, c_p, c_p, c_i, c_p, c_p, c_p, c_p, c_p, c_p,
                                        c_p, c_sz, c_p])
    _sig(L, "evad_ingest_ms_to_idx", c_i, [c_p, c_i64, c_i64, c_p, c_p])
    _sig(L, "evad_ingest_slice_ws_bytes", c_sz, [c_i64])
    _sig(L, "evad_ingest_slice", c_i, [c_p, c_p, c_p, c_p, c_i64, c_i64, c_i64, c_i64, c_i, c_p, c_p, c_p, c_p, c_p, c_sz, c_p])
    _sig(L, "evad_ingest_downsample_ws_bytes", c_sz, [c_i64])
    _sig(L, "evad_ingest_downsample", c_i, [c_p, c_p, c_p, c_p, c_i64, c_i, c_i, c_i, c_i, c_p, c_p, c_p, c_p, c_p, c_p, c_p,
                                            c_p, c_sz, c_p])
    _sig(L, "evad_gemm64_h2_a_bytes", c_sz, [c_i64, c_i])
    _sig(L, "evad_gemm64_h2_w_bytes", c_sz, [c_i])
    _sig(L, "evad_gemm64_h2_ws_bytes", c_sz, [c_i64, c_i])
    _sig(L, "evad_spline_aggregate_h2", c_i, [c_p, c_p, c_i64, c_p, c_p, c_i, c_i, P(PseudoT), P(SplineModeT), c_p, c_p, c_p])
    _sig(L, "evad_gemm64_h2", c_i, [c_p, c_p, c_i64, c_p, c_i, P(EpilogueT), c_p, c_i, c_p, c_sz, c_p])
    _sig(L, "evad_gemm64_tc_ws_bytes", c_sz, [c_i64, c_i])
    _sig(L, "evad_gemm64_tc", c_i, [c_p, c_i64, c_p, c_i64, c_p, c_i, P(EpilogueT), c_p, c_i, c_p, c_sz, c_p])
    _sig(L, "evad_gemm64_tc5_ws_bytes", c_sz, [c_i64, c_i])
    _sig(L, "evad_gemm64_tc5", c_i, [c_p, c_i64, c_p, c_i64, c_i64, c_p, c_i, P(EpilogueT), c_p, c_i, c_p, c_sz, c_p])
    _sig(L, "evad_voxel_pool_ws_bytes", c_sz, [c_i64, c_i64])
    _sig(L, "evad_voxel_pool", c_i, [P(PoolParamsT), c_i64, c_p, c_p, c_i, c_p, c_p, c_i, c_i, c_p, c_p, c_i64, c_i64,
                                     c_p, c_p, c_p, c_i, c_p, c_p, c_p, c_p, c_p, c_p, c_p, c_sz, c_p])
    _sig(L, "evad_coo_to_csr_ws_bytes", c_sz, [c_i64])
    _sig(L, "evad_coo_to_csr", c_i, [c_p, c_i64, c_i64, c_i64, c_p, c_p, c_p, c_p, c_p, c_sz, c_p])
    _sig(L, "evad_csr_to_coo_ws_bytes", c_sz, [c_i64])
    _sig(L, "evad_csr_to_coo", c_i, [c_p, c_p, c_i64, c_p, c_i64, c_p, c_i, c_p, c_i64, c_p, c_p, c_sz, c_p])
    _sig(L, "evad_cartesian", c_i, [c_p, c_i, c_i, c_p, c_i64, c_i64, c_p, c_f, c_i, c_p, c_p])
    _sig(L, "evad_format_events", c_i, [c_p, c_p, c_p, c_p, c_i64, c_i, c_i, c_i, c_p, c_p, c_p, c_p])
    _sig(L, "evad_denormalize_pos", c_i, [c_p, c_i, c_i64, c_i, c_i, c_i, c_p, c_p])
    _sig(L, "evad_asy_masked_lin", c_i, [c_p, c_i64, c_p, c_p, c_p, c_p, c_i, c_i, c_i, c_p])
    _sig(L, "evad_asy_masked_bn", c_i, [c_p, c_i64, c_p, c_p, c_p, c_p, c_p, c_p, c_i, c_f, c_p])
    _sig(L, "evad_asy_masked_isdiff", c_i, [c_p, c_i64, c_p, c_p, c_i, c_f, c_f, c_p])
    _sig(L, "evad_bbox_pool", c_i, [c_p, c_i, c_i, c_p, c_i, c_p, c_i64, c_p, c_p, c_i, c_p, c_i, c_i, c_p, c_p, c_i, c_i, c_p, c_p])
    _sig(L, "evad_sample_concat_h2", c_i, [c_p, c_i, c_i, c_p, c_i, c_p, c_i64, c_p, c_p, c_i, c_i, c_i, c_i, c_f, c_f,
                                           c_i, c_p, c_i, c_p, c_p, c_p])
    _sig(L, "evad_sample_concat", c_i, [c_p, c_i, c_i, c_p, c_i, c_p, c_i64, c_p, c_p, c_i, c_i, c_i, c_i, c_f, c_f,
                                        c_i, c_p, c_i, c_p])


def check(rc: int, what: str = ""):
    if rc != 0:
        raise RuntimeError(f"eventad_b200: {what} failed: {ERRORS.get(rc, rc)}")


def ptr(t):
    """device pointer of a tensor (None -> NULL); the tensor must be a contiguous CUDA tensor."""
    if t is None:
        return None
    if not t.is_cuda:
        raise RuntimeError("eventad_b200: tensor must be a CUDA tensor (no CPU fallback)")
    if not t.is_contiguous():
        raise RuntimeError("eventad_b200: tensor must be contiguous")
    return t.data_ptr()


def ptr_rows(t):
    """device pointer of a 2-D tensor whose rows are contiguous (stride(1) == 1); the row stride is passed separately."""
    if t is None:
        return None
    if not t.is_cuda:
        raise RuntimeError("eventad_b200: tensor must be a CUDA tensor (no CPU fallback)")
    if t.dim() != 2 or (t.shape[1] > 1 and t.stride(1) != 1):
        raise RuntimeError("eventad_b200: tensor rows must be contiguous")
    return t.data_ptr()


def stream():
    return torch.cuda.current_stream().cuda_stream


def check_status(status: torch.Tensor, what: str = ""):
    """Reads the device status word (synchronises) and raises on any set bit.  The word is cleared before raising, so
    one bad batch does not poison the calls that follow (kernels only ever OR bits into it)."""
    s = int(status.item())
    if s:
        status.zero_()
        msgs = [m for b, m in STATUS_BITS.items() if s & b]
        raise RuntimeError(f"eventad_b200: {what}: " + "; ".join(msgs))


class Workspace:
    """Grow-only scratch buffer per device (torch-owned memory; the library never allocates).  A buffer handed out while
    a CUDA graph was being captured is baked into that graph: it is kept alive in `pinned` and never handed to anyone
    else's growth path (`Scratch` gives every captured graph its own Workspace objects in the first place)."""

    def __init__(self):
        self.buf = None
        self.pinned = []

    def get(self, nbytes: int, device) -> torch.Tensor:
        if self.buf is None or self.buf.numel() < nbytes or self.buf.device != torch.device(device):
            self.buf = torch.empty(int(nbytes * 1.25) + 1024, dtype=torch.uint8, device=device)
        if torch.cuda.is_available() and torch.cuda.is_current_stream_capturing() and \
                not any(b is self.buf for b in self.pinned):
            self.pinned.append(self.buf)                     # a replaced buffer stays alive as long as the graph may replay
        return self.buf


class Scratch:
    """Private scratch of one owner (e.g. one captured CUDA graph): the graph builder's workspace, the per-op workspaces
    and the device status word.  engine.GraphedForward holds one, so that neither a later, larger eager call nor a
    second captured graph on the same Net can free or share memory whose address a graph has baked in."""

    def __init__(self):
        self.graph_ws = Workspace()
        self.ops_ws = {}
        self._status = None

    def status(self, device):
        if self._status is None or self._status.device != torch.device(device):
            self._status = torch.zeros(1, dtype=torch.int32, device=device)
        return self._status
