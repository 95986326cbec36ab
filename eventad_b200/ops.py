"""Tensor-level wrappers of the C ABI (include/eventad_b200.h).  Every function only enqueues kernels on the current
torch CUDA stream; sizes that depend on the data stay on the device (int32 cells) unless a caller asks for a tight
tensor.  No function here has a CPU or PyTorch fallback."""
import ctypes
import math
import os

import torch

from . import _lib
from ._lib import EpilogueT, PoolParamsT, PseudoT, SplineModeT, check, ptr, ptr_rows, stream

_WS = {}
_SCRATCH = None     # a _lib.Scratch while its owner (a captured CUDA graph) enqueues work: private workspaces


def _ws(nbytes, device):
    key = (torch.device(device).index, torch.cuda.current_stream().cuda_stream)
    table = _SCRATCH.ops_ws if _SCRATCH is not None else _WS
    if _SCRATCH is not None:
        key = key[0]
    w = table.get(key)
    if w is None:
        w = table[key] = _lib.Workspace()
    return w.get(nbytes, device)


class KernelTimer:
    """Optional CUDA-event timing of the C-ABI calls, on the launching stream (bench.py's per-stage breakdown and the
    roofline of the dominant kernel).  Disabled (None) by default."""

    def __init__(self):
        self.pairs = {}

    def record(self, name):
        ev = (torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True))
        self.pairs.setdefault(name, []).append(ev)
        ev[0].record()
        return ev[1]

    def summary(self):
        """name -> (launch groups, total ms); call after a synchronize."""
        return {k: (len(v), sum(a.elapsed_time(b) for a, b in v)) for k, v in self.pairs.items()}


TIMER = None


class _timed:
    def __init__(self, name):
        self.name = name
        self.stop = None

    def __enter__(self):
        if TIMER is not None:
            self.stop = TIMER.record(self.name)

    def __exit__(self, *a):
        if self.stop is not None:
            self.stop.record()


def new_status(device):
    return torch.zeros(1, dtype=torch.int32, device=device)


# --------------------------------------------------------------------------------------------------- graph formats
def coo_to_csr(edge_index: torch.Tensor, n_nodes: int, status=None):
    """edge_index int64 [2,E] (any order) -> (rowptr int32[n+1], src int32[E], eid int32[E]) dst-major, stable."""
    L = _lib.lib()
    dev = edge_index.device
    e = edge_index.shape[1]
    ei = edge_index.contiguous()
    if ei.dtype != torch.int64:
        ei = ei.long()
    rowptr = torch.empty(n_nodes + 1, dtype=torch.int32, device=dev)
    src = torch.empty(max(e, 1), dtype=torch.int32, device=dev)
    eid = torch.empty(max(e, 1), dtype=torch.int32, device=dev)
    status = status if status is not None else new_status(dev)
    ws = _ws(L.evad_coo_to_csr_ws_bytes(e), dev)
    check(L.evad_coo_to_csr(ptr(ei) if e > 0 else None, ei.stride(0) if e > 0 else 0, e, n_nodes, ptr(rowptr), ptr(src),
                            ptr(eid), ptr(status), ptr(ws), ws.numel(), stream()), "evad_coo_to_csr")
    return rowptr, src[:e], eid[:e]


def csr_to_coo(rowptr, src, n_cap, n_dev, e_cap, num_edges, src_major: bool, want_slots=False):
    """-> edge_index int64 [2, e_cap] (first num_edges columns valid) and optionally slot_of int32[e_cap]."""
    L = _lib.lib()
    dev = rowptr.device
    ei = torch.empty((2, max(e_cap, 1)), dtype=torch.int64, device=dev)
    slot_of = torch.empty(max(e_cap, 1), dtype=torch.int32, device=dev) if want_slots else None
    ws = _ws(L.evad_csr_to_coo_ws_bytes(e_cap), dev)
    check(L.evad_csr_to_coo(ptr(rowptr), ptr(src), n_cap, ptr(n_dev), e_cap, ptr(num_edges), int(src_major), ptr(ei),
                            ei.stride(0), ptr(slot_of), ptr(ws), ws.numel(), stream()), "evad_csr_to_coo")
    return (ei, slot_of) if want_slots else ei


def cartesian(pos, edge_index, two_m: float, clamp01: bool, num_edges=None, dims=None):
    """PyG T.Cartesian(norm=True, cat=False, max_value=two_m/2) (+clamp): [E, dims] float32."""
    L = _lib.lib()
    e_cap = edge_index.shape[1]
    dims = dims or pos.shape[1]
    out = torch.empty((e_cap, dims), dtype=torch.float32, device=pos.device)
    if e_cap == 0:
        return out
    pos = pos.contiguous()
    check(L.evad_cartesian(ptr(pos), pos.stride(0), dims, ptr_rows(edge_index), edge_index.stride(0), e_cap, ptr(num_edges),
                           float(two_m), int(clamp01), ptr(out), stream()), "evad_cartesian")
    return out


def format_events(pos_xy, t, polarity, batch64, width, height, time_window, out=None):
    """utils/buffers.py:33-44 (format_data) in one launch: raw int16 / int32 / int8 / int64 event arrays on the device ->
    (x float32 [n,1], pos float32 [n,3] normalised, batch int32 [n]); `out` = preallocated (x, pos, batch)."""
    L = _lib.lib()
    n = t.shape[0]
    dev = t.device
    assert pos_xy.dtype == torch.int16 and t.dtype == torch.int32 and polarity.dtype == torch.int8
    assert batch64 is None or batch64.dtype == torch.int64
    pos_xy, t, polarity = pos_xy.contiguous(), t.contiguous(), polarity.contiguous()
    if out is None:
        out = (torch.empty((n, 1), dtype=torch.float32, device=dev), torch.empty((n, 3), dtype=torch.float32, device=dev),
               torch.empty(n, dtype=torch.int32, device=dev))
    x, pos, batch = out
    assert x.is_contiguous() and pos.is_contiguous() and batch.is_contiguous() and x.numel() >= n and pos.shape[0] >= n
    with _timed('format_events'):
        check(L.evad_format_events(ptr(pos_xy), ptr(t), ptr(polarity), ptr(batch64.contiguous()) if batch64 is not None else None,
                                   n, int(width), int(height), int(time_window), ptr(x), ptr(pos), ptr(batch), stream()),
              "evad_format_events")
    return out


def denormalize_pos(pos, width, height, time_window):
    """ev_tgn.py:11-16 in one launch: (pos * [W, H, T] + 1e-3).int()."""
    L = _lib.lib()
    pos = pos.contiguous()
    out = torch.empty((pos.shape[0], 3), dtype=torch.int32, device=pos.device)
    check(L.evad_denormalize_pos(ptr(pos), pos.stride(0), pos.shape[0], int(width), int(height), int(time_window), ptr(out),
                                 stream()), "evad_denormalize_pos")
    return out


def to_channel_last(feat: torch.Tensor) -> torch.Tensor:
    """[B,C,h,w] -> contiguous [B,h,w,C] (the layout the sampling kernels read)."""
    return feat.permute(0, 2, 3, 1).contiguous()


def sample_concat(x, pos, batch_i32, feat_nhwc, width, height, append_pos=False, ldo=None, n_dev=None):
    """[x | grid_sample(feat) | pos_xy | 0-pad] -> [n, ldo] (net.py sampling_skip + cat(x, pos[:, :2]))."""
    L = _lib.lib()
    n = pos.shape[0]
    cx = 0 if x is None else x.shape[1]
    cs = 0 if feat_nhwc is None else feat_nhwc.shape[3]
    width_total = cx + cs + (2 if append_pos else 0)
    ldo = ldo or width_total
    out = torch.empty((n, ldo), dtype=torch.float32, device=pos.device)
    if n == 0:
        return out
    if x is not None:
        x = x.contiguous()
    pos = pos.contiguous()
    fb, fh, fw, fc = (feat_nhwc.shape if feat_nhwc is not None else (0, 0, 0, 0))
    with _timed('sample_concat'):
        check(L.evad_sample_concat(ptr(x), x.stride(0) if x is not None else 0, cx, ptr(pos), pos.stride(0), ptr(batch_i32), n,
                                   ptr(n_dev), ptr(feat_nhwc), fb, fh, fw, fc, float(width), float(height), int(append_pos),
                                   ptr(out), ldo, stream()), "evad_sample_concat")
    return out


def sample_concat_h2(x, pos, batch_i32, feat_nhwc, width, height, status, visit=None):
    """sample_concat for the layer-1 row shape ([polarity | 16 sampled channels | pos_xy | 0], rows of 20) written as h2
    words (split_h2 fused into the producer); None when the shape is another one."""
    if (x is None or feat_nhwc is None or x.shape[1] != 1 or feat_nhwc.shape[3] != 16 or pos.shape[0] == 0 or
            feat_nhwc.data_ptr() % 16 != 0):
        return None
    n = pos.shape[0]
    out = torch.empty((n, 20), dtype=torch.float32, device=pos.device)
    x, pos = x.contiguous(), pos.contiguous()
    fb, fh, fw, fc = feat_nhwc.shape
    with _timed('sample_concat'):
        check(_lib.lib().evad_sample_concat_h2(ptr(x), x.stride(0), 1, ptr(pos), pos.stride(0), ptr(batch_i32), n, None,
                                               ptr(feat_nhwc), fb, fh, fw, fc, float(width), float(height), 1, ptr(out), 20,
                                               ptr(visit), ptr(status), stream()), "evad_sample_concat_h2")
    return out


# --------------------------------------------------------------------------------------------------- spline tables
def spline_mode_basis(kernel_size: int):
    return dict(use_lut=0, ks=int(kernel_size), nx=0, ny=0, r=(0.0, 0.0, 0.0, 0.0), tab_k=None, tab_w=None)


def _table_1d(r, M, extent, ks):
    """Per-lattice-coordinate taps of init_lut (/root/reference/src/dagr/model/layers/spline_conv.py:26-33) for one
    dimension: a = d / (2*M*extent) + 0.5 on d = -r..r, then the degree-1 open-spline basis of torch_spline_conv."""
    d = torch.arange(-r, r + 1).float()
    a = d / (2 * M * extent) + 0.5
    v = a * float(ks - 1)
    vi = v.to(torch.long)
    k0 = torch.fmod(vi + 0, ks)
    k1 = torch.fmod(vi + 1, ks)
    f = v - torch.floor(v)
    w0 = 1 - f - 0 + 2 * f * 0
    w1 = 1 - f - 1 + 2 * f * 1
    return torch.stack([k0, k1], 1).int(), torch.stack([w0, w1], 1).float()


def spline_mode_lut(height, width, rx, Mx, ry=None, My=None, kernel_size=5, device="cuda"):
    """message_lut parameters: attr_remapping_matrix (spline_conv.py:21-24) + per-dimension tap tables."""
    ry = ry or rx
    My = My or Mx
    R = torch.Tensor([[2 * Mx * width, 0, -Mx * width + rx], [0, 2 * My * height, -My * height + ry]])
    kx, wx = _table_1d(rx, Mx, width, kernel_size)
    ky, wy = _table_1d(ry, My, height, kernel_size)
    return dict(use_lut=1, ks=int(kernel_size), nx=2 * rx + 1, ny=2 * ry + 1,
                r=(float(R[0, 0]), float(R[0, 2]), float(R[1, 1]), float(R[1, 2])),
                tab_k=torch.cat([kx, ky], 0).contiguous().to(device), tab_w=torch.cat([wx, wy], 0).contiguous().to(device),
                R=R)


def _mode_struct(mode):
    return SplineModeT(use_lut=mode["use_lut"], r00=mode["r"][0], r02=mode["r"][1], r11=mode["r"][2], r12=mode["r"][3],
                       nx=mode["nx"], ny=mode["ny"], tab_k=ptr(mode["tab_k"]), tab_w=ptr(mode["tab_w"]), ks=mode["ks"])


def _pseudo_struct(edge_attr=None, eid=None, pos=None, two_m=1.0, clamp01=False):
    return PseudoT(edge_attr=ptr(edge_attr), eid=ptr(eid), ld_attr=edge_attr.stride(0) if edge_attr is not None else 0,
                   pos=ptr(pos), ld_pos=pos.stride(0) if pos is not None else 0, two_m=float(two_m), clamp01=int(clamp01))


def _epi_struct(scale=None, shift=None, addend=None, relu=False):
    return EpilogueT(scale=ptr(scale), shift=ptr(shift), addend=ptr(addend),
                     ld_add=addend.stride(0) if addend is not None else 0, relu=int(relu))


def fold_bn(bn: torch.nn.BatchNorm1d, conv_bias=None):
    """eval BatchNorm (after an optional conv bias) as y = x*scale + shift; cached on the module until a tensor changes."""
    key = (bn.weight._version, bn.bias._version, bn.running_mean._version, bn.running_var._version, bn.weight.device,
           None if conv_bias is None else conv_bias._version)
    c = getattr(bn, "_evad_fold", None)
    if c is not None and c[0] == key:
        return c[1], c[2]
    with torch.no_grad():
        scale = (bn.weight / torch.sqrt(bn.running_var + bn.eps)).float().contiguous()
        shift = bn.bias - bn.running_mean * scale
        if conv_bias is not None:
            shift = shift + conv_bias.float() * scale
        shift = shift.float().contiguous()
    bn._evad_fold = (key, scale, shift)
    return scale, shift


# --------------------------------------------------------------------------------------------------- SplineConv
def build_lut(weight, root_w, skip_w, cin_pad, mode):
    L = _lib.lib()
    k, cin, cout = weight.shape
    n = mode["nx"] * mode["ny"] + 2
    lut = torch.empty((n, cin_pad, cout), dtype=torch.float32, device=weight.device)
    m = _mode_struct(mode)
    check(L.evad_spline_lut_build(ptr(weight.detach().contiguous()), ptr(root_w.detach().contiguous()),
                                  ptr(skip_w.detach().contiguous()) if skip_w is not None else None, cin, cin_pad, cout,
                                  ctypes.byref(m), ptr(lut), stream()), "evad_spline_lut_build")
    return lut


def spline_conv_small(rowptr, src, n, x, cin, lut, mode, pseudo, epi, status, skip_epi=None, out=None, skip_out=None):
    """Fused LUT conv, cout = 16.  x: [n, ldx] with ldx % 4 == 0, zero padded beyond cin.  Returns (out, skip_out)."""
    L = _lib.lib()
    cin_pad = lut.shape[1]
    cout = lut.shape[2]
    if out is None:
        out = torch.empty((n, cout), dtype=torch.float32, device=x.device)
    if skip_epi is not None and skip_out is None:
        skip_out = torch.empty((n, cout), dtype=torch.float32, device=x.device)
    m = _mode_struct(mode)
    p = _pseudo_struct(**pseudo)
    e = _epi_struct(**epi)
    se = _epi_struct(**skip_epi) if skip_epi is not None else None
    with _timed(f'spline_conv_small_c{cin}'):
        check(L.evad_spline_conv_small(ptr(rowptr), ptr(src), n, ptr(x), x.stride(0), cin, cin_pad, cout, ctypes.byref(p),
                                       ctypes.byref(m), ptr(lut), ctypes.byref(e), ptr(out), out.stride(0),
                                       ctypes.byref(se) if se is not None else None, ptr(skip_out),
                                       skip_out.stride(0) if skip_out is not None else 0, ptr(status), stream()),
              "evad_spline_conv_small")
    return out, skip_out


def offset_class_is_lut_class(mode, two_m, clamp01, width, height):
    """True when, for every pixel and every integer offset d in [-r, r], the reference's fp32 LUT index
    (T.Cartesian attr -> attr_remapping_matrix -> +1e-3 -> long, spline_conv.py:39-42) equals d + r, i.e. when the
    spiral-offset class bytes of evad_graph_build ARE the LUT classes.  Exhaustive over the lattice, same fp32 ops as
    the kernels (positions are pixel / extent in fp32, buffers.py:33-44)."""
    nx, ny = mode["nx"], mode["ny"]
    r00, r02, r11, r12 = mode["r"]
    tm = torch.tensor(float(two_m), dtype=torch.float32)
    for ext, n1, coef, off in ((int(width), nx, r00, r02), (int(height), ny, r11, r12)):
        r = (n1 - 1) // 2
        pix = torch.arange(ext)
        d = torch.arange(-r, r + 1)
        pj_idx = pix[:, None] + d[None, :]
        valid = (pj_idx >= 0) & (pj_idx < ext)
        pi = pix.float() / torch.tensor(float(ext))
        pj = pj_idx.clamp(0, ext - 1).float() / torch.tensor(float(ext))
        a = (pj - pi[:, None]) * (torch.tensor(1.0) / tm) + torch.tensor(0.5)     # CUDA: multiply by the reciprocal
        if clamp01:
            a = a.clamp(0, 1)
        idx = (a * torch.tensor(coef, dtype=torch.float32) + torch.tensor(off, dtype=torch.float32) +
               torch.tensor(1e-3, dtype=torch.float32)).long()
        if not bool((idx == (d + r)[None, :])[valid].all()):
            return False
    return True


def edge_class(rowptr, src, n, mode, pseudo, status):
    """uint8 LUT class (ix*ny + iy) of every CSR slot (spline_conv.py:39-42)."""
    L = _lib.lib()
    cls = torch.empty(max(src.shape[0], 1), dtype=torch.uint8, device=src.device)
    m = _mode_struct(mode)
    p = _pseudo_struct(**pseudo)
    with _timed('edge_class'):
        check(L.evad_edge_class(ptr(rowptr), ptr(src), n, ctypes.byref(p), ctypes.byref(m), ptr(cls), ptr(status), stream()),
              "evad_edge_class")
    return cls


# EVAD_L1_MMA=0: keep the tap aggregation of layer 1 on the FFMA pipe (evad_spline_conv_l1) for A/B measurements
USE_L1_MMA = os.environ.get("EVAD_L1_MMA", "1") != "0"
# EVAD_L1_TC5=1: with the fp16-pair kernel (EVAD_L1_H16), contraction of layer 1 on tcgen05 (csrc/spline_l1_tc5.cu)
# instead of mma.sync (A/B measurements)
USE_L1_TC5 = os.environ.get("EVAD_L1_TC5", "0") == "1"


# EVAD_L1_H16=0: keep layer 1 on the 3xTF32 mma.sync kernel (evad_spline_conv_l1_mma) instead of the fp16-pair kernel
USE_L1_H16 = os.environ.get("EVAD_L1_H16", "1") != "0"


def split_h2(x, status=None):
    """fp32 rows -> "h2 words" (fp16 leading part | fp16 remainder * 2048, include/eventad_b200.h): the operand format of
    evad_spline_conv_l1_h16.  Returned with dtype float32 (same bytes); |x| >= 2048 raises EVAD_STATUS_H16_RANGE."""
    assert x.is_contiguous() and x.numel() % 4 == 0
    out = torch.empty_like(x)
    with _timed('split_h2'):
        check(_lib.lib().evad_split_h2(ptr(x), ptr(out), x.numel(), ptr(status), stream()), "evad_split_h2")
    return out


def l1_kmap(cin_pad):
    """K index of the tcgen05 layer-1 contraction -> (slot, channel); the operand order of csrc/spline_l1_tc5.cu."""
    nch = 5 if cin_pad > 16 else 4
    k = torch.arange(nch * 64)
    u, h, c = k >> 3, (k >> 2) & 1, k & 3
    slot = (u >> 2) + 8 * h
    ch = 4 * (u & 3) + c
    q = k - 256
    slot2 = (q >> 3) + 8 * ((q >> 1) & 1)
    ch2 = 16 + 2 * ((q >> 2) & 1) + (q & 1)
    hi = k >= 256
    return torch.where(hi, slot2, slot), torch.where(hi, ch2, ch)


def split_f16x2(w):
    """w fp32 -> (w1, w2, inv_scale): w * 2^kw = w1 + w2 * 2^-11 with fp16 parts; the power-of-two scale puts max|w| in
    [2^9, 2^10)."""
    w = w.detach().float()
    m = float(w.abs().max())
    kw = 9 - math.floor(math.log2(m)) if m > 0 else 0
    kw = max(min(kw, 100), -100)
    ws = w * (2.0 ** kw)
    w1 = ws.half()
    w2 = ((ws - w1.float()) * 2048.0).half()
    return w1, w2, 2.0 ** (-kw)


def l1_pack16(W, S, cin_pad):
    """fp16 operands of evad_spline_conv_l1_tc5.  W: [16 * cin_pad, 16] fp32 (row = slot * cin_pad + channel, slot 15 =
    root weight); S: [>= cin_pad, 16] skip Linear weight^T (zero padded) or None."""
    slot, ch = l1_kmap(cin_pad)
    Wk = W.detach().float().cpu()[slot * cin_pad + ch]                    # [K, 16]
    w1, w2, inv_sw = split_f16x2(Wk)
    pack = dict(w16=torch.stack([w1.t(), w2.t()]).contiguous().view(torch.int16), inv_sw=inv_sw, s16=None, inv_ss=1.0)
    if S is not None:
        kk = torch.arange(192, 192 + 128)
        valid = kk < slot.shape[0]
        sl = torch.where(valid, slot[kk.clamp(max=slot.shape[0] - 1)], torch.full_like(kk, -1))
        cc = ch[kk.clamp(max=slot.shape[0] - 1)]
        Sk = torch.zeros(128, 16)
        root = sl == 15
        Sk[root] = S.detach().float().cpu()[cc[root]]
        s1, s2, inv_ss = split_f16x2(Sk)
        pack["s16"], pack["inv_ss"] = torch.stack([s1.t(), s2.t()]).contiguous().view(torch.int16), inv_ss
    return pack


def l1_h16_ok(src, cls, ell_stride):
    """True when spline_conv_l1 will take the fp16-pair kernel (fixed-stride rows of 16, aligned)."""
    return (USE_L1_MMA and USE_L1_H16 and ell_stride == 16 and cls is not None and
            src.data_ptr() % 16 == 0 and cls.data_ptr() % 16 == 0)


def spline_conv_l1(rowptr, src, cls, n, x, cin, pack, epi, skip_epi=None, perm=None, ell_stride=0, x_h2=False,
                   out_h2=False, status=None):
    """Tap aggregation + tensor-core contraction, cout = 16.  pack = dict(tapw, whi, wlo, shi, slo, cin_pad).
    ell_stride > 0: fixed-stride rows (graph build_rows): rowptr = row lengths [n], src / cls rows at i * ell_stride.
    x_h2 / out_h2 (fp16-pair kernel only, `l1_h16_ok`): x already holds h2 words / the output is wanted as h2 words."""
    L = _lib.lib()
    out = torch.empty((n, 16), dtype=torch.float32, device=x.device)
    has_skip = skip_epi is not None and pack.get("shi") is not None
    skip_out = torch.empty((n, 16), dtype=torch.float32, device=x.device) if has_skip else None
    e = _epi_struct(**epi)
    se = _epi_struct(**skip_epi) if has_skip else None
    if l1_h16_ok(src, cls, ell_stride):
        if not x_h2:
            x = split_h2(x.contiguous(), status)
        if USE_L1_TC5 and pack.get("w16") is not None:
            with _timed(f'spline_conv_l1_c{cin}'):
                check(L.evad_spline_conv_l1_tc5(ptr(rowptr), ptr(src), ptr(cls), ptr(perm), n, ptr(x), x.stride(0), pack["cin_pad"],
                                                ptr(pack["tapw"]), pack["tapw"].shape[0], ptr(pack["w16"]), float(pack["inv_sw"]),
                                                ptr(pack["s16"]) if has_skip else None, float(pack["inv_ss"]),
                                                ctypes.byref(e), ptr(out), out.stride(0), int(out_h2),
                                                ctypes.byref(se) if has_skip else None, ptr(skip_out),
                                                skip_out.stride(0) if has_skip else 0, ptr(status), stream()),
                      "evad_spline_conv_l1_tc5")
            return out, skip_out
        with _timed(f'spline_conv_l1_c{cin}'):
            check(L.evad_spline_conv_l1_h16(ptr(rowptr), ptr(src), ptr(cls), ptr(perm), n, n, ptr(x), x.stride(0), pack["cin_pad"],
                                            ptr(pack["tapw"]), pack["tapw"].shape[0], ptr(pack["whi"]), ptr(pack["wlo"]),
                                            ptr(pack["shi"]) if has_skip else None, ptr(pack["slo"]) if has_skip else None,
                                            ctypes.byref(e), ptr(out), out.stride(0), int(out_h2),
                                            ctypes.byref(se) if has_skip else None, ptr(skip_out),
                                            skip_out.stride(0) if has_skip else 0, ptr(status), stream()),
                  "evad_spline_conv_l1_h16")
        return out, skip_out
    assert not x_h2 and not out_h2, "h2 words are only understood by evad_spline_conv_l1_h16"
    if (USE_L1_MMA and ell_stride == 16 and src.data_ptr() % 16 == 0 and
            cls.data_ptr() % 16 == 0):
        with _timed(f'spline_conv_l1_c{cin}'):
            check(L.evad_spline_conv_l1_mma(ptr(rowptr), ptr(src), ptr(cls), ptr(perm), n, ptr(x), x.stride(0), pack["cin_pad"],
                                            ptr(pack["tapw"]), pack["tapw"].shape[0], ptr(pack["whi"]), ptr(pack["wlo"]),
                                            ptr(pack["shi"]) if has_skip else None, ptr(pack["slo"]) if has_skip else None,
                                            ctypes.byref(e), ptr(out), out.stride(0), ctypes.byref(se) if has_skip else None,
                                            ptr(skip_out), skip_out.stride(0) if has_skip else 0, stream()),
                  "evad_spline_conv_l1_mma")
        return out, skip_out
    with _timed(f'spline_conv_l1_c{cin}'):
        check(L.evad_spline_conv_l1(ptr(rowptr), ptr(src), ptr(cls), ptr(perm), int(ell_stride), n, ptr(x), x.stride(0), pack["cin_pad"],
                                    ptr(pack["tapw"]), pack["tapw"].shape[0], ptr(pack["whi"]), ptr(pack["wlo"]),
                                    ptr(pack["shi"]) if has_skip else None, ptr(pack["slo"]) if has_skip else None,
                                    ctypes.byref(e), ptr(out), out.stride(0), ctypes.byref(se) if has_skip else None,
                                    ptr(skip_out), skip_out.stride(0) if has_skip else 0, stream()),
              "evad_spline_conv_l1")
    return out, skip_out


def spline_aggregate(rowptr, src, n_cap, n_dev, x, cin, mode, pseudo, status):
    """A[i] = [per-tap aggregated sources (ks*ks*cin) | x_i (cin)]."""
    L = _lib.lib()
    lda = (mode["ks"] ** 2 + 1) * cin
    A = torch.empty((max(n_cap, 1), lda), dtype=torch.float32, device=x.device)
    m = _mode_struct(mode)
    p = _pseudo_struct(**pseudo)
    with _timed('spline_aggregate'):
        check(L.evad_spline_aggregate(ptr(rowptr), ptr(src), n_cap, ptr(n_dev), ptr(x), x.stride(0), cin, ctypes.byref(p),
                                      ctypes.byref(m), ptr(A), lda, ptr(status), stream()), "evad_spline_aggregate")
    return A


# EVAD_GEMM_H2=0: pooled layers back on the fp32 operand + cp.async tcgen05 kernel (evad_gemm64_tc5)
USE_GEMM_H2 = os.environ.get("EVAD_GEMM_H2", "1") != "0"


def _aligned_bytes(nbytes, device, align=1024):
    """uint8 buffer whose data pointer is `align`-byte aligned (SWIZZLE_128B operand tiles are 1024-byte atoms)."""
    buf = torch.empty(int(nbytes) + align, dtype=torch.uint8, device=device)
    off = (-buf.data_ptr()) % align
    return buf[off:off + int(nbytes)]


def spline_aggregate_h2(rowptr, src, n_cap, n_dev, x, cin, mode, pseudo, status):
    """spline_aggregate written as the operand image of gemm64_h2: fp16 pairs, tiled and swizzled per (128 rows, 64
    columns) block (csrc/spline_general.cu: AH2_BLOCK).  Returns the byte buffer."""
    L = _lib.lib()
    k = (mode["ks"] ** 2 + 1) * cin
    A = _aligned_bytes(L.evad_gemm64_h2_a_bytes(max(n_cap, 1), k), x.device)
    m = _mode_struct(mode)
    p = _pseudo_struct(**pseudo)
    with _timed('spline_aggregate'):
        check(L.evad_spline_aggregate_h2(ptr(rowptr), ptr(src), n_cap, ptr(n_dev), ptr(x), x.stride(0), cin, ctypes.byref(p),
                                         ctypes.byref(m), ptr(A), ptr(status), stream()), "evad_spline_aggregate_h2")
    return A


def pack_weight_h2(W):
    """W [K, 64] fp32 -> the weight image of gemm64_h2: per 64-row K chunk 16 KB = [w1 | w2] (W = w1 + w2 * 2^-11 in fp16),
    each a 64 x 128-byte tile (row = output channel, K-major) with 16-byte units swizzled by (row & 7)."""
    W = W.detach().float().cpu()
    K = W.shape[0]
    assert W.shape[1] == 64
    kc = (K + 63) // 64
    Wp = torch.zeros((kc * 64, 64), dtype=torch.float32)
    Wp[:K] = W
    w1 = Wp.half()
    w2 = ((Wp - w1.float()) * 2048.0).half()
    t = torch.stack([w.view(kc, 64, 64).permute(0, 2, 1) for w in (w1, w2)], dim=1).contiguous()   # [kc, part, n, kk]
    n, kk = torch.arange(64)[:, None], torch.arange(64)[None, :]
    dst = (((kk >> 3) ^ (n & 7)) << 3) + (kk & 7)                                                  # position inside row n
    out = torch.empty_like(t)
    out.scatter_(3, dst.expand(kc, 2, 64, 64).contiguous(), t)
    return out.view(torch.int16).reshape(-1)


def gemm64_h2(A_h2, W_h2, m_cap, m_dev, epi, k, out=None):
    """out = epi(A @ W) for the operand images of spline_aggregate_h2 / pack_weight_h2 (cp.async.bulk -> tcgen05.mma
    kind::f16 -> TMEM drain, warp-specialised; csrc/gemm_h2.cu)."""
    L = _lib.lib()
    if out is None:
        out = torch.empty((max(m_cap, 1), 64), dtype=torch.float32, device=A_h2.device)
    e = _epi_struct(**epi)
    ws = _ws(L.evad_gemm64_h2_ws_bytes(m_cap, k), A_h2.device)
    with _timed('gemm_epilogue'):
        check(L.evad_gemm64_h2(ptr(A_h2), ptr(W_h2), m_cap, ptr(m_dev), k, ctypes.byref(e), ptr(out), out.stride(0), ptr(ws),
                               ws.numel(), stream()), "evad_gemm64_h2")
    return out


def gemm_epilogue(A, W, m_cap, m_dev, epi, out=None, k=None):
    """out = epi(A[:, :k] @ W) ; W [k, n] row-major."""
    L = _lib.lib()
    k = k or W.shape[0]
    n = W.shape[1]
    if out is None:
        out = torch.empty((max(m_cap, 1), n), dtype=torch.float32, device=A.device)
    e = _epi_struct(**epi)
    ws = _ws(L.evad_gemm_ws_bytes(m_cap, k, n), A.device)
    with _timed('gemm_epilogue'):
        check(L.evad_gemm_epilogue(ptr(A), A.stride(0), ptr(W), m_cap, ptr(m_dev), k, n, ctypes.byref(e), ptr(out),
                                   out.stride(0), ptr(ws), ws.numel(), stream()), "evad_gemm_epilogue")
    return out


def split_tf32(W):
    """W fp32 -> (hi, lo) with hi = tf32(W), lo = tf32(W - hi): the operand form of evad_gemm64_tc."""
    L = _lib.lib()
    W = W.detach().float().contiguous()
    hi, lo = torch.empty_like(W), torch.empty_like(W)
    check(L.evad_split_tf32(ptr(W), W.numel(), ptr(hi), ptr(lo), stream()), "evad_split_tf32")
    return hi, lo


def gemm64_tc(A, W, m_cap, m_dev, epi, out=None, k=None):
    """out = epi(A[:, :k] @ W) on the tensor cores (3xTF32, fp32 accuracy); W [k, 64] row-major fp32."""
    L = _lib.lib()
    k = k or W.shape[0]
    if out is None:
        out = torch.empty((max(m_cap, 1), 64), dtype=torch.float32, device=A.device)
    e = _epi_struct(**epi)
    ws = _ws(L.evad_gemm64_tc_ws_bytes(m_cap, k), A.device)
    with _timed('gemm_epilogue'):
        check(L.evad_gemm64_tc(ptr(A), A.stride(0), ptr(W), m_cap, ptr(m_dev), k, ctypes.byref(e), ptr(out),
                               out.stride(0), ptr(ws), ws.numel(), stream()), "evad_gemm64_tc")
    return out


def transpose_pad_weight(W):
    """W [k, 64] -> Wt [64, ldw] (ldw = k rounded up to a multiple of 4, zero padded): operand of gemm64_tc5."""
    k = W.shape[0]
    ldw = (k + 3) // 4 * 4
    wt = torch.zeros((W.shape[1], ldw), dtype=torch.float32, device=W.device)
    wt[:, :k] = W.detach().float().t()
    return wt


def gemm64_tc5(A, Wt, m_cap, m_dev, epi, out=None, k=None):
    """out = epi(A[:, :k] @ W) with tcgen05.mma (TMEM accumulator, 3xTF32); Wt = transpose_pad_weight(W)."""
    L = _lib.lib()
    k = k or Wt.shape[1]
    if out is None:
        out = torch.empty((max(m_cap, 1), 64), dtype=torch.float32, device=A.device)
    e = _epi_struct(**epi)
    ws = _ws(L.evad_gemm64_tc5_ws_bytes(m_cap, k), A.device)
    with _timed('gemm_epilogue'):
        check(L.evad_gemm64_tc5(ptr(A), A.stride(0), ptr(Wt), Wt.stride(0), m_cap, ptr(m_dev), k, ctypes.byref(e), ptr(out),
                                out.stride(0), ptr(ws), ws.numel(), stream()), "evad_gemm64_tc5")
    return out


def spline_tap_lists(rowptr, src, n_cap, n_dev, e_cap, mode, pseudo, status):
    """Per-node tap-sorted (source, basis value) entries of a pooled graph: dict(tptr [n_cap, ntap+1], ent [4*e_cap, 2],
    ntap); one pass per graph, shared by the convolutions that see the same graph and pseudo-coordinates."""
    L = _lib.lib()
    ntap = mode["ks"] ** 2
    dev = rowptr.device
    tptr = torch.empty((max(n_cap, 1), ntap + 1), dtype=torch.int32, device=dev)
    ent = torch.empty((4 * max(int(e_cap), 1), 2), dtype=torch.int32, device=dev)
    m = _mode_struct(mode)
    p = _pseudo_struct(**pseudo)
    with _timed('spline_aggregate'):
        check(L.evad_spline_tap_lists(ptr(rowptr), ptr(src), n_cap, ptr(n_dev), ctypes.byref(p), ctypes.byref(m), ptr(tptr),
                                      ptr(ent), ptr(status), stream()), "evad_spline_tap_lists")
    return dict(tptr=tptr, ent=ent, ntap=ntap)


def pack_fused_weight(weight, lin_weight, ldx):
    """[K, Cin, Cout=64] spline weight + [Cout, Cin] root weight -> (Wt_hi, Wt_lo, cp): transposed, tap-major with
    cp = roundup8(ldx) zero-padded columns per tap, split into TF32 parts."""
    ntap, cin, cout = weight.shape
    cp = (int(ldx) + 7) // 8 * 8
    wt = torch.zeros((cout, ntap + 1, cp), dtype=torch.float32, device=weight.device)
    wt[:, :ntap, :cin] = weight.detach().float().permute(2, 0, 1)
    wt[:, ntap, :cin] = lin_weight.detach().float()
    hi, lo = split_tf32(wt.view(cout, (ntap + 1) * cp))
    return hi, lo, cp


def spline_conv64_fused(lists, x, n_cap, n_dev, wpack, epi, out=None):
    """out = epi(SplineConv(x)) with Cout = 64 through evad_spline_conv64_fused; wpack = pack_fused_weight(...)."""
    L = _lib.lib()
    hi, lo, cp = wpack
    if out is None:
        out = torch.empty((max(n_cap, 1), 64), dtype=torch.float32, device=x.device)
    e = _epi_struct(**epi)
    ws = _ws(L.evad_spline_conv64_fused_ws_bytes(n_cap, lists["ntap"]), x.device)
    with _timed('gemm_epilogue'):
        check(L.evad_spline_conv64_fused(ptr(lists["tptr"]), ptr(lists["ent"]), ptr(x), x.stride(0), n_cap, ptr(n_dev),
                                         lists["ntap"], cp, ptr(hi), ptr(lo), hi.stride(0), ctypes.byref(e), ptr(out),
                                         out.stride(0), ptr(ws), ws.numel(), stream()), "evad_spline_conv64_fused")
    return out


# --------------------------------------------------------------------------------------------------- pooling
def pool_geometry(voxel_size4: torch.Tensor, end4: torch.Tensor):
    """num_voxels per dimension exactly as torch_cluster computes them: ((end - start) / size).long() + 1, start = 0."""
    nv = ((end4.cpu() - 0) / voxel_size4.cpu()).to(torch.long) + 1
    return [int(v) for v in nv[:3]]


def pool_params(voxel_size, nv, batch_size, inv_wh, aggr, width, height, feat_nhwc=None, append_pos=False, ell_stride=0,
                self_loop=False):
    """evad_pool_params_t of a Pooling layer (the struct keeps `feat_nhwc` alive only through the caller)."""
    fb, fh, fw, fc = (feat_nhwc.shape if feat_nhwc is not None else (0, 0, 0, 0))
    return PoolParamsT(size_x=float(voxel_size[0]), size_y=float(voxel_size[1]), size_t=float(voxel_size[2]), nvx=nv[0],
                       nvy=nv[1], nvt=nv[2], batch_size=int(batch_size), inv_w=float(inv_wh[0]), inv_h=float(inv_wh[1]),
                       aggr_mean=int(aggr == "mean"), append_pos=int(append_pos), feat=ptr(feat_nhwc), feat_b=fb, feat_h=fh,
                       feat_w=fw, feat_c=fc, width=float(width), height=float(height), ell_stride=int(ell_stride),
                       self_loop=int(bool(self_loop)))


class PoolIncState:
    """Incremental voxel pooling state (csrc/pool.cu: evad_pool_inc_*): per-voxel maxima, position sums, counts and coarse
    adjacency of everything folded in since `reset`."""

    def __init__(self, P, cx, capacity, device):
        L = _lib.lib()
        self.P, self.cx = P, int(cx)
        nbytes = L.evad_pool_inc_state_bytes(ctypes.byref(P), self.cx)
        if nbytes == 0:
            raise RuntimeError("eventad_b200: this Pooling configuration has no incremental kernel (max aggregation, "
                               "no self loops, fixed-stride rows, < 2^24 voxel keys)")
        self.buf = _aligned_bytes(nbytes, device, 256)
        self.node_key = torch.zeros(max(int(capacity), 1), dtype=torch.int32, device=device)

    def reset(self):
        check(_lib.lib().evad_pool_inc_reset(ptr(self.buf), ctypes.byref(self.P), self.cx, stream()), "evad_pool_inc_reset")

    def update(self, n0, n_new, pos, batch_i32, x, count, nbr, status):
        L = _lib.lib()
        ws = _ws(L.evad_pool_inc_update_ws_bytes(n_new), pos.device)
        with _timed('pool_inc_update'):
            check(L.evad_pool_inc_update(ptr(self.buf), ctypes.byref(self.P), int(n0), int(n_new), ptr(pos), pos.stride(0),
                                         ptr(batch_i32), ptr(x), x.stride(0), self.cx, ptr(count), ptr(nbr), ptr(self.node_key),
                                         ptr(status), ptr(ws), ws.numel(), stream()), "evad_pool_inc_update")

    def emit(self, c_cap, e_cap, ldo, n_out, num_edges, status):
        """The pooled graph of all folded nodes, in voxel_pool's return format."""
        L = _lib.lib()
        dev = self.buf.device
        x_out = torch.empty((c_cap, ldo), dtype=torch.float32, device=dev)
        pos_out = torch.empty((c_cap, 3), dtype=torch.float32, device=dev)
        batch_out = torch.empty(c_cap, dtype=torch.int32, device=dev)
        rowptr_out = torch.empty(c_cap + 1, dtype=torch.int32, device=dev)
        src_out = torch.empty(e_cap, dtype=torch.int32, device=dev)
        ws = _ws(L.evad_pool_inc_emit_ws_bytes(ctypes.byref(self.P), c_cap), dev)
        with _timed('voxel_pool'):
            check(L.evad_pool_inc_emit(ptr(self.buf), ctypes.byref(self.P), self.cx, c_cap, e_cap, ptr(n_out), ptr(x_out), ldo,
                                       ptr(pos_out), ptr(batch_out), ptr(rowptr_out), ptr(src_out), ptr(num_edges), ptr(status),
                                       ptr(ws), ws.numel(), stream()), "evad_pool_inc_emit")
        return dict(cluster=None, n_out=n_out, x=x_out, pos=pos_out, batch=batch_out, rowptr=rowptr_out, src=src_out,
                    num_edges=num_edges, c_cap=c_cap, e_cap=e_cap, status=status)


def voxel_pool(x, pos, batch_i32, rowptr, src, n_cap, n_dev, voxel_size, nv, batch_size, inv_wh, aggr, width, height,
               feat_nhwc=None, append_pos=False, ldo=None, c_cap=None, e_cap=None, status=None, n_out=None,
               num_edges=None, ell_stride=0, self_loop=False):
    """Returns dict(cluster, n_out, x, pos, batch, rowptr, src, num_edges); all arrays sized by c_cap / e_cap.
    ell_stride > 0: the fine graph comes as fixed-stride rows (rowptr = row lengths [n], src rows at i * ell_stride)."""
    L = _lib.lib()
    dev = pos.device
    cx = x.shape[1]
    cs = 0 if feat_nhwc is None else feat_nhwc.shape[3]
    ctot = cx + cs + (2 if append_pos else 0)
    ldo = ldo or ctot
    key_bound = nv[0] * nv[1] * nv[2] * (batch_size + 1)
    c_cap = int(c_cap if c_cap is not None else max(1, min(n_cap, key_bound)))
    want_edges = rowptr is not None
    if e_cap is None:
        e_cap = int(src.shape[0]) if want_edges else 0
    e_cap = max(int(e_cap), 1)
    status = status if status is not None else new_status(dev)
    cluster = torch.empty(max(n_cap, 1), dtype=torch.int32, device=dev)
    if n_out is None:                                # callers may pass int32[1] cells (views of one counter block)
        n_out = torch.zeros(1, dtype=torch.int32, device=dev)
    x_out = torch.empty((c_cap, ldo), dtype=torch.float32, device=dev)      # row padding is zeroed by the kernel
    pos_out = torch.empty((c_cap, 3), dtype=torch.float32, device=dev)
    batch_out = torch.empty(c_cap, dtype=torch.int32, device=dev)
    rowptr_out = torch.empty(c_cap + 1, dtype=torch.int32, device=dev) if want_edges else None
    src_out = torch.empty(e_cap, dtype=torch.int32, device=dev) if want_edges else None
    if not want_edges:
        num_edges = None
    elif num_edges is None:
        num_edges = torch.zeros(1, dtype=torch.int32, device=dev)
    fb, fh, fw, fc = (feat_nhwc.shape if feat_nhwc is not None else (0, 0, 0, 0))
    P = PoolParamsT(size_x=float(voxel_size[0]), size_y=float(voxel_size[1]), size_t=float(voxel_size[2]), nvx=nv[0],
                    nvy=nv[1], nvt=nv[2], batch_size=int(batch_size), inv_w=float(inv_wh[0]), inv_h=float(inv_wh[1]),
                    aggr_mean=int(aggr == "mean"), append_pos=int(append_pos), feat=ptr(feat_nhwc), feat_b=fb, feat_h=fh,
                    feat_w=fw, feat_c=fc, width=float(width), height=float(height), ell_stride=int(ell_stride), self_loop=int(bool(self_loop)))
    ws = _ws(L.evad_voxel_pool_ws_bytes(n_cap, c_cap), dev)
    x = x.contiguous()
    pos = pos.contiguous()
    with _timed('voxel_pool'):
        check(L.evad_voxel_pool(ctypes.byref(P), n_cap, ptr(n_dev), ptr(pos), pos.stride(0), ptr(batch_i32), ptr(x),
                                x.stride(0), cx, ptr(rowptr), ptr(src), c_cap, e_cap, ptr(cluster), ptr(n_out), ptr(x_out), ldo,
                                ptr(pos_out), ptr(batch_out), ptr(rowptr_out), ptr(src_out), ptr(num_edges), ptr(status),
                                ptr(ws), ws.numel(), stream()), "evad_voxel_pool")
    return dict(cluster=cluster, n_out=n_out, x=x_out, pos=pos_out, batch=batch_out, rowptr=rowptr_out, src=src_out,
                num_edges=num_edges, c_cap=c_cap, e_cap=e_cap, status=status)
