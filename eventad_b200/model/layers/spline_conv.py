"""MySplineConv / SplineConvToDense with the reference's constructor, attributes and state-dict keys
(/root/reference/src/dagr/model/layers/spline_conv.py:9-118; base class PyG SplineConv 2.6.1: `weight [K,Cin,Cout]`,
`lin.weight [Cout,Cin]`, optional `bias`, persistent buffers `kernel_size`, `is_open_spline`), running on the
sm_100a kernels (csrc/spline_small.cu, csrc/spline_general.cu).  There is no PyTorch fallback for the message passing.
"""
import math

import os

import torch

from ... import ops


USE_TC5 = True      # tcgen05 (UMMA / TMEM) contraction for the cout=64 layers (False: legacy mma.sync kernel)
TC5_MIN_ROWS = 1024 # below this many rows the 128-row tcgen05 tiles leave SMs idle and the mma.sync kernel is faster
                    # (tools/gemm_bench.py: 65 vs 94 us at 17920 x 2132, 11.8 vs 8.5 us at 140 x 3380)
USE_L1 = True       # tap aggregation + tensor-core contraction for the cout=16 layer (False: LUT-gather kernel)
USE_FUSED = os.environ.get("EVAD_FUSED", "0") == "1"   # pooled layers: tap aggregation + tcgen05 contraction in ONE kernel (csrc/spline_tc5.cu): parity-tested, at
                    # best equal to aggregate + GEMM (profiles/r01d_summary.md): the per-tap CTA barrier waits for the longest row
FUSED_MIN_ROWS = 1024


class _PyGLinear(torch.nn.Module):
    """torch_geometric.nn.dense.linear.Linear(in, out, bias=False): key `weight` [out, in]."""

    def __init__(self, in_channels, out_channels):
        super().__init__()
        self.in_channels, self.out_channels = in_channels, out_channels
        self.weight = torch.nn.Parameter(torch.empty(out_channels, in_channels))
        self.register_parameter("bias", None)

    def forward(self, x):
        return torch.nn.functional.linear(x, self.weight)


def no_backward_guard(module, x):
    """The kernels are forward-only (EventAD keeps DAGR frozen, /root/reference/models/EventAD.py:149-150).  Asking for
    gradients through them would silently train nothing, so it is an error."""
    if torch.is_grad_enabled() and (x.requires_grad or any(p.requires_grad for p in module.parameters())):
        raise RuntimeError("eventad_b200: the DAGR event-branch kernels have no backward; run them under torch.no_grad() "
                           "or freeze the module (requires_grad_(False)) as EventADModel does")


def get_csr(data):
    """dst-major CSR of data.edge_index, cached on the Data object (built by evad_coo_to_csr when missing)."""
    c = getattr(data, "_csr", None)
    ei = data.edge_index
    if c is not None and c.get("edge_index_ptr") == (ei.data_ptr(), tuple(ei.shape)):
        return c
    n = data.x.shape[0]
    rowptr, src, eid = ops.coo_to_csr(ei, n)
    c = dict(rowptr=rowptr, src=src, eid=eid, edge_index_ptr=(ei.data_ptr(), tuple(ei.shape)))
    data._csr = c
    return c


class MySplineConv(torch.nn.Module):
    def __init__(self, in_channels, out_channels, args, bias=False, degree=1, **kwargs):
        super().__init__()
        self.reproducible = False
        self.in_channels, self.out_channels = in_channels, out_channels
        self.dim = args.edge_attr_dim
        self.degree = degree
        self.aggr = args.aggr
        self.root_weight = True
        if self.dim != 2 or degree != 1:
            raise NotImplementedError("the EventAD path uses dim=2, degree=1 (spline_conv.py:10-14)")
        if self.aggr not in ("sum", "add"):
            raise NotImplementedError("only sum aggregation is implemented (reference default, SURVEY A.1)")
        ks = int(args.kernel_size)
        self.register_buffer("kernel_size", torch.tensor([ks] * self.dim, dtype=torch.long))
        self.register_buffer("is_open_spline", torch.ones(self.dim, dtype=torch.uint8))
        self.K = ks ** self.dim
        self.weight = torch.nn.Parameter(torch.empty(self.K, in_channels, out_channels))
        self.lin = _PyGLinear(in_channels, out_channels)
        if bias:
            self.bias = torch.nn.Parameter(torch.empty(out_channels))
        else:
            self.register_parameter("bias", None)
        self.reset_parameters()
        self._mode = None           # LUT parameters after init_lut
        self._cache = {}
        self.message = self.message_basis

    def reset_parameters(self):
        bound = 1.0 / math.sqrt(self.in_channels * self.K)
        torch.nn.init.uniform_(self.weight, -bound, bound)
        torch.nn.init.uniform_(self.lin.weight, -bound, bound)
        if self.bias is not None:
            torch.nn.init.zeros_(self.bias)

    # ---- reference API -------------------------------------------------------------------------------------
    def init_lut(self, height, width, rx=None, Mx=None, ry=None, My=None):
        """Switches to message_lut semantics (spline_conv.py:16-37).  The dense [2rx+1, 2ry+1, Cin, Cout] table is
        only materialised for the shared-memory kernel of the small first layer; the general kernel snaps edges to
        the same integer lattice and evaluates the same basis values from two 1-D tables."""
        self._mode = ops.spline_mode_lut(height, width, rx, Mx, ry, My, int(self.kernel_size[0]), self.weight.device)
        self.attr_remapping_matrix = self._mode["R"]
        self._cache = {}
        self.message = self.message_lut

    def message_lut(self, x_j, edge_attr):
        raise RuntimeError("message passing runs inside the fused CUDA kernels; call forward()")

    message_basis = message_lut

    def mode(self):
        m = self._mode
        if m is None:
            m = self._cache.get("basis_mode")
            if m is None:
                m = self._cache["basis_mode"] = ops.spline_mode_basis(int(self.kernel_size[0]))
        elif m["tab_k"].device != self.weight.device:
            m["tab_k"] = m["tab_k"].to(self.weight.device)
            m["tab_w"] = m["tab_w"].to(self.weight.device)
        return m

    # ---- cached weight layouts -----------------------------------------------------------------------------
    def _versions(self, extra=None):
        v = [self.weight._version, self.lin.weight._version, self.weight.device]
        if extra is not None:
            v.append(extra._version)
        return tuple(v)

    def wcat(self):
        """[K*Cin + Cin, Cout]: the K tap matrices stacked, then lin.weight^T (root weight)."""
        key = self._versions()
        c = self._cache.get("wcat")
        if c is None or c[0] != key:
            w = torch.cat([self.weight.detach().reshape(self.K * self.in_channels, self.out_channels),
                           self.lin.weight.detach().t()], 0).contiguous().float()
            c = self._cache["wcat"] = (key, w)
        return c[1]

    def wcat_t(self):
        """wcat() transposed and padded: [Cout, roundup4(K*Cin + Cin)], the operand layout of the tcgen05 kernel."""
        key = self._versions()
        c = self._cache.get("wcat_t")
        if c is None or c[0] != key:
            c = self._cache["wcat_t"] = (key, ops.transpose_pad_weight(self.wcat()))
        return c[1]

    def wcat_h2(self):
        """wcat() as the weight image of evad_gemm64_h2 (fp16 pairs, tiled + swizzled per 64-row K chunk)."""
        key = self._versions()
        c = self._cache.get("wcat_h2")
        if c is None or c[0] != key:
            img = ops.pack_weight_h2(self.wcat()).to(self.weight.device)
            buf = ops._aligned_bytes(img.numel() * 2, self.weight.device)
            buf.view(torch.int16).copy_(img)
            c = self._cache["wcat_h2"] = (key, buf)
        return c[1]

    def fused_pack(self, ldx):
        """(Wt_hi, Wt_lo, cp) of evad_spline_conv64_fused for inputs with row stride ldx."""
        key = self._versions() + (int(ldx),)
        c = self._cache.get("fused")
        if c is None or c[0] != key:
            c = self._cache["fused"] = (key, ops.pack_fused_weight(self.weight, self.lin.weight, ldx))
        return c[1]

    def small_lut(self, skip_weight=None):
        cin_pad = (self.in_channels + 3) // 4 * 4
        key = self._versions(skip_weight)
        c = self._cache.get("lut")
        if c is None or c[0] != key:
            lut = ops.build_lut(self.weight.float(), self.lin.weight.float(),
                                skip_weight.float() if skip_weight is not None else None, cin_pad, self.mode())
            c = self._cache["lut"] = (key, lut)
        return c[1]

    def l1_pack(self, skip_weight=None):
        """Operands of evad_spline_conv_l1, or None when the layer does not fit it (cout 16, Cin_pad in {16, 20}, LUT
        mode, at most 15 distinct kernel taps reachable from the integer lattice)."""
        m = self.mode()
        cin_pad = (self.in_channels + 3) // 4 * 4
        if not (m["use_lut"] == 1 and self.out_channels == 16 and cin_pad in (16, 20) and m["nx"] * m["ny"] <= 256):
            return None
        key = self._versions(skip_weight)
        c = self._cache.get("l1")
        if c is not None and c[0] == key:
            return c[1]
        nx, ny, ks, dev = m["nx"], m["ny"], m["ks"], self.weight.device
        tk, tw = m["tab_k"].cpu().long(), m["tab_w"].cpu().float()
        kx, wx, ky, wy = tk[:nx], tw[:nx], tk[nx:], tw[nx:]
        kk = kx[:, None, :, None] + ks * ky[None, :, None, :]            # [nx, ny, mx, my]
        kk = torch.where(kk < 0, kk + ks * ks, kk)
        b = wx[:, None, :, None] * wy[None, :, None, :]                  # fp32 products, as in init_lut
        used = sorted(set(kk.flatten().tolist()))
        pack = None
        if len(used) <= 15:
            slot = torch.full((ks * ks,), -1, dtype=torch.long)
            slot[torch.tensor(used)] = torch.arange(len(used))
            tapw = torch.zeros(nx * ny, 16)
            tapw.index_put_((torch.arange(nx * ny).view(nx, ny, 1, 1).expand_as(kk), slot[kk]), b, accumulate=True)
            cin = self.in_channels
            W = torch.zeros(16, cin_pad, 16)
            W[:len(used), :cin] = self.weight.detach().float().cpu()[used]
            W[15, :cin] = self.lin.weight.detach().float().cpu().t()
            whi, wlo = ops.split_tf32(W.view(16 * cin_pad, 16).to(dev))
            pack = dict(tapw=tapw.to(dev).contiguous(), whi=whi, wlo=wlo, shi=None, slo=None, cin_pad=cin_pad)
            S = None
            if skip_weight is not None:
                S = torch.zeros((cin_pad + 7) // 8 * 8, 16)
                S[:cin] = skip_weight.detach().float().cpu().t()
                pack["shi"], pack["slo"] = ops.split_tf32(S.to(dev))
            p16 = ops.l1_pack16(W.view(16 * cin_pad, 16), S, cin_pad)        # fp16 operands of the tcgen05 contraction
            pack.update(w16=p16["w16"].to(dev), inv_sw=p16["inv_sw"], inv_ss=p16["inv_ss"],
                        s16=p16["s16"].to(dev) if p16["s16"] is not None else None)
        self._cache["l1"] = (key, pack)
        return pack

    def small_ok(self):
        m = self.mode()
        return (m["use_lut"] == 1 and self.out_channels == 16 and self.in_channels <= 32 and
                (m["nx"] * m["ny"] + 2) * ((self.in_channels + 3) // 4 * 4) * 16 * 4 <= 227 * 1024)

    # ---- kernels ---------------------------------------------------------------------------------------------
    def run(self, x, rowptr, src, n_cap, n_dev, pseudo, epi, status, skip_weight=None, skip_epi=None, cls=None,
            perm=None, ell_stride=0, lists=None, x_h2=False, out_h2=False):
        """x [n_cap, >=Cin] float32 -> epi(conv(x)) [n_cap, Cout] (+ skip_epi(x @ skip_weight^T) when requested)."""
        cin = self.in_channels
        if self.small_ok() and n_dev is None:
            cin_pad = (cin + 3) // 4 * 4
            # rows must be 16-byte aligned and zero beyond Cin; a [n, cin_pad] input is trusted to be zero padded
            if not (x.stride(1) == 1 and x.stride(0) % 4 == 0 and x.data_ptr() % 16 == 0 and x.shape[1] >= cin_pad):
                xp = torch.zeros((x.shape[0], cin_pad), dtype=torch.float32, device=x.device)
                xp[:, :cin] = x[:, :cin]
                x = xp
            pack = self.l1_pack(skip_weight) if USE_L1 else None
            if pack is not None:
                if cls is None:
                    assert ell_stride == 0
                    cls = ops.edge_class(rowptr, src, n_cap, self.mode(), pseudo, status)
                return ops.spline_conv_l1(rowptr, src, cls, n_cap, x, cin, pack, epi, skip_epi, perm=perm,
                                          ell_stride=ell_stride, x_h2=x_h2, out_h2=out_h2, status=status)
            assert not x_h2 and not out_h2
            assert ell_stride == 0, "fixed-stride rows are only read by the layer-1 tensor-core kernel"
            return ops.spline_conv_small(rowptr, src, n_cap, x, cin, self.small_lut(skip_weight), self.mode(), pseudo, epi,
                                         status, skip_epi)
        x = x.contiguous()
        fused = (USE_FUSED and self.out_channels == 64 and self.K <= 32 and n_cap >= FUSED_MIN_ROWS and
                 x.stride(0) % 4 == 0 and x.data_ptr() % 16 == 0 and x.shape[1] == x.stride(0) and x.shape[1] - cin < 4)
        if fused:
            if lists is None:
                lists = ops.spline_tap_lists(rowptr, src, n_cap, n_dev, src.shape[0], self.mode(), pseudo, status)
            out = ops.spline_conv64_fused(lists, x, n_cap, n_dev, self.fused_pack(x.stride(0)), epi)
            A = None
        elif (ops.USE_GEMM_H2 and USE_TC5 and self.out_channels == 64 and n_cap >= TC5_MIN_ROWS and cin <= 160 and
              x.stride(0) % 4 == 0 and x.data_ptr() % 16 == 0 and
              4 * (16 * x.stride(0) + 512) * 4 <= 200 * 1024):
            # large pooled layers: the aggregated operand is written as fp16 pairs in the UMMA operand image and the
            # contraction is a bulk-copy -> tcgen05 pipeline (csrc/gemm_h2.cu); nothing re-formats it in between
            A = None
            Ah = ops.spline_aggregate_h2(rowptr, src, n_cap, n_dev, x, cin, self.mode(), pseudo, status)
            out = ops.gemm64_h2(Ah, self.wcat_h2(), n_cap, n_dev, epi, k=(self.K + 1) * cin)
            fused = True                                     # `out` is final
        else:
            A = ops.spline_aggregate(rowptr, src, n_cap, n_dev, x, cin, self.mode(), pseudo, status)
        tc = self.out_channels == 64 and A is not None and A.stride(0) % 4 == 0
        if fused:
            pass
        elif tc and USE_TC5 and n_cap >= TC5_MIN_ROWS and A.data_ptr() % 16 == 0:
            out = ops.gemm64_tc5(A, self.wcat_t(), n_cap, n_dev, epi, k=A.shape[1])
        elif tc:
            out = ops.gemm64_tc(A, self.wcat(), n_cap, n_dev, epi)
        else:
            out = ops.gemm_epilogue(A, self.wcat(), n_cap, n_dev, epi)
        skip = None
        if skip_epi is not None:
            c = self._cache.get("skip_t")
            if c is None or c[0] != (skip_weight._version, skip_weight.device):
                wt = skip_weight.detach().t().contiguous().float()
                c = self._cache["skip_t"] = ((skip_weight._version, skip_weight.device), wt,
                                             ops.transpose_pad_weight(wt) if wt.shape[1] == 64 else None)
            if c[2] is not None and USE_TC5 and n_cap >= 8 * TC5_MIN_ROWS and x.stride(0) % 4 == 0 and x.data_ptr() % 16 == 0:
                skip = ops.gemm64_tc5(x, c[2], n_cap, n_dev, skip_epi, k=cin)
            elif c[1].shape[1] == 64 and x.stride(0) % 4 == 0 and x.data_ptr() % 16 == 0:
                skip = ops.gemm64_tc(x, c[1], n_cap, n_dev, skip_epi, k=cin)
            else:
                skip = ops.gemm_epilogue(x, c[1], n_cap, n_dev, skip_epi, k=cin)
        return out, skip

    def forward(self, data):
        x = data.x.float()
        n = x.shape[0]
        no_backward_guard(self, x)
        status = ops.new_status(x.device)
        if data.edge_index.numel() > 0:
            csr = get_csr(data)
            rowptr, src, eid = csr["rowptr"], csr["src"], csr["eid"]
            attr = data.edge_attr[:, :self.dim].float().contiguous()
            pseudo = dict(edge_attr=attr, eid=eid)
        else:
            rowptr = torch.zeros(n + 1, dtype=torch.int32, device=x.device)
            src = torch.zeros(1, dtype=torch.int32, device=x.device)
            pseudo = dict(edge_attr=torch.zeros((1, 2), device=x.device), eid=None)
        epi = dict(shift=self.bias.detach().float() if self.bias is not None else None)
        out, _ = self.run(x, rowptr, src, n, None, pseudo, epi, status)
        from ... import _lib
        _lib.check_status(status, "MySplineConv.forward")    # e.g. a LUT offset outside its table: raise, do not clamp silently
        data.x = out[:n]
        return data


def to_dense(self, x, pos, pooling, batch=None, batch_size=None):
    """Nodes of a pooled graph -> dense map [B, C, H, W] (spline_conv.py:80-106; head only, not on the EventAD path).
    Same state as the reference: `self.batch_size` is fixed by the first call (argument, else 1 without `batch`, else
    batch.max() + 1) and `self.dense` is allocated once, zeroed per call and returned (sliced to B samples)."""
    if not hasattr(self, "batch_size"):
        if batch_size is None:
            if batch is None:
                batch = x.new_zeros(len(x), dtype=torch.long)
            batch_size = 1 if batch.numel() == 0 or batch is None else int(batch.max().item()) + 1
        self.batch_size = batch_size
    B = self.batch_size
    cell = pooling[:2]
    if not hasattr(self, "dense"):
        grid_w, grid_h = (1 / cell + 1e-3).long()
        self.dense = x.new_zeros((B, x.shape[-1], int(grid_h), int(grid_w)))
    col, row = (pos[:, :2] / cell).t().long()
    self.dense = self.dense.detach()
    self.dense.zero_()
    view = self.dense[:B] if B < self.dense.shape[0] else self.dense
    view[batch.long(), :, row, col] = x
    return view


class SplineConvToDense(MySplineConv):
    def forward(self, data, batch_size=None):
        data = super().forward(data)
        if data.batch is None:
            data.batch = torch.zeros(len(data.x), dtype=torch.long, device=data.x.device)
        return self.to_dense(data.x, data.pos, data.pooling, data.batch, batch_size=batch_size)

    def to_dense(self, x, pos, pooling, batch=None, batch_size=None):
        return to_dense(self, x, pos, pooling, batch, batch_size=batch_size)
