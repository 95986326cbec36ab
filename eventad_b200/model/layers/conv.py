"""ConvBlock, ConvBlockWithSkip, Layer with the reference's constructors and state-dict keys
(/root/reference/src/dagr/model/layers/conv.py:10-72).  In eval mode `Layer.forward` runs the fused kernels:
conv + folded BatchNorm + ReLU, and conv + BatchNorm + (Linear + BatchNorm skip) + ReLU, each one or two launches."""
import torch

from ... import _lib, ops
from ..utils import shallow_copy
from .components import BatchNormData, Linear
from .spline_conv import MySplineConv, get_csr, no_backward_guard


def _fold(bn_data, conv_bias=None):
    return ops.fold_bn(bn_data.module, conv_bias)


def _same_mode(a, b):
    """True when two convolutions evaluate the same basis for an edge (kernel size, LUT geometry and tables)."""
    if a is b:
        return True
    for k in a:
        va, vb = a[k], b.get(k)
        if isinstance(va, torch.Tensor):
            if not (isinstance(vb, torch.Tensor) and va.shape == vb.shape and bool(torch.equal(va, vb.to(va.device)))):
                return False
        elif va != vb:
            return False
    return True


class ConvBlock(torch.nn.Module):
    def __init__(self, in_channels: int, out_channels: int, args, degree=1) -> None:
        super(ConvBlock, self).__init__()
        self.dim = args.edge_attr_dim
        self.activation = getattr(torch.nn.functional, args.activation, torch.nn.functional.elu)
        self.conv = MySplineConv(in_channels=in_channels, out_channels=out_channels, args=args, bias=False, degree=degree)
        self.norm = BatchNormData(in_channels=out_channels)

    def forward(self, data):
        data = self.conv(data)
        data = self.norm(data)
        data.x = self.activation(data.x)
        return data


class ConvBlockWithSkip(torch.nn.Module):
    def __init__(self, in_channel: int, out_channel: int, skip_in_channel: int, args) -> None:
        super(ConvBlockWithSkip, self).__init__()
        self.dim = args.edge_attr_dim
        self.conv = MySplineConv(in_channels=in_channel, out_channels=out_channel, args=args, bias=False)
        self.activation = getattr(torch.nn.functional, args.activation, torch.nn.functional.elu)
        self.norm = BatchNormData(in_channels=out_channel)
        self.lin = Linear(skip_in_channel, out_channel, bias=False)
        self.norm_skip = BatchNormData(in_channels=out_channel)

    def forward(self, data, data_skip):
        data = self.conv(data)
        data_skip = self.lin(data_skip)
        data_skip = self.norm_skip(data_skip)
        data = self.norm(data)
        data.x = self.activation(data.x + data_skip.x)
        return data


class Layer(torch.nn.Module):
    def __init__(self, in_channels: int, out_channels: int, args) -> None:
        super(Layer, self).__init__()
        self.in_channel = in_channels
        self.out_channel = out_channels
        self.conv_block1 = ConvBlock(in_channels, out_channels, args)
        self.conv_block2 = ConvBlockWithSkip(out_channels, out_channels, in_channels, args=args)

    # fused eval path on raw tensors (used by Layer.forward and by the fused Net.forward)
    def h16_chain_ok(self, src, cls, ell_stride, n_dev=None):
        """True when both convolutions of the layer run on evad_spline_conv_l1_h16, so that features can travel between
        them (and into the first one: run_fused(x_h2=True)) as h2 words."""
        b1, b2 = self.conv_block1, self.conv_block2
        return bool(b1.activation is torch.nn.functional.relu and cls is not None and n_dev is None and
                    ops.l1_h16_ok(src, cls, ell_stride) and b1.conv.small_ok() and b2.conv.small_ok() and
                    b1.conv.l1_pack(b2.lin.mlp.weight) is not None and b2.conv.l1_pack() is not None)

    def run_fused(self, x, rowptr, src, n_cap, n_dev, pseudo, status, perm=None, cls=None, ell_stride=0, x_h2=False):
        b1, b2 = self.conv_block1, self.conv_block2
        relu = b1.activation is torch.nn.functional.relu
        s1, h1 = _fold(b1.norm, b1.conv.bias)
        s2, h2 = _fold(b2.norm, b2.conv.bias)
        ss, hs = _fold(b2.norm_skip, b2.lin.mlp.bias)
        if cls is None and n_dev is None and b1.conv.small_ok() and b2.conv.small_ok() and b1.conv.l1_pack(b2.lin.mlp.weight) is not None:
            # both convolutions of the layer see the same graph and lattice: one class byte per edge serves both
            from .spline_conv import USE_L1
            if USE_L1:
                cls = ops.edge_class(rowptr, src, n_cap, b1.conv.mode(), pseudo, status)
        lists = None
        from .spline_conv import FUSED_MIN_ROWS, USE_FUSED
        m1, m2 = b1.conv.mode(), b2.conv.mode()
        sm = getattr(self, "_same_mode_cache", None)
        if sm is None or sm[0] is not m1 or sm[1] is not m2:  # compared once per LUT (the comparison synchronises)
            sm = self._same_mode_cache = (m1, m2, _same_mode(m1, m2))
        if USE_FUSED and b1.conv.out_channels == 64 and b2.conv.out_channels == 64 and n_cap >= FUSED_MIN_ROWS and sm[2]:
            # both convolutions see the same graph and pseudo-coordinates: one tap-sorted entry list serves both
            lists = ops.spline_tap_lists(rowptr, src, n_cap, n_dev, src.shape[0], b1.conv.mode(), pseudo, status)
        # fp16-pair layer-1 kernel: the first convolution hands its output to the second one as h2 words (ops.split_h2)
        mid_h2 = self.h16_chain_ok(src, cls, ell_stride, n_dev)
        assert mid_h2 or not x_h2
        y, skip = b1.conv.run(x, rowptr, src, n_cap, n_dev, pseudo, dict(scale=s1, shift=h1, relu=relu), status, lists=lists,
                              skip_weight=b2.lin.mlp.weight, skip_epi=dict(scale=ss, shift=hs), cls=cls, perm=perm,
                              ell_stride=ell_stride, x_h2=x_h2, out_h2=mid_h2)
        if not relu:
            y = b1.activation(y)
        out, _ = b2.conv.run(y, rowptr, src, n_cap, n_dev, pseudo, dict(scale=s2, shift=h2, addend=skip, relu=relu), status, lists=lists,
                             cls=cls, perm=perm, ell_stride=ell_stride, x_h2=mid_h2)
        if not relu:
            out = b2.activation(out)
        return out

    def forward(self, data):
        if self.training:
            data_skip = shallow_copy(data)
            data = self.conv_block1(data)
            return self.conv_block2(data, data_skip)
        x = data.x.float()
        n = x.shape[0]
        no_backward_guard(self, x)
        status = ops.new_status(x.device)
        if data.edge_index.numel() > 0:
            csr = get_csr(data)
            rowptr, src = csr["rowptr"], csr["src"]
            pseudo = dict(edge_attr=data.edge_attr[:, :self.conv_block1.dim].float().contiguous(), eid=csr["eid"])
        else:
            rowptr = torch.zeros(n + 1, dtype=torch.int32, device=x.device)
            src = torch.zeros(1, dtype=torch.int32, device=x.device)
            pseudo = dict(edge_attr=torch.zeros((1, 2), device=x.device), eid=None)
        out = self.run_fused(x, rowptr, src, n, None, pseudo, status)
        _lib.check_status(status, "Layer.forward")
        data.x = out[:n]
        return data
