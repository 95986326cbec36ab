"""BatchNormData, Linear, Cartesian with the reference's names, constructor arguments and state-dict keys
(/root/reference/src/dagr/model/layers/components.py:9-35)."""
import torch

from ... import ops


class _PyGBatchNorm(torch.nn.Module):
    """torch_geometric.nn.BatchNorm: a wrapper whose `.module` is nn.BatchNorm1d (keys `module.weight`, ...)."""

    def __init__(self, in_channels, eps=1e-5, momentum=0.1, affine=True, track_running_stats=True):
        super().__init__()
        self.module = torch.nn.BatchNorm1d(in_channels, eps, momentum, affine, track_running_stats)
        self.in_channels = in_channels

    def forward(self, x):
        return self.module(x)


class BatchNormData(_PyGBatchNorm):
    def forward(self, data):
        data.x = _PyGBatchNorm.forward(self, data.x)
        return data


class Linear(torch.nn.Module):
    def __init__(self, ic, oc, bias=True):
        torch.nn.Module.__init__(self)
        self.mlp = torch.nn.Linear(ic, oc, bias=bias)

    def forward(self, data):
        epi = dict(shift=self.mlp.bias.detach() if self.mlp.bias is not None else None)
        x = data.x.contiguous()
        data.x = ops.gemm_epilogue(x, self.mlp.weight.detach().t().contiguous(), x.shape[0], None, epi)[:x.shape[0]]
        return data


class Cartesian(torch.nn.Module):
    """PyG T.Cartesian(norm, cat, max_value) as a module (components.py:25-35); attributes norm, max, cat."""

    def __init__(self, norm=True, max_value=None, cat=True):
        super().__init__()
        self.norm = norm
        self.max = max_value
        self.cat = cat

    def forward(self, data):
        if data.edge_index.shape[1] > 0:
            return apply_cartesian(self, data)
        data.edge_attr = torch.zeros((0, 3), dtype=data.x.dtype, device=data.x.device)
        return data


class CartesianTransform:
    """Plain-callable twin (the `transform=` argument of Pooling, net.py:75-93)."""

    def __init__(self, norm=True, max_value=None, cat=True):
        self.norm = norm
        self.max = max_value
        self.cat = cat

    def __call__(self, data):
        return apply_cartesian(self, data)


def two_m_of(max_value) -> float:
    """fp32(2 * max_value) exactly as `cart / (2 * max_value)` sees it (python float or 0-dim fp32 tensor)."""
    if torch.is_tensor(max_value):
        return float((2 * max_value.detach().float().cpu()).item())
    return float(torch.tensor(2 * max_value, dtype=torch.float32).item())


def apply_cartesian(t, data):
    if not t.norm or t.max is None or t.cat:
        raise NotImplementedError("only Cartesian(norm=True, cat=False, max_value=m) is on the EventAD path")
    pos = data.pos.float()
    data.edge_attr = ops.cartesian(pos, data.edge_index, two_m_of(t.max), False)
    return data
