"""oracle/dagr_layers.py -- torch-CPU restatement of the DAGR layers on the EventAD hot path.

TEST INFRASTRUCTURE ONLY (see oracle/dagr_ops.py).  Each function cites the reference lines it
follows; file names are relative to /root/reference/src/dagr/model.
Parameters are read from a flat state dict that uses the reference's own key names
 relative to `Net`, e.g. "layer2.conv_block1.conv.weight".
"""
import math
from types import SimpleNamespace

import numpy as np
import torch

from . import dagr_ops as ops


# ----------------------------------------------------------------------------- LUT parameters
def lut_params_layer1(radius: float, width: int):
    """DAGR.cache_luts, first layer (networks/dagr.py:37-40)."""
    M = 2 * float(int(radius * width + 2) / width)
    r = int(radius * width + 1)
    return dict(rx=r, ry=r, Mx=M, My=M)


def lut_params_from_pool(voxel_size: torch.Tensor, transform_max, height: int, width: int):
    """voxel_size_to_params (model/utils.py:112-116) as used by cache_luts (dagr.py:42-63)."""
    rx = int(np.ceil(2 * voxel_size[0].cpu().numpy() * width))
    ry = int(np.ceil(2 * voxel_size[1].cpu().numpy() * height))
    return dict(rx=rx, ry=ry, Mx=transform_max, My=transform_max)


def attr_remapping_matrix(height, width, rx, Mx, ry=None, My=None):
    """layers/spline_conv.py:21-24."""
    ry = ry or rx
    My = My or Mx
    return torch.Tensor([[2 * Mx * width, 0, -Mx * width + rx],
                         [0, 2 * My * height, -My * height + ry]])


def lut_rows(weight, kernel_size, is_open_spline, height, width, ix, iy, rx, Mx, ry=None, My=None):
    """Rows (ix, iy) of MySplineConv.init_lut's table (layers/spline_conv.py:26-36), evaluated lazily:
    the table entry only depends on its own (dx, dy), so computing the requested rows is the same arithmetic."""
    ry = ry or rx
    My = My or Mx
    dx = (ix - rx).float() / (2 * Mx * width) + 0.5
    dy = (iy - ry).float() / (2 * My * height) + 0.5
    edge_attr = torch.stack([dx, dy], 1)
    bil_w, indices = ops.spline_basis(edge_attr, kernel_size, is_open_spline, 1)
    return (bil_w[..., None, None] * weight[indices]).sum(1)          # [n, cin, cout]


# ----------------------------------------------------------------------------- MySplineConv
def spline_conv(x, edge_index, edge_attr, sd, prefix, lut=None, height=None, width=None, chunk=1 << 18):
    """MySplineConv.forward/_forward (layers/spline_conv.py:49-78) with PyG propagate (gather x_j =
    x[edge_index[0]], message, sum-aggregate into edge_index[1]).  `lut` = dict(rx,ry,Mx,My) selects
    message_lut (spline_conv.py:39-47, the EventAD inference mode, utils/model.py:34-39); None selects the
    PyG basis message (spline_basis + spline_weighting)."""
    weight = sd[prefix + "weight"]
    kernel_size = sd.get(prefix + "kernel_size", torch.tensor([int(round(math.sqrt(weight.shape[0])))] * 2))
    is_open = sd.get(prefix + "is_open_spline", torch.ones(2, dtype=torch.uint8))
    n, cout = x.shape[0], weight.shape[2]
    if edge_index.numel() > 0:
        attr = edge_attr[:, :2]
        src, dst = edge_index[0], edge_index[1]
        msgs = []
        if lut is not None:
            R = attr_remapping_matrix(height, width, lut["rx"], lut["Mx"], lut["ry"], lut["My"])
            ry_ = lut["ry"] or lut["rx"]
            ix = (attr[:, 0] * R[0, 0] + R[0, -1] + 1e-3).long()
            iy = (attr[:, 1] * R[1, 1] + R[1, -1] + 1e-3).long()
            code = ix * (2 * ry_ + 1) + iy
            ucode, inv = torch.unique(code, return_inverse=True)
            table = lut_rows(weight, kernel_size, is_open, height, width,
                             ucode // (2 * ry_ + 1), ucode % (2 * ry_ + 1), lut["rx"], lut["Mx"], lut["ry"], lut["My"])
            for s in range(0, src.numel(), chunk):
                w = table[inv[s:s + chunk]]
                msgs.append(torch.einsum("nio,ni->no", w, x[src[s:s + chunk]]))
        else:
            for s in range(0, src.numel(), chunk):
                basis, index = ops.spline_basis(attr[s:s + chunk], kernel_size, is_open, 1)
                msgs.append(ops.spline_weighting(x[src[s:s + chunk]], weight, basis, index))
        out = ops.scatter_sum(torch.cat(msgs, 0), dst, n)
    else:
        out = torch.zeros((n, cout), dtype=x.dtype)
    out = out + x @ sd[prefix + "lin.weight"].t()                   # root weight (PyG Linear, no bias)
    if (prefix + "bias") in sd:
        out = out + sd[prefix + "bias"]
    return out


def _bn(x, sd, prefix):
    return ops.batch_norm_eval(x, sd[prefix + "module.weight"], sd[prefix + "module.bias"],
                               sd[prefix + "module.running_mean"], sd[prefix + "module.running_var"])


def layer(x, edge_index, edge_attr, sd, prefix, lut=None, height=None, width=None, act=torch.relu):
    """Layer = ConvBlock + ConvBlockWithSkip (layers/conv.py:10-72)."""
    x_skip = x.clone()
    y = spline_conv(x, edge_index, edge_attr, sd, prefix + "conv_block1.conv.", lut, height, width)
    y = act(_bn(y, sd, prefix + "conv_block1.norm."))
    z = spline_conv(y, edge_index, edge_attr, sd, prefix + "conv_block2.conv.", lut, height, width)
    skip = x_skip @ sd[prefix + "conv_block2.lin.mlp.weight"].t()
    skip = _bn(skip, sd, prefix + "conv_block2.norm_skip.")
    z = _bn(z, sd, prefix + "conv_block2.norm.")
    return act(z + skip)


# ----------------------------------------------------------------------------- Pooling
def consecutive_cluster(src):
    """layers/pooling.py:12-16."""
    unique, inv, counts = torch.unique(src, sorted=True, return_inverse=True, return_counts=True)
    perm = torch.arange(inv.size(0), dtype=inv.dtype)
    perm = inv.new_empty(unique.size(0)).scatter_(0, inv, perm)
    return unique, inv, perm, counts


def pooling(x, pos, batch, edge_index, voxel_size3, width, height, batch_size_ctor, transform_max, aggr="max",
            keep_temporal_ordering=False, self_loop=False):
    """Pooling.forward (layers/pooling.py:51-97), no bn.
    voxel_size3 = [sx, sy, 1] (pooling.py:24).  Returns a namespace with x,pos,batch,edge_index,edge_attr,cluster."""
    voxel_size = voxel_size3
    start = torch.Tensor([0, 0, 0, 0])
    end = torch.Tensor([0.9999999, 0.9999999, 0.9999999, batch_size_ctor - 1])
    wh_inv = 1 / torch.Tensor([[width, height]])
    pos4 = torch.cat([pos, batch.float().view(-1, 1)], dim=-1)
    raw_cluster = ops.grid_cluster(pos4, torch.cat([voxel_size, torch.Tensor([1])]), start, end)
    unique_clusters, cluster, perm, _ = consecutive_cluster(raw_cluster)
    e = cluster[edge_index]
    if self_loop:                                                     # pooling.py:59-60
        e = e.unique(dim=-1)
    else:
        e = e[:, e[0] != e[1]]
        if e.shape[1] > 0:
            e = e.unique(dim=-1)
    new_batch = batch[perm]
    new_pos = ops.scatter_mean(pos, cluster)
    if keep_temporal_ordering:                                        # pooling.py:69-72 (torch_scatter.scatter_max)
        t_max = torch.full((int(cluster.max()) + 1,), float("-inf"), dtype=pos.dtype)
        t_max = t_max.scatter_reduce(0, cluster, pos[:, -1], reduce="amax", include_self=True)
        t_src, t_dst = t_max[e]
        e = e[:, t_dst > t_src]
    if aggr == "max":
        new_x = ops.scatter_max(x, cluster)
    else:
        new_x = ops.scatter_mean(x, cluster)
    xy = new_pos[:, :2].clone()
    torch.div(xy + 1e-5, wh_inv, out=xy, rounding_mode="floor")      # round_to_pixel, pooling.py:47-49
    new_pos[:, :2] = xy * wh_inv
    if e.numel() > 0:
        edge_attr = ops.cartesian(new_pos, e, transform_max)
    else:
        edge_attr = torch.zeros((0, pos.shape[1]), dtype=pos.dtype)
    return SimpleNamespace(x=new_x, pos=new_pos, batch=new_batch, edge_index=e, edge_attr=edge_attr,
                           cluster=cluster, raw_cluster=raw_cluster, unique_clusters=unique_clusters)


# ----------------------------------------------------------------------------- image feature sampling
def sample_features(pos, batch, image_feat, width, height):
    """sample_features/_sample_features (networks/net.py:200-228)."""
    x = pos[:, 0] * width
    y = pos[:, 1] * height
    b = batch.float()
    x = 2 * x / (width - 1) - 1
    y = 2 * y / (height - 1) - 1
    bs = image_feat.shape[0]
    bs = bs if bs > 1 else 2
    b = 2 * b / (bs - 1) - 1
    grid = torch.stack((x, y, b), dim=-1).view(1, 1, 1, -1, 3)
    feat = image_feat.permute(1, 0, 2, 3).unsqueeze(0)
    out = torch.nn.functional.grid_sample(feat, grid=grid, mode="bilinear", align_corners=True)
    return out.view(feat.shape[1], -1).t()
