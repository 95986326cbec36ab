"""Pooling with the reference's constructor, buffers and behaviour (/root/reference/src/dagr/model/layers/pooling.py:
12-97), served by one fused call (csrc/pool.cu): grid_cluster + consecutive_cluster + coarse edges + pool_pos +
scatter_max/mean + round_to_pixel, then the Cartesian transform."""
from typing import Callable, List

import torch

from ... import _lib, ops
from ...data import Batch
from .components import BatchNormData
from .spline_conv import get_csr


class Pooling(torch.nn.Module):
    def __init__(self, size: List[float], width, height, batch_size, transform: Callable, aggr: str = 'max',
                 keep_temporal_ordering=False, dim=2, self_loop=False, in_channels=-1):
        super(Pooling, self).__init__()
        assert aggr in ['mean', 'max']
        self.aggr = aggr
        self.register_buffer("voxel_size", torch.cat([size, torch.Tensor([1])]), persistent=False)
        self.transform = transform
        self.keep_temporal_ordering = keep_temporal_ordering
        self.dim = dim
        self.register_buffer("start", torch.Tensor([0, 0, 0, 0]), persistent=False)
        self.register_buffer("end", torch.Tensor([0.9999999, 0.9999999, 0.9999999, batch_size - 1]), persistent=False)
        self.register_buffer("wh_inv", 1 / torch.Tensor([[width, height]]), persistent=False)
        self.max_num_voxels = batch_size * self.num_grid_cells
        self.register_buffer("sorted_cluster", torch.arange(self.max_num_voxels), persistent=False)
        self.self_loop = self_loop
        self.width, self.height = width, height
        self.bn = None
        if in_channels > 0:
            self.bn = BatchNormData(in_channels)
        # host copies of the geometry (fp32 values exactly as torch_cluster sees them)
        self._vs = [float(v) for v in self.voxel_size]
        self._nv = ops.pool_geometry(self.voxel_size, self.end)
        self._inv_wh = [float(self.wh_inv[0, 0]), float(self.wh_inv[0, 1])]

    @property
    def num_grid_cells(self):
        return (1 / self.voxel_size + 1e-3).int().prod()

    def round_to_pixel(self, pos, wh_inv):
        torch.div(pos + 1e-5, wh_inv, out=pos, rounding_mode='floor')
        return pos * wh_inv

    # device-resident call used by the fused Net.forward
    def run(self, x, pos, batch_i32, rowptr, src, n_cap, n_dev, batch_size, feat_nhwc=None, append_pos=False, ldo=None,
            status=None, e_cap=None, width=None, height=None):
        return ops.voxel_pool(x, pos, batch_i32, rowptr, src, n_cap, n_dev, self._vs, self._nv, batch_size, self._inv_wh,
                              self.aggr, width or self.width, height or self.height, feat_nhwc, append_pos, ldo,
                              e_cap=e_cap, status=status, self_loop=self.self_loop)

    def forward(self, data):
        if data.x.shape[0] == 0:
            return data
        n = data.x.shape[0]
        dev = data.x.device
        x = data.x.float()
        pos = data.pos.float()
        batch = data.batch if data.batch is not None else torch.zeros(n, dtype=torch.long, device=dev)
        batch_i32 = batch.int()
        bsz = int(batch.max().item()) + 1
        if data.edge_index is not None and data.edge_index.numel() > 0:
            csr = get_csr(data)
            rowptr, src = csr["rowptr"], csr["src"]
        else:
            rowptr = torch.zeros(n + 1, dtype=torch.int32, device=dev)
            src = torch.zeros(1, dtype=torch.int32, device=dev)
        r = self.run(x, pos, batch_i32, rowptr, src, n, None, bsz)
        n_out = int(r["n_out"].item())
        e = int(r["num_edges"].item())
        _lib.check_status(r["status"], "voxel pooling")
        ei, slot_of = ops.csr_to_coo(r["rowptr"], r["src"], r["c_cap"], r["n_out"], r["e_cap"], r["num_edges"], True, True)
        edge_index = ei[:, :e]
        new_data = Batch(batch=r["batch"][:n_out].long(), x=r["x"][:n_out], edge_index=edge_index, pos=r["pos"][:n_out])
        eid = torch.empty(max(e, 1), dtype=torch.int32, device=dev)
        eid[slot_of[:e].long()] = torch.arange(e, dtype=torch.int32, device=dev)
        new_data._csr = dict(rowptr=r["rowptr"][:n_out + 1], src=r["src"][:max(e, 1)], eid=eid[:max(e, 1)],
                             edge_index_ptr=(edge_index.data_ptr(), tuple(edge_index.shape)))
        new_data.cluster = r["cluster"][:n].long()
        if self.keep_temporal_ordering:
            # pooling.py:69-72: keep the coarse edges whose destination voxel holds a later event than the source voxel
            t_max = torch.full((max(n_out, 1),), float("-inf"), dtype=pos.dtype, device=dev)
            t_max.scatter_reduce_(0, new_data.cluster, pos[:, -1].contiguous(), reduce="amax", include_self=True)
            if e > 0:
                t_src, t_dst = t_max[edge_index]
                new_data.edge_index = edge_index[:, t_dst > t_src]
                del new_data._csr                            # the cached CSR describes the unfiltered edges
        if hasattr(data, "height"):
            new_data.height = data.height
            new_data.width = data.width
        if self.transform is not None:
            if new_data.edge_index.numel() > 0:
                new_data = self.transform(new_data)
            else:
                new_data.edge_attr = torch.zeros(size=(0, pos.shape[1]), dtype=pos.dtype, device=pos.device)
        if self.bn is not None:
            new_data = self.bn(new_data)
        return new_data
