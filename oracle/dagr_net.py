"""oracle/dagr_net.py -- torch-CPU restatement of the GNN branch of `Net.forward`.

TEST INFRASTRUCTURE ONLY.  Follows /root/reference/src/dagr/model/networks/net.py:106-197 (the event branch;
the ResNet feature maps `image_feat` are an input), with EV_TGN from layers/ev_tgn.py:11-58 driving the C
graph-builder oracle (oracle/evg_oracle.c), and LUT-mode convolutions as `DAGR.cache_luts` configures them
(networks/dagr.py:36-71; EventAD always calls it: utils/model.py:34-39).
"""
from types import SimpleNamespace

import numpy as np
import torch

from . import dagr_layers as L
from . import dagr_ops as ops
from .evg import OracleSlidingWindowGraph


def compute_pooling_at_each_layer(pooling_dim_at_output, num_layers):
    """net.py:19-28."""
    py, px = map(int, pooling_dim_at_output.split("x"))
    pooling_base = torch.tensor([1.0 / px, 1.0 / py, 1.0 / 1])
    poolings = []
    for i in range(num_layers):
        pooling = pooling_base / 2 ** (3 - i)
        pooling[-1] = 1
        poolings.append(pooling)
    return torch.stack(poolings)


def denormalize_pos(pos, width, height, time_window):
    """ev_tgn.py:11-16."""
    denorm = torch.tensor([int(width), int(height), int(time_window)])
    return (denorm.view(1, -1) * pos + 1e-3).int()


def events_to_graph(pos, batch, num_graphs, radius, max_neighbors, width, height, time_window, max_queue_size=128):
    """EV_TGN.init_graph_creator + forward(reset=True) (ev_tgn.py:27-58). Returns edge_index int64 [2,E]."""
    delta_t_us = int(radius * time_window)
    r = int(radius * width + 1)
    g = OracleSlidingWindowGraph(width=width, height=height, max_num_neighbors=max_neighbors,
                                 max_queue_size=max_queue_size, batch_size=num_graphs, radius=r,
                                 delta_t_us=delta_t_us)
    p = denormalize_pos(pos, width, height, time_window)
    e = g.forward(batch.int().numpy(), p.numpy(), delete_nodes=False, collect_edges=True)
    return torch.from_numpy(np.ascontiguousarray(e)).long()


def net_forward(sd, args, x, pos, batch, num_graphs, image_feat, width, height, time_window=1000000,
                use_lut=True, return_intermediate=False, chunk_front_by_sample=False):
    """Event branch of Net.forward (net.py:106-197).  sd: state dict with keys relative to Net.
    x [N,1] polarity, pos [N,3] normalised, batch [N] int64, image_feat: list of 5 maps [B,C,h,w] or None.
    Returns [out3, out4] namespaces (x,pos,batch,edge_index,edge_attr)."""
    use_image = image_feat is not None
    poolings = compute_pooling_at_each_layer(args.pooling_dim_at_output, num_layers=4)
    max_vals = 2 * poolings[:, :2].max(-1).values
    effective_radius = 2 * float(int(args.radius * width + 2) / width)
    inter = {}

    lut = L.lut_params_layer1(args.radius, width) if use_lut else None

    def front(x_, pos_, batch_, graph_batch, n_graphs):
        """graph + CNN sampling + Cartesian + Layer 1 (net.py:113-124) on a set of whole samples."""
        ei = events_to_graph(pos_, graph_batch, n_graphs, args.radius, args.max_neighbors, width, height, time_window)
        if use_image:
            x_ = torch.cat((x_, L.sample_features(pos_, batch_, image_feat[0], width, height)), dim=1)
        ea = ops.cartesian(pos_, ei, effective_radius) if ei.shape[1] > 0 else torch.zeros((0, 3), dtype=x_.dtype)
        ea = torch.clamp(ea, min=0, max=1)
        x_ = torch.cat((x_, pos_[:, :2]), dim=1)
        return ei, ea, L.layer(x_, ei, ea, sd, "conv_block1.", lut, height, width)

    if chunk_front_by_sample and num_graphs > 1:
        # Everything up to pool 1 is independent per sample (graph, sampling and convolution only see a sample's own
        # events; sampling reads the sample's batch index, which is kept), so the samples are processed one at a time:
        # same numbers, bounded memory at the 64 x 100k bench batch.
        eis, xs, off = [], [], 0
        for b in range(num_graphs):
            m = batch == b
            nb = int(m.sum())
            ei, ea, xb = front(x[m], pos[m], batch[m], torch.zeros(nb, dtype=batch.dtype), 1)
            eis.append(ei + off); xs.append(xb)           # the fine edge_attr is not read after layer 1
            off += nb
        assert bool((batch[1:] >= batch[:-1]).all()), "samples must be concatenated sample-major"
        edge_index, edge_attr, x = torch.cat(eis, 1), torch.zeros((0, 3), dtype=x.dtype), torch.cat(xs, 0)
    else:
        edge_index, edge_attr, x = front(x, pos, batch, batch, num_graphs)
    inter["edge_index0"] = edge_index
    inter["x_layer1"] = x

    pool_cart_max = [2 * effective_radius, max_vals[1], max_vals[2], max_vals[3]]
    pool_aggr = [args.pooling_aggr, args.pooling_aggr, args.pooling_aggr, "mean"]
    names = ["layer2.", "layer3.", "layer4.", "layer5."]
    outs = []
    cur = SimpleNamespace(x=x, pos=pos, batch=batch, edge_index=edge_index, edge_attr=edge_attr)
    for k in range(4):
        if use_image:
            cur.x = torch.cat((cur.x, L.sample_features(cur.pos, cur.batch, image_feat[k + 1], width, height)), dim=1)
        voxel3 = poolings[k]
        if cur.x.shape[0] > 0:
            cur = L.pooling(cur.x, cur.pos, cur.batch, cur.edge_index, voxel3, width, height, args.batch_size,
                            pool_cart_max[k], pool_aggr[k],
                            keep_temporal_ordering=getattr(args, "keep_temporal_ordering", False))
        inter[f"pool{k + 1}"] = cur
        lut = L.lut_params_from_pool(torch.cat([voxel3, torch.Tensor([1])]), pool_cart_max[k], height, width) \
            if use_lut else None
        xin = torch.cat((cur.x, cur.pos[:, :2]), dim=1)
        xo = L.layer(xin, cur.edge_index, cur.edge_attr, sd, names[k], lut, height, width)
        cur = SimpleNamespace(x=xo, pos=cur.pos, batch=cur.batch, edge_index=cur.edge_index, edge_attr=cur.edge_attr)
        inter[f"x_layer{k + 2}"] = xo
        if k >= 2:
            outs.append(SimpleNamespace(x=xo.clone(), pos=cur.pos, batch=cur.batch, edge_index=cur.edge_index,
                                        edge_attr=cur.edge_attr, pooling=voxel3))
    if return_intermediate:
        return outs, inter
    return outs
