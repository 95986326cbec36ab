"""DAGR backbone `Net` with the reference's constructor, module tree and state-dict keys
(/root/reference/src/dagr/model/networks/net.py:15-228).

`forward` has two implementations with identical results:
  * the module-by-module composition of the reference (EV_TGN -> Cartesian -> Layer -> Pooling ...), every module
    calling the sm_100a kernels through the C ABI (used in training mode or with `fused=False`);
  * the fused eval path (`forward_fused`): one device-resident pipeline without host synchronisation until the
    outputs are read: graph CSR -> x0 assembly -> layer 1 (2 launches) -> pool (sampling fused) -> layer ... ;
    it is CUDA-graph capturable.
"""
import os

import torch

from ... import _lib, ops
from ...data import Batch, Data
from ..layers.components import Cartesian, CartesianTransform, two_m_of
from ..layers.conv import Layer
from ..layers.ev_tgn import EV_TGN, denormalize_pos
from ..layers.pooling import Pooling
from ..utils import shallow_copy
from .net_img import HookModule

# EVAD_SAMPLE_VISIT=0: the layer-1 input rows are produced in event order instead of pixel order (same rows either way)
SAMPLE_VISIT = os.environ.get("EVAD_SAMPLE_VISIT", "1") != "0"


def sampling_skip(data, image_feat):
    """net.py:15-17 through evad_sample_concat."""
    if data.batch is None or len(data.batch) != len(data.pos):
        data.batch = torch.zeros(len(data.pos), dtype=torch.long, device=data.x.device)
    return ops.sample_concat(data.x.float(), data.pos.float(), data.batch.int(), ops.to_channel_last(image_feat.float()),
                             float(data.width[0]), float(data.height[0]))


def sample_features(data, image_feat, image_sample_mode="bilinear"):
    """net.py:200-209."""
    if data.batch is None or len(data.batch) != len(data.pos):
        data.batch = torch.zeros(len(data.pos), dtype=torch.long, device=data.x.device)
    return ops.sample_concat(None, data.pos.float(), data.batch.int(), ops.to_channel_last(image_feat.float()),
                             float(data.width[0]), float(data.height[0]))


def compute_pooling_at_each_layer(pooling_dim_at_output, num_layers):
    py, px = map(int, pooling_dim_at_output.split("x"))
    pooling_base = torch.tensor([1.0 / px, 1.0 / py, 1.0 / 1])
    poolings = []
    for i in range(num_layers):
        pooling = pooling_base / 2 ** (3 - i)
        pooling[-1] = 1
        poolings.append(pooling)
    return torch.stack(poolings)


class GraphOut(Data):
    """Output of the fused path: `edge_index` (reference order of `unique(dim=-1)`) and `edge_attr` are materialised
    on first access from the dst-major CSR the kernels keep."""

    def __init__(self, **kw):
        self._lazy = None
        super().__init__(**kw)

    @property
    def edge_index(self):
        if self.__dict__.get("_edge_index") is None and self._lazy is not None:
            self._materialise()
        return self.__dict__.get("_edge_index")

    @edge_index.setter
    def edge_index(self, v):
        self.__dict__["_edge_index"] = v

    @property
    def edge_attr(self):
        if self.__dict__.get("_edge_attr") is None and self._lazy is not None:
            self._materialise()
        return self.__dict__.get("_edge_attr")

    @edge_attr.setter
    def edge_attr(self, v):
        self.__dict__["_edge_attr"] = v

    def keys(self):
        return [k for k in ("x", "edge_index", "edge_attr", "pos", "batch") if getattr(self, k) is not None]

    def _materialise(self):
        z = self._lazy
        n, e = self.x.shape[0], z["e"]
        if e > 0:
            ei = ops.csr_to_coo(z["rowptr"], z["src"], n, None, e, z["num_edges"], True)[:, :e]
            attr = ops.cartesian(self.pos, ei, z["two_m"], False)
        else:
            ei = torch.zeros((2, 0), dtype=torch.long, device=self.x.device)
            attr = torch.zeros((0, 3), dtype=torch.float32, device=self.x.device)
        self.__dict__["_edge_index"], self.__dict__["_edge_attr"] = ei, attr
        self._csr = dict(rowptr=z["rowptr"], src=z["src"], eid=None, edge_index_ptr=None)


class Net(torch.nn.Module):
    def __init__(self, args, height, width):
        super().__init__()
        channels = [1, int(args.base_width * 32), int(args.after_pool_width * 64), int(args.net_stem_width * 128),
                    int(args.net_stem_width * 128), int(args.net_stem_width * 128)]
        self.out_channels_cnn = []
        if args.use_image:
            import torchvision.models as tvm
            ctor = getattr(tvm, args.img_net)
            # the reference builds `img_net(pretrained=True)` (net.py:43), which downloads ImageNet weights; the
            # weights that matter come from the DAGR checkpoint (`backbone.net.module.*`), so default to no download
            module = ctor(weights="DEFAULT") if getattr(args, "img_net_pretrained", False) else ctor(weights=None)
            self.out_channels_cnn = [256, 256]
            self.net = HookModule(module, input_channels=3, height=height, width=width,
                                  feature_layers=["conv1", "layer1", "layer2", "layer3", "layer4"],
                                  output_layers=["layer3", "layer4"], feature_channels=channels[1:],
                                  output_channels=self.out_channels_cnn)
        self.use_image = args.use_image
        self.num_scales = args.num_scales
        self.num_classes = dict(dsec=2).get(args.dataset, 2)
        self.events_to_graph = EV_TGN(args)

        output_channels = channels[1:]
        self.out_channels = output_channels[-2:]
        input_channels = channels[:-1]
        if self.use_image:
            input_channels = [input_channels[i] + self.net.feature_channels[i] for i in range(len(input_channels))]

        poolings = compute_pooling_at_each_layer(args.pooling_dim_at_output, num_layers=4)
        max_vals_for_cartesian = 2 * poolings[:, :2].max(-1).values
        self.strides = torch.ceil(poolings[-2:, 1] * height).numpy().astype("int32").tolist()

        effective_radius = 2 * float(int(args.radius * width + 2) / width)
        self.edge_attrs = Cartesian(norm=True, cat=False, max_value=effective_radius)
        self.conv_block1 = Layer(2 + input_channels[0], output_channels[0], args=args)

        kw = dict(width=width, height=height, batch_size=args.batch_size,
                  keep_temporal_ordering=args.keep_temporal_ordering)
        cart1 = CartesianTransform(norm=True, cat=False, max_value=2 * effective_radius)
        self.pool1 = Pooling(poolings[0], transform=cart1, aggr=args.pooling_aggr, **kw)
        self.layer2 = Layer(input_channels[1] + 2, output_channels[1], args=args)
        cart2 = CartesianTransform(norm=True, cat=False, max_value=max_vals_for_cartesian[1])
        self.pool2 = Pooling(poolings[1], transform=cart2, aggr=args.pooling_aggr, **kw)
        self.layer3 = Layer(input_channels[2] + 2, output_channels[2], args=args)
        cart3 = CartesianTransform(norm=True, cat=False, max_value=max_vals_for_cartesian[2])
        self.pool3 = Pooling(poolings[2], transform=cart3, aggr=args.pooling_aggr, **kw)
        self.layer4 = Layer(input_channels[3] + 2, output_channels[3], args=args)
        cart4 = CartesianTransform(norm=True, cat=False, max_value=max_vals_for_cartesian[3])
        self.pool4 = Pooling(poolings[3], transform=cart4, aggr='mean', **kw)
        self.layer5 = Layer(input_channels[4] + 2, output_channels[4], args=args)
        self.cache = []
        self.fused = True
        self.height, self.width = height, width

    def get_output_sizes(self):
        poolings = [self.pool3.voxel_size[:2], self.pool4.voxel_size[:2]]
        return [(1 / p + 1e-3).cpu().int().numpy().tolist()[::-1] for p in poolings]

    # ------------------------------------------------------------------------------------------------ reference
    def forward(self, data, reset=True, image_feat=None):
        image_outputs = None
        if self.use_image and image_feat is None:
            image_feat, image_outputs = self.net(data.image)
        if hasattr(data, 'reset'):
            reset = data.reset
        if self.fused and not self.training and reset and self.fused_supported():
            output = self.forward_fused(data, image_feat)
        else:
            output = self.forward_modules(data, image_feat, reset)
        if self.use_image:
            return output[-self.num_scales:], (image_outputs[-self.num_scales:] if image_outputs is not None else None)
        return output[-self.num_scales:]

    def fused_supported(self):
        """The fused pipeline implements the configuration EventAD runs (pooling.py defaults).  Pooling options it does
        not implement (`keep_temporal_ordering` edge filter pooling.py:69-72, `self_loop`, a pooling BatchNorm, no
        transform) take the module-by-module path, which implements all of them."""
        return all(not p.keep_temporal_ordering and not p.self_loop and p.bn is None and p.transform is not None
                   for p in (self.pool1, self.pool2, self.pool3, self.pool4))

    def forward_modules(self, data, image_feat, reset=True):
        """Module-by-module composition, line by line net.py:113-197."""
        data = self.events_to_graph(data, reset=reset)
        if self.use_image:
            data.x = sampling_skip(data, image_feat[0].detach())
            data.skipped = True
            data.num_image_channels = image_feat[0].shape[1]
        data = self.edge_attrs(data)
        data.edge_attr = torch.clamp(data.edge_attr, min=0, max=1)
        data.x = torch.cat((data.x, data.pos[:, :2]), dim=1)
        data = self.conv_block1(data)
        pools = [self.pool1, self.pool2, self.pool3, self.pool4]
        layers = [self.layer2, self.layer3, self.layer4, self.layer5]
        out3 = None
        for k in range(4):
            if self.use_image:
                data.x = sampling_skip(data, image_feat[k + 1].detach())
            data = pools[k](data)
            if self.use_image:
                data.skipped = True
                data.num_image_channels = image_feat[k + 1].shape[1]
            data.x = torch.cat((data.x, data.pos[:, :2]), dim=1)
            data = layers[k](data)
            if k == 2:
                out3 = shallow_copy(data)
                out3.pooling = self.pool3.voxel_size[:3]
        out4 = data
        out4.pooling = self.pool4.voxel_size[:3]
        return [out3, out4]

    # ------------------------------------------------------------------------------------------------ fused
    def enqueue_fused(self, x, pos, batch_i32, num_graphs, feats_nhwc, width, height, time_window, scratch=None,
                      maps_ready=None):
        """Enqueues the whole event branch; returns device tensors sized by host-known upper bounds plus `counts`
        (int32[12]: n1,E1,n2,E2,n3,E3,n4,E4) and `status`.  No host synchronisation; capturable in a CUDA graph.
        `scratch` (a _lib.Scratch): private workspaces and status word of the caller (one per captured graph).
        `maps_ready` (callable(level)): called right before the first kernel that reads CNN map `level` is enqueued, for a
        caller that produces the maps on another stream (engine.GraphedEventAD(overlap=True) waits on an event there)."""
        self._maps_ready = maps_ready
        try:
            return self._enqueue_fused_scratch(x, pos, batch_i32, num_graphs, feats_nhwc, width, height, time_window, scratch)
        finally:
            self._maps_ready = None

    def _enqueue_fused_scratch(self, x, pos, batch_i32, num_graphs, feats_nhwc, width, height, time_window, scratch):
        g = self.events_to_graph
        if g.graph_creators is None:
            g.init_graph_creator(Data(pos=pos, width=[width], height=[height], time_window=[time_window],
                                      _num_graphs=num_graphs))
        gc = g.graph_creators
        if gc.device is None:
            gc.device = pos.device
            gc.initialize(pos.shape[0], pos.device)
        if scratch is None:
            return self._enqueue_fused(gc, x, pos, batch_i32, num_graphs, feats_nhwc, width, height, time_window)
        saved = (gc._ws, gc._status, ops._SCRATCH)
        gc._ws, gc._status, ops._SCRATCH = scratch.graph_ws, scratch.status(pos.device), scratch
        try:
            return self._enqueue_fused(gc, x, pos, batch_i32, num_graphs, feats_nhwc, width, height, time_window)
        finally:
            gc._ws, gc._status, ops._SCRATCH = saved

    def _enqueue_fused(self, gc, x, pos, batch_i32, num_graphs, feats_nhwc, width, height, time_window):
        dev = pos.device
        n = pos.shape[0]
        gc._status.zero_()                                   # the status word of THIS forward (kernels only OR bits in)
        gc.batch_size = max(gc.batch_size, num_graphs)
        gc.reset()
        pos_int = ops.denormalize_pos(pos, width, height, time_window)
        # When the fp32 LUT index of an integer pixel offset is provably offset + r (exhaustive host check, cached), the
        # spiral walk of the graph build already knows every edge's lattice class: no per-edge float pass is needed.
        c1 = self.conv_block1.conv_block1.conv
        ck = (int(width), int(height))
        if getattr(self, "_offcls_key", None) != ck:
            m = c1.mode()
            self._offcls_ok = bool(m["use_lut"] == 1 and m["nx"] == m["ny"] == 2 * int(gc.radius) + 1 and
                                   c1.l1_pack(self.conv_block1.conv_block2.lin.mlp.weight) is not None and
                                   ops.offset_class_is_lut_class(m, two_m_of(self.edge_attrs.max), True, width, height))
            self._offcls_key = ck
        cin1 = self.conv_block1.in_channel
        f = feats_nhwc if self.use_image else [None] * 5
        pseudo = dict(pos=pos, two_m=two_m_of(self.edge_attrs.max), clamp01=True)
        if self._offcls_ok and n > 0:
            # fixed-stride rows straight from the spiral search: layer 1 and pool 1 read them in place, so the fine graph
            # is never scanned / compacted into a CSR (the reference's `edges[:, edges[1] >= 0]`, graph/utils.py:23)
            graph = gc.build_rows(batch_i32, pos_int, want_cls=True)
            rowptr, src, ell = graph["count"], graph["nbr"].view(-1), graph["K"]
            cls1 = graph["nbr_cls"].view(-1)
        else:
            graph = gc.build_csr(batch_i32, pos_int, want_cells=True, want_cls=False)
            rowptr, src, ell, cls1 = graph["rowptr"], graph["src"], 0, None
        gc.max_index, gc.all_timestamps, gc._history = n, pos_int[:, 2], [(batch_i32, pos_int, 0)]   # AsyncGraph bookkeeping
        status = gc._status
        x0 = None
        mr = getattr(self, "_maps_ready", None)
        if mr is not None:
            mr(0)                                            # the first CNN map is read from here on
        if cls1 is not None and cin1 == 19 and self.conv_block1.h16_chain_ok(src, cls1, ell):
            # fp16-pair layer-1 kernel: x0 is produced as h2 words (ops.split_h2) by the sampling kernel itself
            # threads take the nodes in pixel order: neighbouring threads sample the same map cells
            x0 = ops.sample_concat_h2(x, pos, batch_i32, f[0], width, height, status,
                                      visit=graph["cell_perm"] if SAMPLE_VISIT else None)
        x0_h2 = x0 is not None
        if x0 is None:
            x0 = ops.sample_concat(x, pos, batch_i32, f[0], width, height, append_pos=True, ldo=(cin1 + 3) // 4 * 4)
        # layer 1 walks the nodes in pixel order: neighbouring destinations share their sources, so the row gathers
        # of a tile hit L1 instead of L2
        xk = self.conv_block1.run_fused(x0, rowptr, src, n, None, pseudo, status, perm=graph["cell_perm"], cls=cls1,
                                        ell_stride=ell, x_h2=x0_h2)
        cur = dict(x=xk, pos=pos, batch=batch_i32, rowptr=rowptr, src=src, n_cap=n, n_dev=None, e_cap=src.shape[0], ell=ell)
        return self._enqueue_tail(cur, f, num_graphs, width, height, status)

    def _enqueue_tail(self, cur, f, num_graphs, width, height, status, pool1=None):
        """pool1 .. layer5 on the layer-1 output `cur` (net.py:129-187); also the dense tail of the incremental update
        (asynchronous/incremental.py), where `cur` holds all resident nodes and only its newest rows were recomputed.
        pool1: optional callable(c_cap, e_cap, ldo, n_out, num_edges, status) that yields the first pooling's result in
        ops.voxel_pool's format (the incremental pooling state emits it instead of re-pooling all nodes)."""
        dev = cur["x"].device
        counts = torch.zeros(12, dtype=torch.int32, device=dev)
        pools = [self.pool1, self.pool2, self.pool3, self.pool4]
        layers = [self.layer2, self.layer3, self.layer4, self.layer5]
        outs = []
        for k in range(4):
            p = pools[k]
            key_bound = p._nv[0] * p._nv[1] * p._nv[2] * (num_graphs + 1)
            c_cap = max(1, min(cur["n_cap"], key_bound))
            e_cap = max(1, min(cur["e_cap"], c_cap * 64))
            cin_next = layers[k].in_channel
            mr = getattr(self, "_maps_ready", None)
            if mr is not None:
                mr(k + 1)
            if k == 0 and pool1 is not None:
                r = pool1(c_cap, e_cap, (cin_next + 3) // 4 * 4, counts[0:1], counts[1:2], status)
            else:
                r = ops.voxel_pool(cur["x"], cur["pos"], cur["batch"], cur["rowptr"], cur["src"], cur["n_cap"], cur["n_dev"],
                                   p._vs, p._nv, num_graphs, p._inv_wh, p.aggr, width, height, f[k + 1], append_pos=True,
                                   ldo=(cin_next + 3) // 4 * 4, c_cap=c_cap, e_cap=e_cap, status=status,
                                   n_out=counts[2 * k:2 * k + 1], num_edges=counts[2 * k + 1:2 * k + 2],
                                   ell_stride=cur.get("ell", 0))
            pseudo = dict(pos=r["pos"], two_m=two_m_of(p.transform.max), clamp01=False)
            xo = layers[k].run_fused(r["x"], r["rowptr"], r["src"], c_cap, r["n_out"], pseudo, status)
            cur = dict(x=xo, pos=r["pos"], batch=r["batch"], rowptr=r["rowptr"], src=r["src"], n_cap=c_cap,
                       n_dev=r["n_out"], e_cap=e_cap)
            if k >= 2:
                outs.append(dict(x=xo, pos=r["pos"], batch=r["batch"], rowptr=r["rowptr"], src=r["src"],
                                 num_edges=r["num_edges"], two_m=two_m_of(p.transform.max), k=k))
        # counts and status side by side, gathered as the last node of the enqueued work (inside a captured graph): the
        # caller's single device->host read needs no launch of its own
        return dict(outs=outs, counts=counts, status=status, readback=torch.cat([counts, status.view(-1)]))

    def finalize_fused(self, res):
        """One device->host read (counts + status), then tight views."""
        c = res["readback"].cpu().tolist()
        if c[12]:
            _lib.check_status(res["status"], "fused DAGR forward")
        datas = []
        pool_sizes = [self.pool3.voxel_size[:3], self.pool4.voxel_size[:3]]
        for o, ps in zip(res["outs"], pool_sizes):
            k = o["k"]
            nk, ek = c[2 * k], c[2 * k + 1]
            d = GraphOut(x=o["x"][:nk], pos=o["pos"][:nk], batch=o["batch"][:nk].long())
            d._lazy = dict(rowptr=o["rowptr"][:nk + 1], src=o["src"], num_edges=o["num_edges"], e=ek, two_m=o["two_m"])
            d.pooling = ps
            d.skipped = True
            datas.append(d)
        return datas

    def forward_fused(self, data, image_feat):
        n = data.pos.shape[0]
        dev = data.pos.device
        if data.batch is None:
            data.batch = torch.zeros(n, dtype=torch.long, device=dev)
        num_graphs = data.num_graphs
        feats = [ops.to_channel_last(fm.detach().float()) for fm in image_feat] if self.use_image else None
        res = self.enqueue_fused(data.x.float().contiguous(), data.pos.float().contiguous(), data.batch.int(), num_graphs,
                                 feats, int(data.width[0]), int(data.height[0]), int(data.time_window[0]))
        outs = self.finalize_fused(res)
        for o in outs:
            o.width, o.height = data.width, data.height
        return outs
