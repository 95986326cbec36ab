"""oracle/ref_py.py -- runs the reference's OWN Python text for the hot path on the CPU.

TEST INFRASTRUCTURE ONLY (imported by tests/ and by tests/golden/make_ref_py_golden.py; never by eventad_b200/).

The reference's event branch (`/root/reference/src/dagr/model/{layers,networks}`, `graph/`, `model/utils.py`,
`utils/buffers.py`) is pure torch once six things exist that this container lacks: `torch_geometric`,
`torch_scatter`, `torch_cluster`, `torch_spline_conv`, `yolox` and the compiled `ev_graph_cuda` extension.
`install()` registers stand-ins for exactly those in `sys.modules` and puts `/root/reference/src` on `sys.path`;
after that `import dagr.model.networks.dagr` executes the reference files where they lie, UNMODIFIED, and
`DAGR(args, h, w)`, `.cache_luts(...)`, `.backbone(data)`, `Pooling.forward`, `MySplineConv.init_lut / message_lut /
_forward`, `consecutive_cluster`, `round_to_pixel`, `sample_features`, `compute_pooling_at_each_layer`,
`denormalize_pos`, `EV_TGN`, `AsyncGraph` / `SlidingWindowGraph`, `_insert_events_into_queue`, `_search_for_edges`,
`format_data`, `voxel_size_to_params`, `shallow_copy` are the reference's code, not a retyped copy.

What the stand-ins are (the ONLY restated arithmetic; everything else is the reference's text):
  torch_spline_conv.spline_basis / spline_weighting   -> oracle/dagr_ops.py   (library 1.2.2)  UNPINNED
  torch_cluster.grid_cluster                          -> oracle/dagr_ops.py   (library 1.6.3)    UNPINNED
  torch_scatter.scatter_max / scatter_sum             -> oracle/dagr_ops.py   (library 2.1.2)    UNPINNED
      (with `out=` both continue IN PLACE and scatter_max returns the argmax: what dagr/asynchronous relies on)
  torch_geometric: SplineConv (parameters + propagate = gather / message / scatter-add), BatchNorm (.module =
      BatchNorm1d), T.Cartesian, pool_pos, _avg_pool_x, Data / Batch (attribute bags)  (2.6.1)  UNPINNED
  ev_graph_cuda.{fill_edges_cuda, insert_in_queue_cuda, insert_in_queue_single_cuda} on CPU tensors
      -> oracle/evg_oracle.c kernel-level entry points; PINNED: the same C bodies are checked on the B200 against the
         unmodified ev_graph.cu (tests/test_ref_kernel_gpu.py) and against its golden outputs (tests/golden/ref_graph_*.npz)
  asy_tools.{masked_lin, masked_lin_no_bias, masked_isdiff, masked_inplace_BN} on CPU tensors
      -> oracle/asy_oracle.c; PINNED on the B200 against the unmodified asy_tools/main.cu (tests/test_asy_gpu.py)
  yolox.models.{YOLOX, YOLOXHead, IOUloss}            -> empty nn.Module shells (the head is never run by EventAD)
  torchvision resnet50(pretrained=True)               -> same constructor with weights=None (no network)

Nothing here runs on the GPU box: /root/reference does not exist there.  tests/golden/make_ref_py_golden.py uses this
module to write the committed fixtures tests/golden/ref_py_*.npz.
"""
import copy
import ctypes
import os
import sys
import types

import torch

from . import dagr_ops, evg

REF_ROOT = "/root/reference"
REF_SRC = os.path.join(REF_ROOT, "src")
_installed = False


def available() -> bool:
    return os.path.isdir(os.path.join(REF_SRC, "dagr", "model", "layers"))


# ------------------------------------------------------------------------------------------------ torch_geometric
def _pyg_modules():
    from eventad_b200.data import Batch, Data          # the product's attribute bags double as PyG's Data / Batch

    class PyGLinear(torch.nn.Module):
        """torch_geometric.nn.dense.linear.Linear(in, out, bias=False): parameter `weight` [out, in]."""

        def __init__(self, in_channels, out_channels, bias=False, weight_initializer=None):
            super().__init__()
            self.in_channels, self.out_channels = in_channels, out_channels
            self.weight = torch.nn.Parameter(torch.empty(out_channels, in_channels))
            self.register_parameter("bias", None)
            torch.nn.init.uniform_(self.weight, -1.0, 1.0)

        def forward(self, x):
            return torch.nn.functional.linear(x, self.weight)

    class SplineConv(torch.nn.Module):
        """torch_geometric.nn.conv.SplineConv 2.6.1: parameters / buffers and MessagePassing.propagate (flow
        source_to_target: x_j = x[edge_index[0]], message, sum into edge_index[1] with dim_size = size[1])."""

        def __init__(self, in_channels, out_channels, dim, kernel_size, is_open_spline=True, degree=1, aggr="mean",
                     root_weight=True, bias=True, **kwargs):
            super().__init__()
            self.in_channels, self.out_channels, self.dim, self.degree = in_channels, out_channels, dim, degree
            self.aggr, self.root_weight = aggr, root_weight
            ks = torch.tensor([kernel_size] * dim if isinstance(kernel_size, int) else list(kernel_size), dtype=torch.long)
            self.register_buffer("kernel_size", ks)
            iso = [is_open_spline] * dim if isinstance(is_open_spline, bool) else list(is_open_spline)
            self.register_buffer("is_open_spline", torch.tensor(iso, dtype=torch.uint8))
            self.K = int(ks.prod())
            self.weight = torch.nn.Parameter(torch.empty(self.K, in_channels, out_channels))
            if root_weight:
                self.lin = PyGLinear(in_channels, out_channels, bias=False)
            if bias:
                self.bias = torch.nn.Parameter(torch.zeros(out_channels))
            else:
                self.register_parameter("bias", None)
            torch.nn.init.uniform_(self.weight, -1.0, 1.0)

        def message(self, x_j, edge_attr):
            basis, index = dagr_ops.spline_basis(edge_attr, self.kernel_size, self.is_open_spline, self.degree)
            return dagr_ops.spline_weighting(x_j, self.weight, basis, index)

        def propagate(self, edge_index, x=None, edge_attr=None, size=None):
            assert self.aggr in ("sum", "add"), "the EventAD configuration aggregates with sum (SURVEY A.1)"
            x_src = x[0] if isinstance(x, (tuple, list)) else x
            msg = self.message(x_src.index_select(0, edge_index[0]), edge_attr)
            n = size[1] if size is not None else x_src.shape[0]
            return dagr_ops.scatter_sum(msg, edge_index[1], n)

        def aggregate(self, inputs, index, ptr=None, dim_size=None):
            """MessagePassing.aggregate with aggr='sum' (called directly by asynchronous/conv.py:14)."""
            assert self.aggr in ("sum", "add")
            return dagr_ops.scatter_sum(inputs, index, dim_size)

    class GCNConv(torch.nn.Module):
        """Only named in an isinstance test of asynchronous/__init__.py; never constructed by EventAD."""

    class BatchNorm(torch.nn.Module):
        """torch_geometric.nn.BatchNorm: wrapper whose `.module` is nn.BatchNorm1d."""

        def __init__(self, in_channels, eps=1e-5, momentum=0.1, affine=True, track_running_stats=True, **kw):
            super().__init__()
            self.module = torch.nn.BatchNorm1d(in_channels, eps, momentum, affine, track_running_stats)

        def forward(self, x):
            return self.module(x)

    class Cartesian:
        """torch_geometric.transforms.Cartesian; BaseTransform.__call__ works on a shallow copy."""

        def __init__(self, norm=True, max_value=None, cat=True, interval=(0.0, 1.0)):
            self.norm, self.max, self.cat, self.interval = norm, max_value, cat, interval

        def __call__(self, data):
            return self.forward(copy.copy(data))

        def forward(self, data):
            pseudo = data.edge_attr
            cart = data.pos[data.edge_index[0]] - data.pos[data.edge_index[1]]
            cart = cart.view(-1, 1) if cart.dim() == 1 else cart
            if self.norm and cart.numel() > 0:
                max_value = cart.abs().max() if self.max is None else self.max
                cart = cart * dagr_ops.cuda_scalar_reciprocal(2 * max_value) + 0.5    # the division as CUDA does it
            if pseudo is not None and self.cat:
                pseudo = pseudo.view(-1, 1) if pseudo.dim() == 1 else pseudo
                data.edge_attr = torch.cat([pseudo, cart.type_as(pseudo)], dim=-1)
            else:
                data.edge_attr = cart
            return data

    class ToSparseTensor:
        def __init__(self, *a, **k):
            pass

        def __call__(self, data):
            raise NotImplementedError("the `reproducible` (SparseTensor) route of MySplineConv is off in EventAD")

    def pool_pos(cluster, pos):
        return dagr_ops.scatter_mean(pos, cluster)

    def _avg_pool_x(cluster, x, size=None):
        return dagr_ops.scatter_mean(x, cluster, size)

    def mod(name, **attrs):
        m = types.ModuleType(name)
        m.__dict__.update(attrs)
        m.__path__ = []
        return m

    tg = mod("torch_geometric")
    data = mod("torch_geometric.data", Data=Data, Batch=Batch)
    nn_conv = mod("torch_geometric.nn.conv", SplineConv=SplineConv, GCNConv=GCNConv)
    avg = mod("torch_geometric.nn.pool.avg_pool", _avg_pool_x=_avg_pool_x)
    pool = mod("torch_geometric.nn.pool.pool", pool_pos=pool_pos)
    nn_pool = mod("torch_geometric.nn.pool", avg_pool=avg, pool=pool)
    nn_norm = mod("torch_geometric.nn.norm", BatchNorm=BatchNorm)
    nn = mod("torch_geometric.nn", BatchNorm=BatchNorm, conv=nn_conv, pool=nn_pool, SplineConv=SplineConv, norm=nn_norm)
    tst = mod("torch_geometric.transforms.to_sparse_tensor", ToSparseTensor=ToSparseTensor)
    tr = mod("torch_geometric.transforms", Cartesian=Cartesian, to_sparse_tensor=tst)
    tg.data, tg.nn, tg.transforms = data, nn, tr
    return {m.__name__: m for m in (tg, data, nn, nn_conv, nn_norm, nn_pool, avg, pool, tr, tst)}


# ------------------------------------------------------------------------------------------------ ev_graph_cuda on the CPU
def _ev_graph_module():
    L = evg.lib()
    c_p, c_i, c_i64, c_f = ctypes.c_void_p, ctypes.c_int, ctypes.c_int64, ctypes.c_float
    L.evgo_k_insert.argtypes = [c_p, c_i, c_i, c_i, c_i, c_p, c_p, c_p, c_i64]
    L.evgo_k_insert_single.argtypes = [c_p, c_i, c_i, c_i, c_p, c_p]
    L.evgo_k_fill_edges.argtypes = [c_p, c_p, c_p, c_p, c_p, c_i64, c_i, c_i, c_i, c_i, c_f, c_f, c_p, c_i64, c_i]

    def _chk(t, dtype):
        assert t.device.type == "cpu" and t.is_contiguous() and t.dtype == dtype, (t.dtype, t.is_contiguous())
        return t.data_ptr()

    def insert_in_queue_cuda(indices, unique_coords, cumsum_counts, queue):
        B, Q, H, W = queue.shape
        uc = unique_coords.int().contiguous()               # the kernel is instantiated for int32 (ev_graph.cu:259)
        L.evgo_k_insert(_chk(queue, torch.int32), B, Q, H, W, _chk(indices.contiguous(), torch.int32),
                        _chk(uc, torch.int32), _chk(cumsum_counts.contiguous(), torch.int32), uc.shape[0])
        return queue

    def insert_in_queue_single_cuda(indices, events, queue):
        B, Q, H, W = queue.shape
        L.evgo_k_insert_single(_chk(queue, torch.int32), Q, H, W, _chk(indices.int().contiguous(), torch.int32),
                               _chk(events.int().contiguous(), torch.int32))
        return queue

    def fill_edges_cuda(batch, pos, all_timestamps, event_queue, indices, max_num_neighbors, radius, delta_t_us, edges,
                        min_index):
        B, Q, H, W = event_queue.shape
        n = batch.shape[0]
        assert edges.shape[0] == 2 and edges.is_contiguous()
        L.evgo_k_fill_edges(_chk(batch, torch.int32), _chk(pos.contiguous(), torch.int32),
                            _chk(all_timestamps, torch.int32), _chk(event_queue, torch.int32),
                            _chk(indices.contiguous(), torch.int32), n, Q, H, W, int(max_num_neighbors), float(radius),
                            float(delta_t_us), _chk(edges, torch.int64), edges.shape[1], int(min_index))

    m = types.ModuleType("ev_graph_cuda")
    m.insert_in_queue_cuda = insert_in_queue_cuda
    m.insert_in_queue_single_cuda = insert_in_queue_single_cuda
    m.fill_edges_cuda = fill_edges_cuda
    return m


def _asy_tools_module():
    """CPU stand-in of the reference's `asy_tools` pybind module (asy_tools/main.cu:239-244) over oracle/asy_oracle.c."""
    A = evg.asy_lib()

    def _p(t, dtype):
        assert t.device.type == "cpu" and t.is_contiguous() and t.dtype == dtype, (t.dtype, t.is_contiguous())
        return t.data_ptr()

    def masked_isdiff(indices, x_old, x_new, atol, rtol):
        A.asyo_masked_isdiff(_p(indices, torch.int64), _p(x_old, torch.float32), _p(x_new, torch.float32),
                             indices.shape[0], x_old.shape[1], float(atol), float(rtol))
        return indices[indices > -1]

    def masked_inplace_BN(indices, x, x_out, running_mean, running_var, weight, bias, eps):
        A.asyo_masked_inplace_bn(_p(indices, torch.int64), _p(x, torch.float32), _p(x_out, torch.float32),
                                 _p(running_mean, torch.float32), _p(running_var, torch.float32),
                                 _p(weight.detach(), torch.float32), _p(bias.detach(), torch.float32), indices.shape[0],
                                 x.shape[1], float(eps))

    def masked_lin(indices, x_in, x_out, weight, bias, add):
        A.asyo_masked_lin(_p(indices, torch.int64), _p(x_in, torch.float32), _p(x_out, torch.float32),
                          _p(weight.detach(), torch.float32), _p(bias.detach(), torch.float32), indices.shape[0],
                          weight.shape[1], weight.shape[0], int(bool(add)))

    def masked_lin_no_bias(indices, x_in, x_out, weight, add):
        A.asyo_masked_lin(_p(indices, torch.int64), _p(x_in, torch.float32), _p(x_out, torch.float32),
                          _p(weight.detach(), torch.float32), None, indices.shape[0], weight.shape[1], weight.shape[0],
                          int(bool(add)))

    m = types.ModuleType("asy_tools")
    m.masked_isdiff, m.masked_inplace_BN, m.masked_lin, m.masked_lin_no_bias = (masked_isdiff, masked_inplace_BN,
                                                                                masked_lin, masked_lin_no_bias)
    return m


def _misc_modules():
    def scatter_max(src, index, dim=0, out=None, dim_size=None):
        """torch_scatter 2.1.2 scatter_max (csrc/cpu/scatter_cpu.cpp): values and argmax.  With `out` the reduction
        continues from its contents IN PLACE and an entry is replaced only by a strictly greater value; argmax is the
        FIRST source row that holds the maximum, src.size(0) where no row replaced the initial value (the CPU kernel's
        order; the CUDA kernel's tie choice is a race).  Without `out`, untouched entries are 0."""
        assert dim == 0
        if out is None:
            return dagr_ops.scatter_max(src, index, dim_size), None      # dense path: the argmax is never read
        n_src = src.shape[0]
        idx = index.view(-1, *([1] * (src.dim() - 1))).expand_as(src)
        before = out.clone()
        out.scatter_reduce_(0, idx, src, reduce="amax", include_self=True)
        rows = torch.arange(n_src, device=src.device).view(-1, *([1] * (src.dim() - 1))).expand_as(src)
        cand = torch.where(src == out.gather(0, idx), rows, torch.full_like(rows, n_src))
        arg = torch.full(out.shape, n_src, dtype=torch.long, device=src.device)
        arg.scatter_reduce_(0, idx, cand, reduce="amin", include_self=True)
        arg[~(out > before)] = n_src
        return out, arg

    def scatter_sum(src, index, dim=0, out=None, dim_size=None):
        """torch_scatter 2.1.2 scatter_sum: with `out`, adds into it IN PLACE (the asynchronous layers rely on that)."""
        assert dim == 0
        if out is not None:
            idx = index.view(-1, *([1] * (src.dim() - 1))).expand_as(src)
            return out.scatter_add_(0, idx, src)
        n = dim_size if dim_size is not None else (int(index.max()) + 1 if index.numel() else 0)
        return dagr_ops.scatter_sum(src, index, n)

    ts = types.ModuleType("torch_scatter")
    ts.scatter_max, ts.scatter_sum = scatter_max, scatter_sum
    tc = types.ModuleType("torch_cluster")
    tc.grid_cluster = lambda pos, size, start=None, end=None: dagr_ops.grid_cluster(pos, size, start, end)
    tsc = types.ModuleType("torch_spline_conv")
    tsc.spline_basis, tsc.spline_weighting = dagr_ops.spline_basis, dagr_ops.spline_weighting

    class YOLOX(torch.nn.Module):
        def __init__(self, backbone=None, head=None):
            super().__init__()
            self.backbone, self.head = backbone, head

    class YOLOXHead(torch.nn.Module):
        def __init__(self, num_classes=80, width=1.0, strides=(8, 16, 32), in_channels=(256, 512, 1024), act="silu",
                     depthwise=False):
            super().__init__()

    class IOUloss(torch.nn.Module):
        def __init__(self, reduction="none", loss_type="iou"):
            super().__init__()

    yx = types.ModuleType("yolox")
    yx.__path__ = []
    ym = types.ModuleType("yolox.models")
    ym.YOLOX, ym.YOLOXHead, ym.IOUloss = YOLOX, YOLOXHead, IOUloss
    yx.models = ym
    coco = types.ModuleType("dagr.utils.coco_eval")
    coco.evaluate_detection = None
    return {"torch_scatter": ts, "torch_cluster": tc, "torch_spline_conv": tsc, "yolox": yx, "yolox.models": ym,
            "dagr.utils.coco_eval": coco}


def install():
    """Registers the stand-ins and makes `import dagr...` / `import models.EventAD` resolve to the reference tree."""
    global _installed
    if _installed:
        return
    if not available():
        raise FileNotFoundError(f"{REF_SRC}/dagr not present (the reference is only mounted in the build container)")
    mods = {}
    mods.update(_pyg_modules())
    mods.update(_misc_modules())
    mods["ev_graph_cuda"] = _ev_graph_module()
    mods["asy_tools"] = _asy_tools_module()
    for k, v in mods.items():
        assert k not in sys.modules or getattr(sys.modules[k], "__file__", None) is None, f"real {k} is importable: use it"
        sys.modules[k] = v
    import torchvision.models as tvm
    for name in ("resnet18", "resnet34", "resnet50"):
        ctor = getattr(tvm, name)
        if getattr(ctor, "_evad_offline", False):
            continue

        def offline(pretrained=False, _ctor=ctor, **kw):   # net.py:41-43 asks for ImageNet weights: no network here
            kw.pop("weights", None)
            return _ctor(weights=None, **kw)
        offline._evad_offline = True
        setattr(tvm, name, offline)
    for p in (REF_SRC, REF_ROOT):
        if p not in sys.path:
            sys.path.insert(0, p)
    _installed = True


# ------------------------------------------------------------------------------------------------ convenience drivers
def build_dagr(args, height, width, state_dict=None, cache_luts=True, skip_head_luts=True):
    """The reference's DAGR(args, height, width) (dagr.py:14-34) with `state_dict` (keys `backbone.*`) loaded strictly
    into its backbone, then `cache_luts` (dagr.py:36-71) as utils/model.py:34-39 calls it.  skip_head_luts: the
    detection head's LUTs (never read by EventAD) are not materialised."""
    install()
    import argparse
    from dagr.model.networks.dagr import DAGR
    ns = argparse.Namespace(**vars(args))
    model = DAGR(ns, height=height, width=width)
    if state_dict is not None:
        bb = {k[len("backbone."):]: v for k, v in state_dict.items() if k.startswith("backbone.")}
        model.backbone.load_state_dict(bb, strict=True)
    model.eval()
    if cache_luts:
        if skip_head_luts:
            for name in ("stem1", "cls_conv1", "reg_conv1", "stem2", "cls_conv2", "reg_conv2"):
                getattr(model.head, name).conv.init_lut = lambda **k: None
            for name in ("cls_pred1", "reg_pred1", "obj_pred1", "cls_pred2", "reg_pred2", "obj_pred2"):
                getattr(model.head, name).init_lut = lambda **k: None
        radius = args.radius
        with torch.no_grad():
            model.cache_luts(radius=radius, height=height, width=width)     # utils/model.py:36-39 argument order
    return model


def run_backbone(model, data, image_feat=None):
    """Net.forward of the reference (net.py:106-197) on a CPU Batch.  image_feat: when given, the HookModule is
    replaced for this call by one that returns these maps (the event branch is what is under test)."""
    net = model.backbone
    saved = net.net if getattr(net, "use_image", False) else None
    if image_feat is not None:
        class _Feats(torch.nn.Module):
            def forward(self, image):
                return list(image_feat), [image_feat[-2], image_feat[-1]]
        net.net = _Feats()
        if not hasattr(data, "image") or data.image is None:
            data.image = torch.zeros(1)
    try:
        with torch.no_grad():
            out = net(data)
    finally:
        if saved is not None:
            net.net = saved
    return out[0] if getattr(net, "use_image", False) else out
