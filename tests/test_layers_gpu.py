"""Layer-level parity of the sm_100a kernels with the torch-CPU oracle (oracle/dagr_layers.py, oracle/dagr_ops.py),
through the reference-shaped modules.  Integer results (cluster ids, coarse edges, batch) are compared bit-exactly;
fp32 features within rtol 1e-4 / atol 1e-5 (BASELINE.json north_star)."""
import numpy as np
import pytest
import torch

from tests.util import assert_close, make_backbone, make_data, net_state_dict

pytestmark = pytest.mark.gpu


def _graph(data_cpu, args, width=320, height=240):
    from oracle.dagr_net import events_to_graph
    return events_to_graph(data_cpu.pos, data_cpu.batch, data_cpu.num_graphs, args.radius, args.max_neighbors, width,
                           height, 1000000)


def test_cartesian_bit_exact():
    from eventad_b200 import ops
    from oracle import dagr_ops
    model, args = make_backbone(use_image=False)
    d = make_data(20_000, 2, seed=1)
    ei = _graph(d, args)
    for m, clamp in ((0.03125, True), (0.0625, False), (torch.tensor(0.1) * 2, False)):
        ref = dagr_ops.cartesian(d.pos, ei, m)
        if clamp:
            ref = torch.clamp(ref, 0, 1)
        from eventad_b200.model.layers.components import two_m_of
        got = ops.cartesian(d.pos.cuda(), ei.cuda(), two_m_of(m), clamp)
        assert torch.equal(got.cpu(), ref)


@pytest.mark.parametrize("B", [1, 3])
def test_sample_features(B):
    from eventad_b200 import ops, synthetic
    from oracle import dagr_layers
    d = make_data(5000, B, seed=2)
    feats = synthetic.make_image_features(B, seed=3)
    for f in (feats[0], feats[1], feats[4]):
        ref = dagr_layers.sample_features(d.pos, d.batch, f, 320, 240)
        got = ops.sample_concat(None, d.pos.cuda(), d.batch.int().cuda(), ops.to_channel_last(f.cuda()), 320, 240)
        assert_close(got, ref, rtol=1e-5, atol=1e-6, what="sample_features")


def _conv_case(cin, cout, mode, n=3000, seed=0, pad4=False):
    """One MySplineConv on a real event graph; mode in {'lut', 'basis'}.  pad4: the GPU input rows are zero-padded to a
    multiple of four floats, the row shape the fused forward hands to the pooled layers."""
    from types import SimpleNamespace
    from eventad_b200.data import Data
    from eventad_b200.model.layers.spline_conv import MySplineConv
    from oracle import dagr_layers, dagr_ops
    model, args = make_backbone(use_image=False)
    d = make_data(n, 1, seed=seed, dist="C")
    ei = _graph(d, args)
    torch.manual_seed(seed)
    conv = MySplineConv(cin, cout, args, bias=(seed % 2 == 1))
    if conv.bias is not None:
        conv.bias.data.uniform_(-0.5, 0.5)
    x = torch.randn(n, cin)
    attr = torch.clamp(dagr_ops.cartesian(d.pos, ei, 0.03125), 0, 1)
    lut = dagr_layers.lut_params_layer1(args.radius, 320) if mode == "lut" else None
    sd = {"c." + k: v.detach() for k, v in conv.state_dict().items()}
    ref = dagr_layers.spline_conv(x, ei, attr, sd, "c.", lut, 240, 320)
    conv = conv.cuda()
    if mode == "lut":
        conv.init_lut(height=240, width=320, **{k: lut[k] for k in ("rx", "Mx")})
    xg = x.cuda()
    if pad4 and cin % 4:
        xg = torch.nn.functional.pad(xg, (0, 4 - cin % 4))
    data = Data(x=xg, edge_index=ei.cuda(), edge_attr=attr.cuda(), pos=d.pos.cuda())
    out = conv(data).x
    assert_close(out, ref, what=f"MySplineConv {cin}->{cout} {mode}")


@pytest.mark.parametrize("cin,cout,mode", [(19, 16, "lut"), (16, 16, "lut"), (19, 16, "basis"), (82, 64, "lut"),
                                           (64, 64, "basis"), (130, 64, "lut"), (5, 16, "lut"), (7, 3, "basis")])
def test_spline_conv(cin, cout, mode):
    _conv_case(cin, cout, mode, seed=cin % 2)


def test_spline_conv_shuffled_edges_and_empty():
    """edge_index in arbitrary order (the CSR builder must keep per-destination order) and a graph without edges."""
    from eventad_b200.data import Data
    from eventad_b200.model.layers.spline_conv import MySplineConv
    from oracle import dagr_layers, dagr_ops
    model, args = make_backbone(use_image=False)
    d = make_data(2000, 1, seed=4)
    ei = _graph(d, args)
    perm = torch.randperm(ei.shape[1], generator=torch.Generator().manual_seed(0))
    ei = ei[:, perm]
    attr = dagr_ops.cartesian(d.pos, ei, 0.03125)
    torch.manual_seed(0)
    conv = MySplineConv(12, 64, args)
    x = torch.randn(2000, 12)
    sd = {"c." + k: v.detach() for k, v in conv.state_dict().items()}
    ref = dagr_layers.spline_conv(x, ei, attr, sd, "c.", None)
    out = conv.cuda()(Data(x=x.cuda(), edge_index=ei.cuda(), edge_attr=attr.cuda())).x
    assert_close(out, ref, what="shuffled edges")
    empty = torch.zeros((2, 0), dtype=torch.long)
    ref0 = dagr_layers.spline_conv(x, empty, torch.zeros((0, 3)), sd, "c.", None)
    out0 = conv(Data(x=x.cuda(), edge_index=empty.cuda(), edge_attr=torch.zeros((0, 3)).cuda())).x
    assert_close(out0, ref0, what="no edges")


@pytest.mark.parametrize("which", ["conv_block1", "layer2"])
def test_layer_fused(which):
    from eventad_b200.data import Data
    from oracle import dagr_layers, dagr_ops
    model, args = make_backbone(use_image=False, seed=1)
    model.cache_luts(320, 240, args.radius)
    sd = net_state_dict(model)
    d = make_data(4000, 1, seed=5, dist="C")
    ei = _graph(d, args)
    layer = getattr(model.backbone, which)
    cin = layer.in_channel
    x = torch.randn(4000, cin, generator=torch.Generator().manual_seed(1))
    if which == "conv_block1":
        attr = torch.clamp(dagr_ops.cartesian(d.pos, ei, 0.03125), 0, 1)
        lut = dagr_layers.lut_params_layer1(args.radius, 320)
    else:   # layer-2 LUT geometry on the fine graph: offsets stay inside the 25x25 table
        attr = dagr_ops.cartesian(d.pos, ei, 0.0625)
        p = model.backbone.pool1
        lut = dagr_layers.lut_params_from_pool(p.voxel_size.cpu(), p.transform.max, 240, 320)
    ref = dagr_layers.layer(x, ei, attr, sd, which + ".", lut, 240, 320)
    layer = layer.cuda()
    out = layer(Data(x=x.cuda(), edge_index=ei.cuda(), edge_attr=attr.cuda(), pos=d.pos.cuda())).x
    assert_close(out, ref, what=which)


@pytest.mark.parametrize("B,n,level,aggr", [(1, 20_000, 0, "max"), (3, 9_000, 0, "max"), (1, 50_000, 1, "max"),
                                            (2, 6_000, 3, "mean"), (1, 300, 2, "max"),
                                            # <= 4608 nodes and <= 8192 voxel keys: the single-CTA front end of pool.cu
                                            (1, 4_608, 0, "max"), (2, 2_000, 3, "mean"), (3, 1_500, 1, "max"), (1, 1, 0, "max")])
def test_pooling(B, n, level, aggr):
    """Pooling.forward vs the oracle: cluster ids, coarse edge_index, batch bit-exact; pos bit-exact; x max exact."""
    from eventad_b200.data import Batch
    from oracle import dagr_layers
    from oracle.dagr_net import compute_pooling_at_each_layer
    model, args = make_backbone(use_image=False)
    d = make_data(n, B, seed=6, dist="C")          # the last event of every sample has t == 1.0 (cluster quirk)
    ei = _graph(d, args)
    pool = [model.backbone.pool1, model.backbone.pool2, model.backbone.pool3, model.backbone.pool4][level]
    assert pool.aggr == aggr
    n = d.pos.shape[0]                              # make_data takes events PER window
    x = torch.randn(n, 24, generator=torch.Generator().manual_seed(2))
    voxel3 = compute_pooling_at_each_layer(args.pooling_dim_at_output, 4)[level]
    ref = dagr_layers.pooling(x, d.pos, d.batch, ei, voxel3, 320, 240, args.batch_size, pool.transform.max, aggr)
    data = Batch(x=x.cuda(), pos=d.pos.cuda(), batch=d.batch.cuda(), edge_index=ei.cuda())
    data.width, data.height = d.width, d.height
    out = pool.cuda()(data)
    assert torch.equal(out.cluster.cpu(), ref.cluster)
    assert torch.equal(out.edge_index.cpu(), ref.edge_index)
    assert torch.equal(out.batch.cpu(), ref.batch)
    assert torch.equal(out.pos.cpu(), ref.pos)
    if aggr == "max":
        assert torch.equal(out.x.cpu(), ref.x)
    else:
        assert_close(out.x, ref.x, what="mean pool")
    assert torch.equal(out.edge_attr.cpu(), ref.edge_attr)


@pytest.mark.parametrize("B,n,deg", [(3, 16_000, 16), (1, 30_000, 3)])
def test_pooling_long_range_edges(B, n, deg):
    """Coarse edges from a fine graph whose edges are random node pairs: the sources of a cluster lie all over the id
    range (outside the 4096-id window of the one-pass bitmap) and rows hold more than 64 distinct sources, i.e. the
    general path of the edge kernels (counting by warp, filling by CTA) instead of the parked-row fast path."""
    from eventad_b200.data import Batch
    from oracle import dagr_layers
    from oracle.dagr_net import compute_pooling_at_each_layer
    model, args = make_backbone(use_image=False)
    d = make_data(n, B, seed=9, dist="U")
    N = d.pos.shape[0]
    g = torch.Generator().manual_seed(4)
    dst = torch.arange(N).repeat_interleave(deg)
    src = torch.randint(0, N, (N * deg,), generator=g)
    same = d.batch[src] == d.batch[dst]                  # edges stay inside a sample, like the reference's graph
    ei = torch.stack([src[same], dst[same]]).long()
    pool = model.backbone.pool1
    x = torch.randn(N, 8, generator=g)
    voxel3 = compute_pooling_at_each_layer(args.pooling_dim_at_output, 4)[0]
    ref = dagr_layers.pooling(x, d.pos, d.batch, ei, voxel3, 320, 240, args.batch_size, pool.transform.max, "max")
    data = Batch(x=x.cuda(), pos=d.pos.cuda(), batch=d.batch.cuda(), edge_index=ei.cuda())
    data.width, data.height = d.width, d.height
    out = pool.cuda()(data)
    assert torch.equal(out.cluster.cpu(), ref.cluster)
    assert torch.equal(out.edge_index.cpu(), ref.edge_index)
    deg_c = torch.bincount(ref.edge_index[1])
    assert int(deg_c.max()) > 64 or deg < 8              # rows beyond the parked-row capacity were exercised
    assert torch.equal(out.x.cpu(), ref.x) and torch.equal(out.pos.cpu(), ref.pos)


@pytest.mark.parametrize("level,self_loop,keep_t", [(0, True, False), (0, False, True), (1, True, True), (2, True, False)])
def test_pooling_options(level, self_loop, keep_t):
    """Pooling(self_loop=..., keep_temporal_ordering=...) (pooling.py:59-63,69-72) vs the oracle, bit-exact edges.  The
    level-1 / level-2 inputs are coarse graphs without self loops, where (c, c) appears only if two members are linked."""
    from eventad_b200.data import Batch
    from eventad_b200.model.layers.components import Cartesian
    from eventad_b200.model.layers.pooling import Pooling
    from oracle import dagr_layers
    from oracle.dagr_net import compute_pooling_at_each_layer
    model, args = make_backbone(use_image=False)
    B, n = 2, 12_000
    d = make_data(n, B, seed=8, dist="C")
    ei = _graph(d, args)
    voxels = compute_pooling_at_each_layer(args.pooling_dim_at_output, 4)
    x = torch.randn(d.pos.shape[0], 8, generator=torch.Generator().manual_seed(3))
    pos, batch = d.pos, d.batch
    for lv in range(level):                          # walk down to the requested level with the oracle (default options)
        tm = [model.backbone.pool1, model.backbone.pool2, model.backbone.pool3][lv].transform.max
        r = dagr_layers.pooling(x, pos, batch, ei, voxels[lv], 320, 240, args.batch_size, tm, "max")
        x, pos, batch, ei = r.x, r.pos, r.batch, r.edge_index
    tmax = [model.backbone.pool1, model.backbone.pool2, model.backbone.pool3][level].transform.max
    ref = dagr_layers.pooling(x, pos, batch, ei, voxels[level], 320, 240, args.batch_size, tmax, "max",
                              keep_temporal_ordering=keep_t, self_loop=self_loop)
    pool = Pooling(voxels[level], width=320, height=240, batch_size=args.batch_size,
                   transform=Cartesian(norm=True, cat=False, max_value=tmax), aggr="max",
                   keep_temporal_ordering=keep_t, self_loop=self_loop).cuda()
    data = Batch(x=x.cuda(), pos=pos.cuda(), batch=batch.cuda(), edge_index=ei.cuda())
    data.width, data.height = d.width, d.height
    out = pool(data)
    assert torch.equal(out.cluster.cpu(), ref.cluster)
    assert torch.equal(out.edge_index.cpu(), ref.edge_index)
    assert torch.equal(out.x.cpu(), ref.x) and torch.equal(out.pos.cpu(), ref.pos)
    assert torch.equal(out.edge_attr.cpu(), ref.edge_attr)
    if self_loop and not keep_t:                     # (c, c) has t_dst == t_src and does not survive the temporal filter
        assert bool((ref.edge_index[0] == ref.edge_index[1]).any())


@pytest.mark.parametrize("dist,B", [("U", 1), ("C", 3)])
def test_spiral_offset_class_equals_lut_class(dist, B):
    """The class bytes evad_graph_build derives from the spiral offset equal the reference's fp32 LUT index
    (spline_conv.py:39-42) computed per edge by evad_edge_class, and the exhaustive host check agrees."""
    from eventad_b200 import ops
    from eventad_b200.graph.ev_graph import SlidingWindowGraph
    from eventad_b200.model.layers.components import two_m_of
    model, args = make_backbone(use_image=True)
    model.cache_luts(320, 240, args.radius)
    conv = model.backbone.conv_block1.conv_block1.conv.cuda()
    d = make_data(30_000, B, seed=11, dist=dist, device="cuda")
    pos_int = ops.denormalize_pos(d.pos, 320, 240, 1_000_000)
    g = SlidingWindowGraph(width=320, height=240, batch_size=B, max_num_neighbors=16, max_queue_size=128, radius=4,
                           delta_t_us=10_000)
    out = g.build_csr(d.batch.int(), pos_int, want_cls=True)
    e = int(out["num_edges"].item())
    two_m = two_m_of(model.backbone.edge_attrs.max)
    assert ops.offset_class_is_lut_class(conv.mode(), two_m, True, 320, 240)
    status = ops.new_status(d.pos.device)
    ref = ops.edge_class(out["rowptr"], out["src"], d.pos.shape[0], conv.mode(),
                         dict(pos=d.pos.contiguous(), two_m=two_m, clamp01=True), status)
    assert int(status.item()) == 0
    assert torch.equal(out["edge_cls"][:e], ref[:e])


def test_spline_conv_to_dense():
    """SplineConvToDense.forward (spline_conv.py:80-118): the convolution, then every node written into the dense
    [B, C, H, W] map at (pos / pooling).long(); against the oracle convolution scattered independently (numpy loop)."""
    import numpy as np
    from eventad_b200.data import Data
    from eventad_b200.model.layers.spline_conv import SplineConvToDense
    from oracle import dagr_layers, dagr_ops
    model, args = make_backbone(use_image=False)
    B, Wc, Hc = 2, 10, 8                                        # a pooled graph: one node per occupied cell of a 10 x 8 grid
    g = torch.Generator().manual_seed(5)
    cells = torch.stack([torch.randperm(Wc * Hc, generator=g)[:40] for _ in range(B)])
    batch = torch.arange(B).repeat_interleave(40)
    cx, cy = (cells % Wc).flatten(), (cells // Wc).flatten()
    pooling = torch.tensor([1.0 / Wc, 1.0 / Hc, 1.0])
    pos = torch.stack([(cx + 0.5) / Wc, (cy + 0.5) / Hc, torch.rand(B * 40, generator=g)], 1)
    n = pos.shape[0]
    src, dst = torch.randint(0, 40, (2, 300), generator=g)
    off = (torch.arange(300) % B) * 40
    ei = torch.stack([src + off, dst + off])
    attr = dagr_ops.cartesian(pos, ei, 1.0).clamp(0, 1)
    torch.manual_seed(1)
    conv = SplineConvToDense(12, 8, args)
    x = torch.randn(n, 12, generator=g)
    sd = {"c." + k: v.detach() for k, v in conv.state_dict().items()}
    ref = dagr_layers.spline_conv(x, ei, attr, sd, "c.", None).numpy()
    want = np.zeros((B, 8, Hc, Wc), dtype=np.float32)
    for i in range(n):
        want[int(batch[i]), :, int(cy[i]), int(cx[i])] = ref[i]
    d = Data(x=x.cuda(), edge_index=ei.cuda(), edge_attr=attr.cuda(), pos=pos.cuda(), batch=batch.cuda())
    d.pooling = pooling.cuda()
    dense = conv.cuda()(d)
    assert tuple(dense.shape) == (B, 8, Hc, Wc)
    assert_close(dense.cpu(), torch.from_numpy(want), what="SplineConvToDense")
    # second call reuses (and clears) the cached map, as the reference does
    d2 = Data(x=x[:40].cuda(), edge_index=ei[:, ei[0] < 40][:, :50].cuda(), edge_attr=attr[ei[0] < 40][:50].cuda(),
              pos=pos[:40].cuda(), batch=batch[:40].cuda())
    d2.pooling = pooling.cuda()
    dense2 = conv(d2)
    assert float(dense2[1].abs().max()) == 0.0 and float(dense2[0].abs().max()) > 0.0


@pytest.mark.parametrize("cin,mode", [(82, "lut"), (64, "basis"), (130, "lut"), (18, "lut")])
def test_spline_conv_large_rows_h2_gemm(cin, mode):
    """>= 8 192 rows: the aggregated operand is written as fp16 pairs in the UMMA operand image (evad_spline_aggregate_h2)
    and contracted by the bulk-copy / tcgen05 pipeline (evad_gemm64_h2); K = 26 * cin with and without a K tail, and
    against the fp32-operand tcgen05 kernel (EVAD_GEMM_H2=0) on the same inputs."""
    from eventad_b200 import ops
    assert ops.USE_GEMM_H2
    calls = []
    real = ops.gemm64_h2
    ops.gemm64_h2 = lambda *a, **k: (calls.append(1), real(*a, **k))[1]
    try:
        _conv_case(cin, 64, mode, n=20_000, seed=cin % 2, pad4=True)
        assert calls == [1], "the h2 path did not run"
        ops.USE_GEMM_H2 = False
        _conv_case(cin, 64, mode, n=20_000, seed=cin % 2, pad4=True)
        assert calls == [1]
    finally:
        ops.USE_GEMM_H2, ops.gemm64_h2 = True, real
