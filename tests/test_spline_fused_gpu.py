"""Fused tap aggregation + tcgen05 contraction (csrc/spline_tc5.cu) against the two-kernel path it replaces
(evad_spline_aggregate + fp64 matmul) on random pooled-graph-like inputs: LUT-free basis mode and ragged rows,
device-side row counts, the tap-split path for small graphs, determinism."""
import pytest
import torch

pytestmark = pytest.mark.gpu


def _random_graph(n, deg_max, seed):
    g = torch.Generator().manual_seed(seed)
    deg = torch.randint(0, deg_max + 1, (n,), generator=g)
    rowptr = torch.zeros(n + 1, dtype=torch.int32)
    rowptr[1:] = torch.cumsum(deg, 0).int()
    e = int(rowptr[-1])
    src = torch.randint(0, n, (max(e, 1),), generator=g).int()
    attr = torch.rand((max(e, 1), 2), generator=g)
    return rowptr, src, attr, e


@pytest.mark.parametrize("n,cin,deg_max", [(3000, 82, 12), (20_000, 64, 9), (1500, 130, 20), (129, 64, 5)])
def test_fused_conv_matches_aggregate_then_fp64_matmul(n, cin, deg_max):
    from eventad_b200 import ops
    rowptr, src, attr, e = _random_graph(n, deg_max, seed=n + cin)
    g = torch.Generator().manual_seed(7)
    ldx = (cin + 3) // 4 * 4
    x = torch.zeros(n, ldx)
    x[:, :cin] = torch.randn(n, cin, generator=g)
    weight = torch.randn(25, cin, 64, generator=g) / (4 * cin) ** 0.5
    lin = torch.randn(64, cin, generator=g) / cin ** 0.5
    scale, shift = torch.rand(64, generator=g) + 0.5, torch.randn(64, generator=g)
    addend = torch.randn(n, 64, generator=g)
    mode = ops.spline_mode_basis(5)
    dev = "cuda"
    pseudo = dict(edge_attr=attr.to(dev), eid=None)
    status = ops.new_status(torch.device(dev))
    rp, sr, xd = rowptr.to(dev), src.to(dev), x.to(dev)
    A = ops.spline_aggregate(rp, sr, n, None, xd, cin, mode, pseudo, status)
    wcat = torch.cat([weight.reshape(25 * cin, 64), lin.t()], 0)
    ref = A[:, :26 * cin].double().cpu() @ wcat.double()
    lists = ops.spline_tap_lists(rp, sr, n, None, sr.shape[0], mode, pseudo, status)
    wp = ops.pack_fused_weight(weight.to(dev), lin.to(dev), ldx)
    got = ops.spline_conv64_fused(lists, xd, n, None, wp, dict())
    err = (got.cpu().double() - ref).abs()
    assert float((err / (2e-6 + 2e-5 * ref.abs())).max()) < 1.0, float(err.max())
    epi = dict(scale=scale.to(dev), shift=shift.to(dev), addend=addend.to(dev), relu=True)
    got2 = ops.spline_conv64_fused(lists, xd, n, None, wp, epi)
    ref2 = (ref * scale.double() + shift.double() + addend.double()).clamp(min=0)
    err2 = (got2.cpu().double() - ref2).abs()
    assert float((err2 / (4e-6 + 2e-5 * ref2.abs())).max()) < 1.0, float(err2.max())
    assert torch.equal(got2, ops.spline_conv64_fused(lists, xd, n, None, wp, epi))      # fixed summation order
    assert int(status.item()) == 0


def test_fused_conv_device_row_count():
    from eventad_b200 import ops
    n_cap, n_live, cin = 2600, 2100, 64
    rowptr, src, attr, e = _random_graph(n_cap, 8, seed=3)
    g = torch.Generator().manual_seed(9)
    x = torch.randn(n_cap, cin, generator=g).cuda()
    weight = (torch.randn(25, cin, 64, generator=g) / (4 * cin) ** 0.5).cuda()
    lin = (torch.randn(64, cin, generator=g) / cin ** 0.5).cuda()
    mode = ops.spline_mode_basis(5)
    pseudo = dict(edge_attr=attr.cuda(), eid=None)
    status = ops.new_status(torch.device("cuda"))
    n_dev = torch.tensor([n_live], dtype=torch.int32, device="cuda")
    src = src.clamp(max=n_live - 1)
    lists = ops.spline_tap_lists(rowptr.cuda(), src.cuda(), n_cap, n_dev, src.shape[0], mode, pseudo, status)
    out = torch.full((n_cap, 64), -3.0, device="cuda")
    ops.spline_conv64_fused(lists, x, n_cap, n_dev, ops.pack_fused_weight(weight, lin, cin), dict(), out=out)
    A = ops.spline_aggregate(rowptr.cuda(), src.cuda(), n_cap, n_dev, x, cin, mode, pseudo, status)
    wcat = torch.cat([weight.reshape(25 * cin, 64), lin.t()], 0)
    ref = A[:n_live].double() @ wcat.double()
    assert float((out[:n_live].double() - ref).abs().max()) < 2e-5
    assert torch.all(out[n_live:] == -3.0)
