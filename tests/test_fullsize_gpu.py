"""BASELINE.json's full sizes: 1M-event windows (configs[4], north_star "50k-1M events per window") checked bit-exactly
against the C oracle, the whole branch at 1M events against the torch-CPU oracle, and the bench batch (64 windows x
100k events = 6.4M events, configs[2]) through size-independent properties of the reference's graph
(ev_graph.cu:15-80: self loop first, causal sources, radius and time gates, neighbour cap, no duplicates) and of the
pooled outputs (sorted batch vector, bounded cluster counts, finite features, a per-window result that does not depend
on which other windows share the batch)."""
import numpy as np
import pytest
import torch

from eventad_b200 import synthetic
from tests.util import assert_close, make_backbone, make_data, net_state_dict

pytestmark = pytest.mark.gpu

W, H, T, R, DT, K, Q = 320, 240, 1_000_000, 4, 10_000, 16, 128


def _pos_int(raw):
    return np.concatenate([raw["pos"].astype(np.int32), raw["t"].reshape(-1, 1)], 1).astype(np.int32)


@pytest.mark.parametrize("dist", ["U", "C"])
def test_graph_1m_events_bit_exact(dist):
    """One 1M-event window; the clustered one puts > Q = 128 events on single pixels, so the FIFO cap is live."""
    from eventad_b200.graph.ev_graph import SlidingWindowGraph
    from oracle.evg import OracleSlidingWindowGraph
    raw = synthetic.make_batch(1_000_000, 1, seed=4, dist=dist)
    pos, batch = _pos_int(raw), raw["batch"].astype(np.int32)
    if dist == "C":
        per_pixel = np.bincount(pos[:, 0] + W * pos[:, 1], minlength=W * H)
        assert per_pixel.max() > Q
    g = SlidingWindowGraph(width=W, height=H, batch_size=1, max_num_neighbors=K, max_queue_size=Q, radius=R, delta_t_us=DT)
    e = g.forward(torch.from_numpy(batch).cuda(), torch.from_numpy(pos).cuda(), delete_nodes=False)
    o = OracleSlidingWindowGraph(width=W, height=H, batch_size=1, max_num_neighbors=K, max_queue_size=Q, radius=R,
                                 delta_t_us=DT)
    eo = o.forward(batch, pos, delete_nodes=False)
    assert e.shape[1] == eo.shape[1]
    assert np.array_equal(e.cpu().numpy(), eo)


def test_net_forward_1m_events_matches_oracle():
    from oracle.dagr_net import net_forward
    model, args = make_backbone(seed=2, use_image=True)
    model.cache_luts(W, H, args.radius)
    sd = net_state_dict(model)
    d = make_data(1_000_000, 1, seed=5, dist="C")
    feats = synthetic.make_image_features(1, seed=2)
    ref = net_forward(sd, args, d.x, d.pos, d.batch, 1, feats, W, H)
    net = model.cuda().backbone
    with torch.no_grad():
        outs = net(make_data(1_000_000, 1, seed=5, dist="C", device="cuda"), image_feat=[f.cuda() for f in feats])[0]
    for o, r, name in zip(outs, ref, ("out3", "out4")):
        assert torch.equal(o.batch.cpu(), r.batch) and torch.equal(o.edge_index.cpu(), r.edge_index), name
        assert torch.equal(o.pos[:, :2].cpu(), r.pos[:, :2]), name
        assert_close(o.x, r.x, what=f"1M events {name} x")


def test_bench_batch_graph_properties():
    """64 windows x 100k events: the properties every reference graph has, evaluated on the GPU over all 6.4M rows."""
    from eventad_b200.graph.ev_graph import SlidingWindowGraph
    B, n = 64, 100_000
    raw = synthetic.make_batch(n, B, seed=9, dist="U")
    pos = torch.from_numpy(_pos_int(raw)).cuda()
    batch = torch.from_numpy(raw["batch"].astype(np.int32)).cuda()
    g = SlidingWindowGraph(width=W, height=H, batch_size=B, max_num_neighbors=K, max_queue_size=Q, radius=R, delta_t_us=DT)
    rows = g.build_rows(batch, pos, want_cls=True)
    nbr, cnt, cls = rows["nbr"].long(), rows["count"].long(), rows["nbr_cls"].long()
    N = B * n
    assert int(cnt.min()) >= 1 and int(cnt.max()) <= K
    assert torch.equal(nbr[:, 0], torch.arange(N, device="cuda"))            # self loop first (ev_graph.cu:43-46)
    live = torch.arange(K, device="cuda")[None, :] < cnt[:, None]
    dst = torch.arange(N, device="cuda")[:, None].expand(N, K)[live]
    src = nbr[live]
    assert bool((src <= dst).all())                                           # causal: only older (or own) indices
    assert bool((batch[src] == batch[dst]).all())                             # never across windows
    dx, dy = pos[src, 0] - pos[dst, 0], pos[src, 1] - pos[dst, 1]
    assert int(dx.abs().max()) <= R and int(dy.abs().max()) <= R
    assert bool(((pos[dst, 2] - pos[src, 2]).float() <= float(DT)).all())     # time gate (ev_graph.cu:69)
    assert torch.equal(cls[live], (dx + R) * (2 * R + 1) + (dy + R))          # offset class bytes of the spiral walk
    key = dst * N + src                                                      # no duplicate edges
    assert int(torch.unique(key).numel()) == int(key.numel())
    # a row that stopped short of K had no further candidate: its count equals the brute-force count for a sample
    short = (cnt < K).nonzero().view(-1)[:200]
    for i in short.tolist():
        b0 = int(batch[i])
        m = (batch == b0) & (torch.arange(N, device="cuda") < i) & ((pos[:, 0] - pos[i, 0]).abs() <= R) & \
            ((pos[:, 1] - pos[i, 1]).abs() <= R) & ((pos[i, 2] - pos[:, 2]).float() <= float(DT))
        assert int(m.sum()) + 1 == int(cnt[i]), i


def test_bench_batch_window_independence():
    """A window's out3 / out4 rows do not depend on the other windows of the batch (windows shard freely), up to
    the t == T quirk that moves a window's last events into the next sample's voxel (A.7-2): compare windows whose
    successor is the same in both batches."""
    model, args = make_backbone(seed=1, use_image=True)
    model.cache_luts(W, H, args.radius)
    net = model.cuda().backbone
    n = 100_000
    big, small = 8, 3
    feats = [f.cuda() for f in synthetic.make_image_features(big, seed=3)]
    with torch.no_grad():
        ob = net(make_data(n, big, seed=6, dist="U", device="cuda"), image_feat=feats)[0]
        osm = net(make_data(n, small, seed=6, dist="U", device="cuda"), image_feat=[f[:small] for f in feats])[0]
    for a, b in zip(ob, osm):
        assert bool((a.batch[1:] >= a.batch[:-1]).all())
        assert bool(torch.isfinite(a.x).all())
        for w in range(small - 1):                           # windows 0, 1: same successor in both batches
            ma, mb = a.batch == w, b.batch == w
            assert int(ma.sum()) == int(mb.sum())
            assert torch.equal(a.pos[ma], b.pos[mb])
            assert_close(a.x[ma], b.x[mb], what=f"window {w}")


@pytest.mark.slow
@pytest.mark.parametrize("dist", ["U", "C"])
def test_bench_batch_matches_oracle(dist):
    """The batch `bench.py` times (64 windows x 100k events, `synthetic.make_batch(100_000, 64, seed=1000*rank+s)`, rank 0,
    s 0) through the whole fused branch against the CPU oracle on the same batch: out3 / out4 structure bit-exact,
    features rtol 1e-4 / atol 1e-5.  The oracle walks the per-sample front end (graph, layer 1) one window at a time
    (tests/test_oracle_cpu.py::test_chunked_front_equals_unchunked), ~2 s of CPU per window."""
    from oracle.dagr_net import net_forward
    B, n = 64, 100_000
    model, args = make_backbone(seed=0, use_image=True)
    model.cache_luts(W, H, args.radius)
    sd = net_state_dict(model)
    d = make_data(n, B, seed=0, dist=dist)
    feats = synthetic.make_image_features(B, seed=0)
    ref = net_forward(sd, args, d.x, d.pos, d.batch, B, feats, W, H, chunk_front_by_sample=True)
    net = model.cuda().backbone
    with torch.no_grad():
        outs = net(make_data(n, B, seed=0, dist=dist, device="cuda"), image_feat=[f.cuda() for f in feats])[0]
    for o, r, name in zip(outs, ref, ("out3", "out4")):
        assert torch.equal(o.batch.cpu(), r.batch), name
        assert torch.equal(o.edge_index.cpu(), r.edge_index), name
        assert torch.equal(o.pos.cpu(), r.pos), name
        assert_close(o.x, r.x, what=f"bench batch {dist} {name} x")
