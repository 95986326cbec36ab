"""Streaming (append) mode of the graph builder against the C oracle: micro-batches of 1..10^3 events on top of resident
events, single-event calls (the reference's b=0 insertion quirk), sliding-window deletion, Q-cap dependence on the cuts."""
import numpy as np
import pytest
import torch

from eventad_b200 import synthetic

pytestmark = pytest.mark.gpu


def _mk(n, B, W, H, dist="C", seed=0):
    raw = synthetic.make_batch(n, B, seed=seed, dist=dist, width=W, height=H)
    pos = np.concatenate([raw["pos"].astype(np.int32), raw["t"].reshape(-1, 1)], 1).astype(np.int32)
    return raw["batch"].astype(np.int32), pos


@pytest.mark.parametrize("Q,cuts", [(16, [0, 4000, 4001, 4002, 4100, 5100, 6000]), (4, [0, 1, 2, 500, 6000]),
                                    (128, [0, 3000, 6000])])
def test_append_matches_oracle(Q, cuts):
    from eventad_b200.graph.ev_graph import SlidingWindowGraph
    from oracle.evg import OracleSlidingWindowGraph
    W, H = 64, 48
    batch, pos = _mk(6000, 1, W, H)
    kw = dict(width=W, height=H, batch_size=1, max_num_neighbors=16, max_queue_size=Q, radius=4, delta_t_us=10000)
    g, o = SlidingWindowGraph(**kw), OracleSlidingWindowGraph(**kw)
    for a, b in zip(cuts[:-1], cuts[1:]):
        e = g.forward(torch.from_numpy(batch[a:b]).cuda(), torch.from_numpy(pos[a:b]).cuda(), delete_nodes=False)
        eo = o.forward(batch[a:b], pos[a:b], delete_nodes=False)
        assert np.array_equal(e.cpu().numpy(), eo), (a, b)
    assert g.max_index == o.max_index == 6000
    assert np.array_equal(g.edges.cpu().numpy(), o.edges)
    assert np.array_equal(g.all_timestamps.cpu().numpy(), o.all_timestamps)


def test_sliding_window_delete_nodes():
    from eventad_b200.graph.ev_graph import SlidingWindowGraph
    from oracle.evg import OracleSlidingWindowGraph
    W, H = 40, 30
    batch, pos = _mk(5000, 1, W, H, seed=3)
    kw = dict(width=W, height=H, batch_size=1, max_num_neighbors=8, max_queue_size=8, radius=3, delta_t_us=20000)
    g, o = SlidingWindowGraph(**kw), OracleSlidingWindowGraph(**kw)
    for a, b in [(0, 2000), (2000, 3000), (3000, 3001), (3001, 5000)]:
        r = g.forward(torch.from_numpy(batch[a:b]).cuda(), torch.from_numpy(pos[a:b]).cuda(), delete_nodes=True,
                      return_total_edges=True, return_node_counts=True)
        ro = o.forward(batch[a:b], pos[a:b], delete_nodes=True, return_total_edges=True, return_node_counts=True)
        assert np.array_equal(r[0].cpu().numpy(), ro[0]), (a, b)
        assert np.array_equal(r[1].cpu().numpy(), ro[1]), (a, b)           # deleted edges
        assert np.array_equal(r[2].cpu().numpy(), ro[2]), (a, b)           # total edges before deletion
        assert r[3] == ro[3]
        assert g.min_index == o.min_index
    assert np.array_equal(g.edges.cpu().numpy(), o.edges)


def test_multi_sample_append_and_single_event_quirk():
    from eventad_b200.graph.ev_graph import SlidingWindowGraph
    from oracle.evg import OracleSlidingWindowGraph
    W, H = 32, 24
    batch, pos = _mk(1500, 2, W, H, seed=5)                      # 3000 events, sample-major
    order = np.argsort(pos[:, 2], kind="stable")                 # interleave the two samples in time
    batch, pos = batch[order], pos[order]
    kw = dict(width=W, height=H, batch_size=2, max_num_neighbors=16, max_queue_size=16, radius=4, delta_t_us=10000)
    g, o = SlidingWindowGraph(**kw), OracleSlidingWindowGraph(**kw)
    cuts = [0, 1000, 2000, 2001, 2002, 2003, 3000]               # single events: inserted into sample 0 whatever batch says
    for a, b in zip(cuts[:-1], cuts[1:]):
        e = g.forward(torch.from_numpy(batch[a:b]).cuda(), torch.from_numpy(pos[a:b]).cuda(), delete_nodes=False)
        eo = o.forward(batch[a:b], pos[a:b], delete_nodes=False)
        assert np.array_equal(e.cpu().numpy(), eo), (a, b)


def test_ev_tgn_reset_false_after_fused_forward():
    """EV_TGN.forward(reset=False) after a fused Net.forward: the graph is appended (asynchronous mode entry)."""
    from oracle.evg import OracleSlidingWindowGraph
    from tests.util import make_backbone, make_data
    model, args = make_backbone(use_image=False)
    net = model.cuda().backbone
    d = make_data(5000, 1, seed=2, dist="C", device="cuda")
    with torch.no_grad():
        net(d)
    d2 = make_data(300, 1, seed=9, dist="C", device="cuda")
    out = net.events_to_graph(d2, reset=False)
    o = OracleSlidingWindowGraph(width=320, height=240, batch_size=1, max_num_neighbors=16, max_queue_size=128, radius=4,
                                 delta_t_us=10000)
    for dd in (d, d2):
        p = (torch.tensor([320, 240, 1000000]) * dd.pos.cpu() + 1e-3).int().numpy()
        eo = o.forward(dd.batch.cpu().int().numpy(), p, delete_nodes=False)
    assert np.array_equal(out.edge_index.cpu().numpy(), eo)


@pytest.mark.parametrize("dist", ["U", "C"])
def test_streaming_at_baseline_config5_size(dist):
    """BASELINE configs[4]: 320x240, 10^6 resident events (clustered: > Q = 128 events on single pixels, the FIFO cap is
    live), then micro-batches of 1, 10, 10^2, 10^3, 10^4 events with delete_nodes=True (sliding window), bit-exact against
    the C oracle AND against the unmodified reference kernels (oracle/_ref/ev_graph_cuda.so) on every micro-batch
    (reference: ev_graph.py:75-103,121-136, ev_graph.cu:15-80,130-212)."""
    from eventad_b200.graph.ev_graph import SlidingWindowGraph
    from oracle.evg import OracleSlidingWindowGraph
    from tests.util import RefGraph, load_ref_graph_module
    W, H, N0 = 320, 240, 1_000_000
    sizes = [1, 10, 100, 1000, 10_000, 1, 10, 100, 1000]
    total = N0 + sum(sizes)
    batch, pos = _mk(total, 1, W, H, dist=dist, seed=21)          # one time-sorted stream of `total` events
    kw = dict(width=W, height=H, batch_size=1, max_num_neighbors=16, max_queue_size=128, radius=4, delta_t_us=10000)
    g, o = SlidingWindowGraph(**kw), OracleSlidingWindowGraph(**kw)
    m = load_ref_graph_module()
    ref = RefGraph(m, W, H, 1, 16, 128, 4, 10000) if m is not None else None
    cuts = np.cumsum([0, N0] + sizes)
    if dist == "C":
        px = pos[:N0, 0].astype(np.int64) + W * pos[:N0, 1].astype(np.int64)
        assert np.bincount(px).max() > 128
    for i, (a, b) in enumerate(zip(cuts[:-1], cuts[1:])):
        bt, pt = torch.from_numpy(batch[a:b]).cuda(), torch.from_numpy(pos[a:b]).cuda()
        n_delete = (b - a) if i > 0 else 0
        r = g.forward(bt, pt, delete_nodes=True, return_node_counts=True)
        ro = o.forward(batch[a:b], pos[a:b], delete_nodes=True, return_node_counts=True)
        assert np.array_equal(r[0].cpu().numpy(), ro[0]), (dist, a, b)
        if i > 0:
            assert np.array_equal(r[1].cpu().numpy(), ro[1]), (dist, a, b)       # edges that left the window
        assert r[2] == ro[2] and g.min_index == o.min_index and g.max_index == o.max_index
        if ref is not None:
            er = ref.forward(bt, pt)
            assert torch.equal(er, r[0]), ("reference kernel", dist, a, b)
            ref.delete_nodes(n_delete)
    assert np.array_equal(g.edges.cpu().numpy(), o.edges)
    assert m is not None, "oracle/_ref/ev_graph_cuda.so missing: the reference-kernel half of this test did not run"
