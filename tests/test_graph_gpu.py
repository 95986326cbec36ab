"""Bit-exact parity of the sm_100a event-graph builder with the C oracle (oracle/evg_oracle.c), through the
reference-shaped SlidingWindowGraph API (reference: src/dagr/graph/ev_graph.py, ev_graph.cu)."""
import numpy as np
import pytest
import torch

from eventad_b200 import synthetic

pytestmark = pytest.mark.gpu


def _pos_int(raw):
    return np.concatenate([raw["pos"].astype(np.int32), raw["t"].reshape(-1, 1)], 1).astype(np.int32)


def _run_both(batch, pos, W, H, B, K=16, Q=128, r=4, dt=10000):
    from eventad_b200.graph.ev_graph import SlidingWindowGraph
    from oracle.evg import OracleSlidingWindowGraph
    g = SlidingWindowGraph(width=W, height=H, batch_size=B, max_num_neighbors=K, max_queue_size=Q, radius=r,
                           delta_t_us=dt)
    e = g.forward(torch.from_numpy(batch).int().cuda(), torch.from_numpy(pos).int().cuda(), delete_nodes=False)
    o = OracleSlidingWindowGraph(width=W, height=H, batch_size=B, max_num_neighbors=K, max_queue_size=Q, radius=r,
                                 delta_t_us=dt)
    eo = o.forward(batch, pos, delete_nodes=False)
    return e, eo, g


@pytest.mark.parametrize("n,dist,B", [(1, "U", 1), (2, "U", 1), (500, "U", 1), (5000, "C", 1), (50_000, "U", 1),
                                      (50_000, "C", 1), (100_000, "U", 1), (30_000, "C", 3), (200_000, "C", 2)])
def test_edge_index_bit_exact(n, dist, B):
    raw = synthetic.make_batch(n, B, seed=n % 7, dist=dist)
    pos = _pos_int(raw)
    batch = raw["batch"].astype(np.int32)
    e, eo, g = _run_both(batch, pos, raw["width"], raw["height"], B)
    assert e.dtype == torch.int64 and e.shape[0] == 2
    assert e.shape[1] == eo.shape[1]
    assert np.array_equal(e.cpu().numpy(), eo)
    rowptr, src, ne = g.csr
    assert int(ne.item()) == eo.shape[1]
    assert np.array_equal(src.cpu().numpy().astype(np.int64), eo[0])
    deg = np.diff(rowptr.cpu().numpy())
    assert np.array_equal(np.repeat(np.arange(len(batch)), deg), eo[1])


def test_queue_cap_and_ties():
    """> Q events on one pixel (FIFO cap hides the oldest), equal timestamps, border pixels, tiny sensor."""
    rng = np.random.default_rng(3)
    W, H, n = 12, 9, 4000
    x = rng.integers(0, W, n); y = rng.integers(0, H, n)
    x[:1500] = 5; y[:1500] = 4                         # 1500 events on one pixel, Q = 16
    x[1500:1600] = 0; y[1500:1600] = 0
    t = np.sort(rng.integers(0, 3000, n)) // 7 * 7       # ties
    perm = rng.permutation(n)
    pos = np.stack([x[perm], y[perm], t], 1).astype(np.int32)
    batch = np.zeros(n, np.int32)
    for K, Q, r, dt in [(16, 16, 4, 500), (8, 4, 2, 100), (32, 128, 1, 10_000), (5, 3, 3, 0)]:
        e, eo, _ = _run_both(batch, pos, W, H, 1, K=K, Q=Q, r=r, dt=dt)
        assert np.array_equal(e.cpu().numpy(), eo), (K, Q, r, dt)


def test_unsorted_timestamps_and_negative_dt():
    rng = np.random.default_rng(5)
    W, H, n = 40, 30, 20_000
    pos = np.stack([rng.integers(0, W, n), rng.integers(0, H, n), rng.integers(0, 50_000, n)], 1).astype(np.int32)
    batch = np.sort(rng.integers(0, 2, n)).astype(np.int32)
    e, eo, _ = _run_both(batch, pos, W, H, 2, dt=7000)
    assert np.array_equal(e.cpu().numpy(), eo)


def test_empty_and_errors():
    from eventad_b200.graph.ev_graph import SlidingWindowGraph
    g = SlidingWindowGraph(width=32, height=32, batch_size=1, max_queue_size=8, radius=2, delta_t_us=100)
    e = g.forward(torch.zeros(0, dtype=torch.int32, device="cuda"), torch.zeros((0, 3), dtype=torch.int32, device="cuda"),
                  delete_nodes=False)
    assert e.shape == (2, 0)
    with pytest.raises(RuntimeError):
        g.forward(torch.zeros(4, dtype=torch.int32), torch.zeros((4, 3), dtype=torch.int32), delete_nodes=False)
    with pytest.raises(AssertionError):
        g.forward(torch.zeros(4, dtype=torch.int64, device="cuda"), torch.zeros((4, 3), dtype=torch.int32, device="cuda"),
                  delete_nodes=False)
    bad = torch.tensor([[40, 1, 5], [1, 1, 6]], dtype=torch.int32, device="cuda")      # x outside the sensor
    with pytest.raises(RuntimeError):
        g.forward(torch.zeros(2, dtype=torch.int32, device="cuda"), bad, delete_nodes=False)


@pytest.mark.parametrize("n,B,dist", [(30_000, 2, "C"), (100_000, 1, "U"), (1, 1, "U")])
def test_fixed_stride_rows_equal_the_compacted_graph(n, B, dist):
    """evad_graph_build_rows (no scan / compaction) holds exactly the rows of evad_graph_build, slot for slot, and the
    offset-class bytes of both agree."""
    from eventad_b200.graph.ev_graph import SlidingWindowGraph
    raw = synthetic.make_batch(n, B, seed=5, dist=dist)
    pos = np.concatenate([raw["pos"].astype(np.int32), raw["t"].reshape(-1, 1)], 1).astype(np.int32)
    bt, pt = torch.from_numpy(raw["batch"].astype(np.int32)).cuda(), torch.from_numpy(pos).cuda()
    mk = lambda: SlidingWindowGraph(width=raw["width"], height=raw["height"], batch_size=B, max_num_neighbors=16,
                                    max_queue_size=128, radius=4, delta_t_us=10000)
    g1, g2 = mk(), mk()
    g1.reset(); g2.reset()
    csr = g1.build_csr(bt, pt, want_cells=True, want_cls=True)
    rows = g2.build_rows(bt, pt, want_cls=True)
    N = bt.shape[0]
    rowptr = csr["rowptr"].cpu().numpy()
    cnt = rows["count"].cpu().numpy()[:N]
    assert np.array_equal(np.diff(rowptr[:N + 1]), cnt)
    E = int(csr["num_edges"].item())
    mask = np.arange(16)[None, :] < cnt[:, None]
    assert np.array_equal(rows["nbr"].cpu().numpy()[:N][mask], csr["src"].cpu().numpy()[:E])
    assert np.array_equal(rows["nbr_cls"].cpu().numpy()[:N][mask], csr["edge_cls"].cpu().numpy()[:E])
    assert torch.equal(rows["cell_perm"][:N], csr["cell_perm"][:N])


@pytest.mark.parametrize("W,H,n,K,Q,r,dt", [(12, 9, 4000, 16, 128, 4, 10_000), (12, 9, 4000, 16, 128, 4, 300),
                                             (20, 16, 9000, 32, 64, 5, 2000), (7, 5, 600, 2, 128, 3, 10_000),
                                             (16, 16, 6000, 9, 48, 2, 1500), (30, 20, 5000, 16, 128, 7, 3000),
                                             (30, 20, 5000, 12, 128, 6, 3000), (9, 9, 700, 16, 200, 1, 10_000),
                                             (40, 30, 3000, 16, 128, 4, 0)])
@pytest.mark.parametrize("span_limit", [32, 8, 0])
def test_row_span_search_saturated_lists(W, H, n, K, Q, r, dt, span_limit):
    """Tens of events per pixel: the K-1 list overflows on most events, so what is kept and in which order is the
    spiral-position / newest-first rule; offset classes are recomputed from the coordinates of both ends.  The search
    kernel picks, per event, the row-span search (all spans <= min(span_limit, Q, 32) records) or the per-cell walk:
    32 = default mix, 8 = more events walk, 0 = all walk."""
    from eventad_b200 import _lib
    old = _lib.lib().evad_graph_set_span_limit(span_limit)
    try:
        _row_span_case(W, H, n, K, Q, r, dt)
    finally:
        _lib.lib().evad_graph_set_span_limit(old)


def _row_span_case(W, H, n, K, Q, r, dt):
    from eventad_b200.graph.ev_graph import SlidingWindowGraph
    rng = np.random.default_rng(W * H + K)
    x, y = rng.integers(0, W, n), rng.integers(0, H, n)
    t = np.sort(rng.integers(0, 20_000, n)) // 3 * 3
    pos = np.stack([x, y, t], 1).astype(np.int32)
    batch = np.zeros(n, np.int32)
    e, eo, _ = _run_both(batch, pos, W, H, 1, K=K, Q=Q, r=r, dt=dt)
    assert np.array_equal(e.cpu().numpy(), eo)
    g = SlidingWindowGraph(width=W, height=H, batch_size=1, max_num_neighbors=K, max_queue_size=Q, radius=r, delta_t_us=dt)
    g.reset()
    rows = g.build_rows(torch.from_numpy(batch).cuda(), torch.from_numpy(pos).cuda(), want_cls=True)
    cnt = rows["count"].cpu().numpy()[:n]
    mask = np.arange(K)[None, :] < cnt[:, None]
    nbr = rows["nbr"].cpu().numpy()[:n].reshape(n, K)
    assert np.array_equal(nbr[mask].astype(np.int64), eo[0])
    dst = np.repeat(np.arange(n), cnt)
    want = (x[eo[0]] - x[dst] + r) * (2 * r + 1) + (y[eo[0]] - y[dst] + r)
    assert np.array_equal(rows["nbr_cls"].cpu().numpy()[:n].reshape(n, K)[mask].astype(np.int64), want)


@pytest.mark.parametrize("Q", [128, 16])
def test_sparse_scene_with_hot_pixels(Q):
    """A sparse scene (row-span search) with one pixel far beyond the FIFO capacity and one whose row span exceeds the
    32-record masks: the events around them go through the walk list, everything else through the row spans."""
    rng = np.random.default_rng(17)
    W, H, n = 64, 48, 4000
    x, y = rng.integers(0, W, n), rng.integers(0, H, n)
    x[:300] = 20; y[:300] = 11                          # 300 events on one pixel
    x[300:345] = 40; y[300:345] = 30                    # 45 on another
    t = np.sort(rng.integers(0, 30_000, n)) // 5 * 5
    perm = rng.permutation(n)
    pos = np.stack([x[perm], y[perm], t], 1).astype(np.int32)
    batch = np.zeros(n, np.int32)
    e, eo, _ = _run_both(batch, pos, W, H, 1, K=16, Q=Q, r=4, dt=10_000)
    assert np.array_equal(e.cpu().numpy(), eo)


def test_graph_search_random_configs():
    """40 seeded random configurations (sensor size, density, K, Q, radius, time gate, batch size, ties, unsorted time):
    whatever mix of row-span search and per-cell walk the build picks, edge_index equals the C oracle's."""
    rng = np.random.default_rng(2024)
    for case in range(40):
        W, H = int(rng.integers(3, 90)), int(rng.integers(3, 70))
        B = int(rng.integers(1, 4))
        dens = float(rng.choice([0.2, 0.8, 1.5, 3.0, 8.0]))
        n = max(1, min(int(W * H * B * dens), 30_000))
        K = int(rng.integers(2, 33))
        Q = int(rng.choice([1, 3, 8, 31, 32, 33, 128, 300]))
        r = int(rng.integers(1, 9))                       # 8: beyond the row-span kernel, walk only
        dt = int(rng.choice([0, 50, 700, 5000, 10_000, 40_000]))
        x, y = rng.integers(0, W, n), rng.integers(0, H, n)
        if case % 3 == 0:                                 # a few hot pixels
            hot = rng.integers(0, n, max(1, n // 10))
            x[hot], y[hot] = x[hot[0]], y[hot[0]]
        t = rng.integers(0, 30_000, n)
        if case % 4:
            t = np.sort(t)
        if case % 5 == 0:
            t = t // 16 * 16
        batch = np.sort(rng.integers(0, B, n)).astype(np.int32)
        pos = np.stack([x, y, t], 1).astype(np.int32)
        e, eo, _ = _run_both(batch, pos, W, H, B, K=K, Q=Q, r=r, dt=dt)
        assert np.array_equal(e.cpu().numpy(), eo), (case, W, H, B, n, K, Q, r, dt)
