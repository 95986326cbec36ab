"""Pins the C oracle (oracle/evg_oracle.c) to the UNMODIFIED reference kernels: oracle/_ref/ev_graph_cuda.so is
/root/reference/src/dagr/graph/ev_graph.cu compiled as-is for sm_100a by oracle/build_ref.py.  The host glue around
the three pybind functions restates src/dagr/graph/utils.py:6-23 and ev_graph.py:63-103 (sort, run lengths, mask)."""
import importlib.util
import os

import numpy as np
import pytest
import torch

from eventad_b200 import synthetic

pytestmark = pytest.mark.gpu
REF = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "oracle", "_ref", "ev_graph_cuda.so")


def _ref_module():
    if not os.path.exists(REF):
        pytest.skip("oracle/_ref/ev_graph_cuda.so not built (python oracle/build_ref.py)")
    spec = importlib.util.spec_from_file_location("ev_graph_cuda", REF)
    m = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(m)
    return m


class RefGraph:
    """AsyncGraph.forward driven with the reference's own kernels."""

    def __init__(self, m, W, H, B, K, Q, r, dt):
        self.m, self.K, self.r, self.dt = m, K, r, dt
        self.queue = torch.full((B, Q, H, W), -1, dtype=torch.int32, device="cuda")
        self.all_ts = torch.zeros(0, dtype=torch.int32, device="cuda")
        self.max_index = 0
        self.min_index = 0

    def forward(self, batch, pos):
        n = len(batch)
        self.all_ts = torch.cat([self.all_ts, pos[:, 2]])
        indices = (self.max_index + torch.arange(n, dtype=torch.int32, device="cuda")).int()
        self.max_index += n
        H, W = self.queue.shape[-2:]
        if n > 1:
            lin = pos[:, 0] + W * pos[:, 1] + W * H * batch
            s, si = torch.sort(lin, stable=True)
            sidx = indices[si].int()
            u, c = torch.unique_consecutive(s, return_counts=True)
            cs = torch.cumsum(c, 0).int()
            self.queue = self.m.insert_in_queue_cuda(sidx, u, cs, self.queue)
        else:
            self.queue = self.m.insert_in_queue_single_cuda(indices, pos, self.queue)
        edges = torch.full((2, self.K * n), -1, dtype=torch.int64, device="cuda")
        self.m.fill_edges_cuda(batch, pos, self.all_ts.contiguous(), self.queue, indices, self.K, self.r, self.dt,
                               edges, self.min_index)
        return edges[:, edges[1] >= 0]


@pytest.mark.parametrize("n,dist,B,Q", [(2000, "U", 1, 128), (50_000, "C", 1, 128), (100_000, "U", 2, 128),
                                        (30_000, "C", 1, 8)])
def test_oracle_and_kernel_match_reference_kernel(n, dist, B, Q):
    from eventad_b200.graph.ev_graph import SlidingWindowGraph
    from oracle.evg import OracleSlidingWindowGraph
    m = _ref_module()
    raw = synthetic.make_batch(n, B, seed=1, dist=dist)
    W, H = raw["width"], raw["height"]
    pos = np.concatenate([raw["pos"].astype(np.int32), raw["t"].reshape(-1, 1)], 1).astype(np.int32)
    batch = raw["batch"].astype(np.int32)
    bt, pt = torch.from_numpy(batch).cuda(), torch.from_numpy(pos).cuda()
    ref = RefGraph(m, W, H, B, 16, Q, 4, 10000).forward(bt, pt).cpu().numpy()
    o = OracleSlidingWindowGraph(width=W, height=H, batch_size=B, max_num_neighbors=16, max_queue_size=Q, radius=4,
                                 delta_t_us=10000)
    eo = o.forward(batch, pos, delete_nodes=False)
    assert np.array_equal(ref, eo)
    g = SlidingWindowGraph(width=W, height=H, batch_size=B, max_num_neighbors=16, max_queue_size=Q, radius=4,
                           delta_t_us=10000)
    e = g.forward(bt, pt, delete_nodes=False).cpu().numpy()
    assert np.array_equal(ref, e)


def test_streaming_oracle_matches_reference_kernel():
    """Append mode (reset=False): three micro-batches, then single events (the b=0 quirk path)."""
    from oracle.evg import OracleSlidingWindowGraph
    m = _ref_module()
    raw = synthetic.make_batch(6000, 1, seed=2, dist="C", width=64, height=48)
    W, H = 64, 48
    pos = np.concatenate([raw["pos"].astype(np.int32), raw["t"].reshape(-1, 1)], 1).astype(np.int32)
    batch = raw["batch"].astype(np.int32)
    ref = RefGraph(m, W, H, 1, 16, 16, 4, 10000)
    o = OracleSlidingWindowGraph(width=W, height=H, batch_size=1, max_num_neighbors=16, max_queue_size=16, radius=4,
                                 delta_t_us=10000)
    cuts = [0, 3000, 5000, 5990] + list(range(5991, 6001))
    for a, b in zip(cuts[:-1], cuts[1:]):
        er = ref.forward(torch.from_numpy(batch[a:b]).cuda(), torch.from_numpy(pos[a:b]).cuda()).cpu().numpy()
        eo = o.forward(batch[a:b], pos[a:b], delete_nodes=False)
        assert np.array_equal(er, eo), (a, b)
