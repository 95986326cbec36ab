"""Pins the C oracle (oracle/evg_oracle.c) to the UNMODIFIED reference kernels: oracle/_ref/ev_graph_cuda.so is
/root/reference/src/dagr/graph/ev_graph.cu compiled as-is for sm_100a by oracle/build_ref.py.  The host glue around
the three pybind functions restates src/dagr/graph/utils.py:6-23 and ev_graph.py:63-103 (sort, run lengths, mask)."""
import numpy as np
import pytest
import torch

from eventad_b200 import synthetic

pytestmark = pytest.mark.gpu
from tests.util import REF_GRAPH_SO as REF, RefGraph, load_ref_graph_module


def _ref_module():
    m = load_ref_graph_module()
    if m is None:
        pytest.skip("oracle/_ref/ev_graph_cuda.so not built (python oracle/build_ref.py)")
    return m


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
