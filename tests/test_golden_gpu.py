"""sm_100a graph builder against the golden vectors of the UNMODIFIED reference kernels (tests/golden/*.npz)."""
import glob
import os

import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu
GOLD = sorted(glob.glob(os.path.join(os.path.dirname(__file__), "golden", "ref_graph_*.npz")))


@pytest.mark.parametrize("path", GOLD, ids=[os.path.basename(p) for p in GOLD])
def test_kernel_matches_reference_golden(path):
    from eventad_b200.graph.ev_graph import SlidingWindowGraph
    z = np.load(path)
    W, H, B, Q, K, r, dt = [int(v) for v in z["params"]]
    g = SlidingWindowGraph(width=W, height=H, batch_size=B, max_num_neighbors=K, max_queue_size=Q, radius=r, delta_t_us=dt)
    cuts, off = z["cuts"], 0
    for (a, b), ne in zip(zip(cuts[:-1], cuts[1:]), z["edge_counts"]):
        e = g.forward(torch.from_numpy(z["batch"][a:b]).cuda(), torch.from_numpy(z["pos"][a:b]).cuda(), delete_nodes=False)
        assert np.array_equal(e.cpu().numpy(), z["edges"][:, off:off + ne].astype(np.int64)), (path, a, b)
        off += ne
