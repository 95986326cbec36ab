"""Shared helpers of the parity tests: seeded models, inputs and oracle calls."""
import numpy as np
import torch

from eventad_b200 import synthetic
from eventad_b200.config import default_args
from eventad_b200.data import Batch


from eventad_b200.synthetic import make_backbone, make_data, net_state_dict, randomize_bn  # noqa: E402,F401


def assert_close(a, b, rtol=1e-4, atol=1e-5, what=""):
    a = a.detach().cpu().double()
    b = b.detach().cpu().double()
    assert a.shape == b.shape, (what, a.shape, b.shape)
    if a.numel() == 0:
        return
    err = (a - b).abs()
    tol = atol + rtol * b.abs()
    bad = err > tol
    assert not bad.any(), f"{what}: {int(bad.sum())}/{a.numel()} beyond rtol={rtol} atol={atol}; max abs err {err.max():.3e}"


# ---------------------------------------------------------------- the reference's own CUDA graph builder (oracle/_ref)
import importlib.util as _ilu
import os as _os

REF_GRAPH_SO = _os.path.join(_os.path.dirname(_os.path.dirname(_os.path.abspath(__file__))), "oracle", "_ref",
                             "ev_graph_cuda.so")


def load_ref_graph_module():
    """oracle/_ref/ev_graph_cuda.so = /root/reference/src/dagr/graph/ev_graph.cu compiled unmodified by
    oracle/build_ref.py; None when it was not built."""
    if not _os.path.exists(REF_GRAPH_SO):
        return None
    spec = _ilu.spec_from_file_location("ev_graph_cuda", REF_GRAPH_SO)
    m = _ilu.module_from_spec(spec)
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

    def delete_nodes(self, n):
        """SlidingWindowGraph.delete_nodes bookkeeping the kernels see (ev_graph.py:121-125)."""
        self.all_ts = self.all_ts[n:]
        self.min_index += n
