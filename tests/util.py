"""Shared helpers of the parity tests: seeded models, inputs and oracle calls."""
import numpy as np
import torch

from eventad_b200 import synthetic
from eventad_b200.config import default_args
from eventad_b200.data import Batch


def randomize_bn(module, seed=0):
    """BN running stats / affine away from identity so that eval BatchNorm is a real transform."""
    g = torch.Generator().manual_seed(seed)
    for m in module.modules():
        if isinstance(m, torch.nn.BatchNorm1d):
            m.running_mean.copy_(torch.randn(m.num_features, generator=g) * 0.1)
            m.running_var.copy_(torch.rand(m.num_features, generator=g) + 0.5)
            m.weight.data.copy_(torch.rand(m.num_features, generator=g) + 0.5)
            m.bias.data.copy_(torch.randn(m.num_features, generator=g) * 0.1)


def make_backbone(seed=0, use_image=True, height=240, width=320, **over):
    from eventad_b200.model.networks.dagr import DAGR
    torch.manual_seed(seed)
    args = default_args(use_image=use_image, **over)
    model = DAGR(args, height, width)
    randomize_bn(model.backbone, seed)
    model.eval()
    return model, args


def net_state_dict(model):
    """state dict with keys relative to Net (what oracle/dagr_net.py reads), on the CPU."""
    return {k[len("backbone."):]: v.detach().cpu().clone() for k, v in model.state_dict().items()
            if k.startswith("backbone.") and not k.startswith("backbone.net.")}


def make_data(n_events, batch_size=1, seed=0, dist="U", width=320, height=240, device="cpu"):
    """Formatted input (after format_data): x [N,1] f32, pos [N,3] f32 normalised, batch int64."""
    raw = synthetic.make_batch(n_events, batch_size, seed=seed, dist=dist, width=width, height=height)
    pos = torch.from_numpy(raw["pos"].astype(np.float32))
    t = torch.from_numpy(raw["t"].astype(np.float32)).view(-1, 1)
    normalizer = torch.tensor([raw["width"], raw["height"], raw["time_window"]])
    d = Batch(x=torch.from_numpy(raw["x"].astype(np.float32)), pos=torch.cat([pos, t], -1) / normalizer,
              batch=torch.from_numpy(raw["batch"]))
    d.num_graphs = batch_size
    d.width = torch.tensor([width] * batch_size)
    d.height = torch.tensor([height] * batch_size)
    d.time_window = torch.tensor([raw["time_window"]] * batch_size)
    return d.to(device) if device != "cpu" else d


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
