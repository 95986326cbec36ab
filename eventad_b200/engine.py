"""Serving-side runner of the event branch: host event buffers in, per-window graph features out.

`EventBranchRunner.infer(raw)` is the end-to-end call `bench.py` times for the `e2e` number: pinned host buffers with
the reference's raw Data layout (pos int16 [N,2], t int32 [N], x int8 [N,1] polarity, batch int64 [N];
/root/reference/src/dagr/data/utils.py:6-20) -> one H2D copy each -> `format_data`
(/root/reference/src/dagr/utils/buffers.py:33-44) on the device -> fused `Net.forward` -> D2H of the two output graphs
(`out3`, `out4`: x, pos, batch) that `EventADModel.extract_features` consumes (/root/reference/models/EventAD.py:357-367).
"""
import numpy as np
import torch

from .data import Batch


def pin_raw(raw):
    """numpy raw batch (synthetic.make_batch) -> dict of pinned host tensors."""
    out = {}
    for k in ("pos", "t", "x", "batch"):
        t = torch.from_numpy(np.ascontiguousarray(raw[k]))
        out[k] = t.pin_memory() if torch.cuda.is_available() else t
    for k in ("num_graphs", "width", "height", "time_window"):
        out[k] = raw[k]
    return out


def raw_bytes(raw_pinned):
    return sum(raw_pinned[k].numel() * raw_pinned[k].element_size() for k in ("pos", "t", "x", "batch"))


class EventBranchRunner:
    def __init__(self, dagr_model, device="cuda"):
        self.model = dagr_model.to(device).eval()
        self.net = self.model.backbone
        self.device = torch.device(device)

    def to_device(self, raw_pinned):
        """H2D of one raw batch + format_data on the device -> Batch ready for Net.forward."""
        dev = self.device
        pos = raw_pinned["pos"].to(dev, non_blocking=True)
        t = raw_pinned["t"].to(dev, non_blocking=True)
        x = raw_pinned["x"].to(dev, non_blocking=True)
        batch = raw_pinned["batch"].to(dev, non_blocking=True)
        b = raw_pinned["num_graphs"]
        normalizer = torch.tensor([raw_pinned["width"], raw_pinned["height"], raw_pinned["time_window"]], device=dev)
        d = Batch(x=x.float(), pos=torch.cat([pos, t.view(-1, 1)], dim=-1) / normalizer, batch=batch)
        d.num_graphs = b
        d.width = [raw_pinned["width"]] * b
        d.height = [raw_pinned["height"]] * b
        d.time_window = [raw_pinned["time_window"]] * b
        return d

    @torch.no_grad()
    def forward_device(self, data, image_feat=None):
        res = self.net(data, image_feat=image_feat)
        return res[0] if self.net.use_image else res

    @torch.no_grad()
    def infer(self, raw_pinned, image_feat=None):
        """End to end: host buffers -> device -> event branch -> host results.  Returns (results, d2h_bytes)."""
        data = self.to_device(raw_pinned)
        outs = self.forward_device(data, image_feat)
        host, nbytes = [], 0
        for o in outs:
            h = dict(x=o.x.cpu(), pos=o.pos.cpu(), batch=o.batch.cpu())
            nbytes += sum(v.numel() * v.element_size() for v in h.values())
            host.append(h)
        return host, nbytes
