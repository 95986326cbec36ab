"""Synthetic ROL-shaped event windows (SURVEY.md section 8d / BASELINE.md 3.3).

W x H = 320 x 240, T = 1_000_000 us, a window = the last 50 ms (t in [950000, 1000000], sorted, t[-1] = T,
mirroring /root/reference/src/dagr/data/dsec_data.py:124-130), polarity +-1, two spatial distributions:
  "U"  uniform pixels;
  "C"  ROL-like clustered: 70 % of the events from 24 drifting Gaussian blobs (sigma 6 px), 30 % uniform.
Raw layout follows the reference's Data before `format_data` (src/dagr/utils/buffers.py:33-44):
pos int16 [N,2] (x,y), t int32 [N], x int8 [N,1] (polarity), batch int64 [N], sample-major.
"""
from types import SimpleNamespace

import numpy as np

W_DEFAULT, H_DEFAULT, T_DEFAULT = 320, 240, 1_000_000


def make_window(n_events: int, seed: int = 0, dist: str = "U", width: int = W_DEFAULT, height: int = H_DEFAULT,
                time_window: int = T_DEFAULT, t_span: int = 50_000):
    rng = np.random.default_rng(seed)
    n = int(n_events)
    t = np.sort(rng.integers(time_window - t_span, time_window + 1, n)).astype(np.int32)
    if n > 0:
        t[-1] = time_window
    if dist == "U":
        x = rng.integers(0, width, n)
        y = rng.integers(0, height, n)
    elif dist == "C":
        n_blob = int(0.7 * n)
        k = 24
        cx = rng.uniform(0, width, k)
        cy = rng.uniform(0, height, k)
        vx = rng.uniform(-8, 8, k)
        vy = rng.uniform(-8, 8, k)
        which = rng.integers(0, k, n_blob)
        frac = (t[:n_blob].astype(np.float64) - (time_window - t_span)) / max(t_span, 1)
        bx = cx[which] + vx[which] * frac + rng.normal(0, 6.0, n_blob)
        by = cy[which] + vy[which] * frac + rng.normal(0, 6.0, n_blob)
        ux = rng.integers(0, width, n - n_blob)
        uy = rng.integers(0, height, n - n_blob)
        x = np.concatenate([np.clip(np.rint(bx), 0, width - 1).astype(np.int64), ux])
        y = np.concatenate([np.clip(np.rint(by), 0, height - 1).astype(np.int64), uy])
        perm = rng.permutation(n)                      # blob / uniform events interleaved in time
        x, y = x[perm], y[perm]
    else:
        raise ValueError(dist)
    p = (rng.integers(0, 2, n) * 2 - 1).astype(np.int8)
    return SimpleNamespace(x=x.astype(np.int16), y=y.astype(np.int16), t=t, p=p)


def make_batch(n_events_per_window, batch_size: int = 1, seed: int = 0, dist: str = "U",
               width: int = W_DEFAULT, height: int = H_DEFAULT, time_window: int = T_DEFAULT):
    """Raw (un-normalised) batch: dict of numpy arrays, windows concatenated sample-major."""
    if np.isscalar(n_events_per_window):
        n_events_per_window = [int(n_events_per_window)] * batch_size
    xs, ys, ts, ps, bs = [], [], [], [], []
    for b, n in enumerate(n_events_per_window):
        w = make_window(n, seed * 1000 + b, dist, width, height, time_window)
        xs.append(w.x); ys.append(w.y); ts.append(w.t); ps.append(w.p)
        bs.append(np.full(len(w.x), b, np.int64))
    return dict(pos=np.stack([np.concatenate(xs), np.concatenate(ys)], 1).astype(np.int16),
                t=np.concatenate(ts).astype(np.int32),
                x=np.concatenate(ps).astype(np.int8).reshape(-1, 1),
                batch=np.concatenate(bs),
                num_graphs=len(n_events_per_window), width=width, height=height, time_window=time_window)


def make_image_features(batch_size: int, channels=(16, 64, 64, 64, 64), width: int = W_DEFAULT,
                        height: int = H_DEFAULT, seed: int = 0):
    """Stand-in for HookModule's five 1x1-projected ResNet maps (net_img.py:122-135, shapes SURVEY A.6):
    strides 2,4,8,16,32 of the (H,W) image, float32, seeded."""
    import torch
    g = torch.Generator().manual_seed(seed)
    feats = []
    for i, c in enumerate(channels):
        s = 2 ** (i + 1)
        h, w = -(-height // s), -(-width // s)
        feats.append(torch.randn((batch_size, c, h, w), generator=g, dtype=torch.float32))
    return feats


# ------------------------------------------------------------------------------------- seeded models and formatted inputs
def randomize_bn(module, seed=0):
    """BN running stats / affine away from identity so that eval BatchNorm is a real transform."""
    import torch
    g = torch.Generator().manual_seed(seed)
    for m in module.modules():
        if isinstance(m, torch.nn.BatchNorm1d):
            m.running_mean.copy_(torch.randn(m.num_features, generator=g) * 0.1)
            m.running_var.copy_(torch.rand(m.num_features, generator=g) + 0.5)
            m.weight.data.copy_(torch.rand(m.num_features, generator=g) + 0.5)
            m.bias.data.copy_(torch.randn(m.num_features, generator=g) * 0.1)


def make_backbone(seed=0, use_image=True, height=H_DEFAULT, width=W_DEFAULT, **over):
    """Random-init dagr-s DAGR (there are no checkpoints on disk) with non-trivial BatchNorm statistics."""
    import torch
    from .config import default_args
    from .model.networks.dagr import DAGR
    torch.manual_seed(seed)
    args = default_args(use_image=use_image, **over)
    model = DAGR(args, height, width)
    randomize_bn(model.backbone, seed)
    model.eval()
    return model, args


def net_state_dict(model):
    """state dict with keys relative to Net (event branch only, no ResNet), on the CPU."""
    return {k[len("backbone."):]: v.detach().cpu().clone() for k, v in model.state_dict().items()
            if k.startswith("backbone.") and not k.startswith("backbone.net.")}


def make_data(n_events, batch_size=1, seed=0, dist="U", width=W_DEFAULT, height=H_DEFAULT, device="cpu"):
    """Formatted input (after format_data, buffers.py:33-44): x [N,1] f32, pos [N,3] f32 normalised, batch int64."""
    import torch
    from .data import Batch
    raw = make_batch(n_events, batch_size, seed=seed, dist=dist, width=width, height=height)
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


def make_boxes(batch_size, seed=0, n_boxes=8, width=W_DEFAULT, height=H_DEFAULT, drop=0.0, max_track=8):
    """ROL-shaped box annotations for a batch of consecutive windows: per sample `n_boxes` boxes
    [x, y, w, h, class, track_id] with w, h in [15, 80] px, track ids drawn from 1..max_track (each box kept with
    probability 1 - drop, so tracks appear and disappear along the clip), for the current frame (`bbox`) and the previous
    one (`bbox0`).  Returns dict(bbox [M,6] f32, bbox_batch [M] i64, bbox0, bbox0_batch)."""
    import torch
    rng = np.random.default_rng(seed)
    out = {}
    for key in ("bbox", "bbox0"):
        rows, bidx = [], []
        for b in range(batch_size):
            ids = rng.permutation(max_track)[:n_boxes] + 1
            for t in ids:
                if rng.random() < drop:
                    continue
                w, h = rng.uniform(15, 80, 2) * (width / W_DEFAULT)   # 15..80 px at the 320-wide ROL sensor
                x, y = rng.uniform(0, width - w), rng.uniform(0, height - h)
                rows.append([x, y, w, h, float(rng.integers(0, 2)), float(t)])
                bidx.append(b)
        out[key] = torch.tensor(np.asarray(rows, np.float32).reshape(-1, 6))
        out[key + "_batch"] = torch.tensor(np.asarray(bidx, np.int64))
    return out
