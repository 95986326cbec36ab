"""Generates tests/golden/ref_eventad_*.npz: per-frame anomaly logits of THE REFERENCE'S OWN EventADModel
(/root/reference/models/EventAD.py, imported unmodified) on top of the reference's own DAGR Python (oracle/ref_py.py), on the
CPU.  north_star: "final per-frame anomaly scores must match within rtol 1e-4 / atol 1e-5".

    python tests/golden/make_ref_eventad_golden.py          (build container only: needs /root/reference; ~1 min)

The CNN feature maps are a seeded input on both sides (the ResNet branch is out of scope and runs cuDNN on the GPU), weights
and data are regenerated from seeds (eventad_b200/synthetic.py); `sd_sha` guards the weights."""
import hashlib
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
from eventad_b200 import synthetic  # noqa: E402

# name, W, H, events/window, B (consecutive windows of one clip), dist, seed, boxes per frame, drop probability
CASES = [
    ("clip_s", 64, 48, 2500, 5, "C", 1, 6, 0.3),
    ("clip_f", 320, 240, 20_000, 6, "C", 2, 8, 0.2),       # ROL-shaped: 320x240, batch 6 (config/eventad_config.py:46), 8 boxes
]


class FixedFeats(torch.nn.Module):
    """Stands in for HookModule (net_img.py:122-135) on both sides: returns seeded CNN maps."""

    def __init__(self, feats):
        super().__init__()
        self.feats = feats
        self.feature_channels = [f.shape[1] for f in feats]

    def forward(self, image):
        return list(self.feats), [self.feats[-2], self.feats[-1]]


def sd_sha(sd):
    h = hashlib.sha256()
    for k in sorted(sd):
        h.update(k.encode())
        h.update(sd[k].detach().cpu().contiguous().numpy().tobytes())
    return h.hexdigest()


def case_inputs(case):
    name, W, H, n, B, dist, seed, nb, drop = case
    model, args = synthetic.make_backbone(seed=seed, use_image=True, height=H, width=W)
    torch.manual_seed(seed + 100)
    from eventad_b200.models.EventAD import EventADModel
    ead = EventADModel(model).eval()
    d = synthetic.make_data(n, B, seed=seed + 3, dist=dist, width=W, height=H)
    for k, v in synthetic.make_boxes(B, seed=seed + 5, n_boxes=nb, width=W, height=H, drop=drop).items():
        setattr(d, k, v)
    d.image = torch.zeros((B, 3, H, W))
    feats = synthetic.make_image_features(B, seed=seed + 7, width=W, height=H)
    labels = d.bbox[:, 4]
    return ead, args, d, feats, labels


def flatten(outputs, labels):
    lo = torch.cat([o for fr in outputs for o in fr], 0) if outputs else torch.zeros((0, 2))
    la = torch.cat([l.view(-1) for fr in labels for l in fr], 0) if labels else torch.zeros((0,), dtype=torch.long)
    return lo, la, np.array([len(fr) for fr in outputs], np.int64)


def main():
    from oracle import ref_py
    ref_py.install()
    from models.EventAD import EventADModel as RefEventADModel            # /root/reference/models/EventAD.py
    torch.set_grad_enabled(False)
    out_dir = os.path.dirname(os.path.abspath(__file__))
    for case in CASES:
        name, W, H, n, B, dist, seed, nb, drop = case
        ead, args, d, feats, labels = case_inputs(case)
        sd = {k: v.detach().clone() for k, v in ead.state_dict().items()}
        ref_dagr = ref_py.build_dagr(args, H, W, {k[len("dagr_model."):]: v for k, v in sd.items() if k.startswith("dagr_model.")})
        ref_dagr.backbone.net = FixedFeats(feats)
        ref = RefEventADModel(ref_dagr).eval()
        head_sd = {k: v for k, v in sd.items() if not k.startswith("dagr_model.")}
        ref_head_keys = {k for k in ref.state_dict() if not k.startswith("dagr_model.")}
        assert ref_head_keys == set(head_sd), (ref_head_keys ^ set(head_sd))
        missing, unexpected = ref.load_state_dict(head_sd, strict=False)
        assert not unexpected and all(k.startswith("dagr_model.") for k in missing)
        losses, outputs, out_labels = ref(d, labels, testing=True)
        bf = ref.batch_feature_cache
        lo, la, per = flatten(outputs, out_labels)
        np.savez_compressed(os.path.join(out_dir, f"ref_eventad_{name}.npz"), sd_sha=np.array(sd_sha(sd)), logits=lo.numpy(),
                            labels=la.numpy(), per_frame=per, loss=np.array(float(losses["cross_entropy"])),
                            batch_features=bf.numpy())
        print(name, tuple(lo.shape), per.tolist(), float(losses["cross_entropy"]), flush=True)


if __name__ == "__main__":
    main()
