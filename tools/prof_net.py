"""One fused forward at B windows x N events (default 8 x 100k, U): the workload of the per-kernel ncu captures.
Usage: python tools/prof_net.py [B] [N] [dist]"""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from eventad_b200 import synthetic  # noqa: E402

B = int(sys.argv[1]) if len(sys.argv) > 1 else 8
N = int(sys.argv[2]) if len(sys.argv) > 2 else 100_000
dist = sys.argv[3] if len(sys.argv) > 3 else "U"
torch.set_grad_enabled(False)
model, args = synthetic.make_backbone(seed=0, use_image=True)
model.cache_luts(320, 240, args.radius)
net = model.cuda().backbone
feats = [f.cuda() for f in synthetic.make_image_features(B, seed=0)]
for it in range(3):
    d = synthetic.make_data(N, B, seed=it, dist=dist, device="cuda")
    outs = net(d, image_feat=feats)[0]
torch.cuda.synchronize()
print("ok", [tuple(o.x.shape) for o in outs])
