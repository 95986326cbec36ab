"""Generates tests/golden/ref_async_*.npz: outputs of THE REFERENCE'S OWN ASYNCHRONOUS MODEL (`make_model_asynchronous`,
/root/reference/src/dagr/asynchronous, executed unmodified on the CPU through oracle/ref_py.py) after a reset and after
one micro-batch (per micro-batch size) of a streamed window.  tests/test_incremental_gpu.py compares IncrementalBackbone with them.

    python tests/golden/make_ref_async_golden.py        (build container only: needs /root/reference; ~1 min)
"""
import copy
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
from eventad_b200 import synthetic  # noqa: E402

W, H = 64, 48
# name, model seed, data seed, distribution, events of the reset call, micro-batches
CASES = [("u", 0, 3, "U", 2000, (1, 10, 100, 889)), ("c", 2, 5, "C", 1500, (3, 50, 700))]


def case_inputs(case):
    name, mseed, dseed, dist, n0, steps = case
    model, args = synthetic.make_backbone(seed=mseed, use_image=True, height=H, width=W)
    # one event more than is ever streamed: the window's last event (t == 1.0) gets a voxel of its own in the dense 3-D
    # grid_cluster but not in the asynchronous pooling's 2-D cell index (max_pool.py:245-252); the reference's
    # two modes disagree on it, so it is kept out of the comparison
    d = synthetic.make_data(n0 + max(steps) + 1, 1, seed=dseed, dist=dist, width=W, height=H)
    feats = synthetic.make_image_features(1, seed=dseed + 100, width=W, height=H)
    return model, args, d, feats


def part(d, a, b):
    p = copy.copy(d)
    p.x, p.pos, p.batch = d.x[a:b].clone(), d.pos[a:b].clone(), d.batch[a:b].clone()
    p.image = torch.zeros(1)
    return p


def main():
    from oracle import ref_py
    torch.set_grad_enabled(False)
    out_dir = os.path.dirname(os.path.abspath(__file__))
    for case in CASES:
        name, mseed, dseed, dist, n0, steps = case
        model, args, d, feats = case_inputs(case)
        sd = {k: v.detach().clone() for k, v in model.state_dict().items()}
        ref = ref_py.build_dagr(args, H, W, sd, cache_luts=True)
        from dagr.asynchronous import make_model_asynchronous, make_model_synchronous

        class _Feats(torch.nn.Module):
            def forward(self, image):
                return list(feats), [feats[-2], feats[-1]]
        ref.backbone.net = _Feats()
        # the reference's asynchronous layers support ONE update after their initialisation (asynchronous/cartesian.py:40
        # concatenates the new positions to the INITIAL graph's): every micro-batch size is its own reset + update pair
        z = {}
        for s, m in enumerate(steps):
            net = make_model_asynchronous(ref.backbone)
            try:
                net(part(d, 0, n0), reset=True)
                out = net(part(d, n0, n0 + m), reset=False)
            finally:
                make_model_synchronous(ref.backbone)
            out = out[0] if isinstance(out, tuple) else out
            for lvl, o in zip((3, 4), out):
                z[f"s{s}_out{lvl}_pos"], z[f"s{s}_out{lvl}_x"] = o.pos.numpy().copy(), o.x.numpy().copy()
        done = n0
        np.savez_compressed(os.path.join(out_dir, f"ref_async_{name}.npz"), **z)
        print(name, done, [z[f"s{s}_out4_x"].shape for s in range(len(steps))])


if __name__ == "__main__":
    main()
