"""Generates tests/golden/ref_py_*.npz by EXECUTING THE REFERENCE'S OWN PYTHON (oracle/ref_py.py: /root/reference/src/dagr
imported unmodified; only the uninstallable third-party ops and the CUDA-only graph extension are stand-ins) on the CPU.

    python tests/golden/make_ref_py_golden.py            (build container only: needs /root/reference; ~2 min, ~15 GB RAM)

The fixtures pin oracle/dagr_layers.py / dagr_net.py (tests/test_ref_py_cpu.py, also live against the reference when it is
mounted) and the sm_100a path (tests/test_ref_py_golden_gpu.py) to outputs of the reference's text.  Inputs and weights are
regenerated from seeds (eventad_b200/synthetic.py); `sd_sha` guards the weights."""
import hashlib
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
from eventad_b200 import synthetic  # noqa: E402

# name, W, H, events/window, B, dist, model seed, data seed, use_image, cache_luts, pooling overrides
CASES = [
    ("s_u3000_b1", 64, 48, 3000, 1, "U", 0, 3, True, True, {}),
    ("s_c2500_b3", 64, 48, 2500, 3, "C", 1, 4, True, True, {}),          # t == 1.0 quirk across samples
    ("s_c3000_b2_noimg", 64, 48, 3000, 2, "C", 2, 5, False, True, {}),
    ("s_c3000_b1_basis", 64, 48, 3000, 1, "C", 3, 6, True, False, {}),   # PyG basis message (no cache_luts)
    ("s_c3000_b2_kto", 64, 48, 3000, 2, "C", 4, 7, True, True, {"keep_temporal_ordering": True}),
    ("f_u50000_b1", 320, 240, 50_000, 1, "U", 0, 11, True, True, {}),    # BASELINE configs[0]
    ("f_c20000_b2", 320, 240, 20_000, 2, "C", 1, 12, True, True, {}),
]


def sd_sha(sd):
    h = hashlib.sha256()
    for k in sorted(sd):
        h.update(k.encode())
        h.update(sd[k].detach().cpu().contiguous().numpy().tobytes())
    return h.hexdigest()


def case_inputs(case):
    name, W, H, n, B, dist, mseed, dseed, use_image, luts, over = case
    model, args = synthetic.make_backbone(seed=mseed, use_image=use_image, height=H, width=W, **over)
    d = synthetic.make_data(n, B, seed=dseed, dist=dist, width=W, height=H)
    feats = synthetic.make_image_features(B, seed=dseed + 100, width=W, height=H) if use_image else None
    return model, args, d, feats


def edges_sha(ei):
    return hashlib.sha256(np.ascontiguousarray(ei.cpu().numpy().astype(np.int64)).tobytes()).hexdigest()


def main():
    from oracle import ref_py
    torch.set_grad_enabled(False)
    out_dir = os.path.dirname(os.path.abspath(__file__))
    for case in CASES:
        name, W, H, n, B, dist, mseed, dseed, use_image, luts, over = case
        model, args, d, feats = case_inputs(case)
        sd = {k: v.detach().clone() for k, v in model.state_dict().items()}
        ref = ref_py.build_dagr(args, H, W, sd, cache_luts=luts)
        cap = {}
        hooks = [ref.backbone.events_to_graph.register_forward_hook(
            lambda m, i, o: cap.__setitem__("edge_index0", o.edge_index.clone()))]
        for k in range(1, 5):
            hooks.append(getattr(ref.backbone, f"pool{k}").register_forward_hook(
                lambda m, i, o, k=k: cap.__setitem__(f"pool{k}", (o.pos.clone(), o.batch.clone(), o.edge_index.clone(),
                                                                  o.x.clone()))))
        hooks.append(ref.backbone.conv_block1.register_forward_hook(lambda m, i, o: cap.__setitem__("x1", o.x.clone())))
        outs = ref_py.run_backbone(ref, d, feats)
        for h in hooks:
            h.remove()
        z = dict(sd_sha=np.array(sd_sha(sd)), n_edges0=np.array(cap["edge_index0"].shape[1]),
                 edges0_sha=np.array(edges_sha(cap["edge_index0"])),
                 x1_sample=cap["x1"][:: max(1, cap["x1"].shape[0] // 2048)].numpy(),
                 x1_stride=np.array(max(1, cap["x1"].shape[0] // 2048)))
        if W <= 64:
            z["edge_index0"] = cap["edge_index0"].numpy().astype(np.int32)
        for k in range(1, 5):
            pos, batch, ei, x = cap[f"pool{k}"]
            z[f"pool{k}_pos"], z[f"pool{k}_batch"], z[f"pool{k}_edge_index"] = pos.numpy(), batch.numpy().astype(np.int32), ei.numpy().astype(np.int32)
            if k > 1:
                z[f"pool{k}_x"] = x.numpy()
        for nm, o in zip(("out3", "out4"), outs):
            z[f"{nm}_x"], z[f"{nm}_pos"], z[f"{nm}_batch"] = o.x.numpy(), o.pos.numpy(), o.batch.numpy().astype(np.int32)
            z[f"{nm}_edge_index"], z[f"{nm}_edge_attr"] = o.edge_index.numpy().astype(np.int32), o.edge_attr.numpy()
        np.savez_compressed(os.path.join(out_dir, f"ref_py_{name}.npz"), **z)
        print(name, int(z["n_edges0"]), [tuple(o.x.shape) for o in outs], flush=True)
        del ref


if __name__ == "__main__":
    main()
