"""The sm_100a event branch against outputs of THE REFERENCE'S OWN PYTHON (tests/golden/ref_py_*.npz: /root/reference/src/dagr
executed unmodified on the CPU by oracle/ref_py.py; generator tests/golden/make_ref_py_golden.py).  Graph structure,
cluster assignment (through pos / batch / coarse edges) bit-exact; fp32 features rtol 1e-4 / atol 1e-5 (north_star)."""
import importlib.util
import os

import numpy as np
import pytest
import torch

from tests.util import assert_close

pytestmark = pytest.mark.gpu
HERE = os.path.dirname(os.path.abspath(__file__))
spec = importlib.util.spec_from_file_location("make_ref_py_golden", os.path.join(HERE, "golden", "make_ref_py_golden.py"))
gold = importlib.util.module_from_spec(spec)
spec.loader.exec_module(gold)
CASES = {c[0]: c for c in gold.CASES}


@pytest.mark.parametrize("fused", [True, False])
@pytest.mark.parametrize("name", sorted(CASES))
def test_cuda_path_matches_reference_python_golden(name, fused):
    case = CASES[name]
    _, W, H, n, B, dist, mseed, dseed, use_image, luts, over = case
    z = np.load(os.path.join(HERE, "golden", f"ref_py_{name}.npz"))
    model, args, d, feats = gold.case_inputs(case)
    assert gold.sd_sha({k: v for k, v in model.state_dict().items()}) == str(z["sd_sha"])
    if luts:
        model.cache_luts(W, H, args.radius)
    net = model.cuda().backbone
    net.fused = fused
    cap = {}
    hooks = []
    if not fused or not net.fused_supported():
        hooks.append(net.events_to_graph.register_forward_hook(lambda m, i, o: cap.__setitem__("e0", o.edge_index.clone())))
        for k in range(1, 5):
            hooks.append(getattr(net, f"pool{k}").register_forward_hook(
                lambda m, i, o, k=k: cap.__setitem__(f"pool{k}", (o.pos.clone(), o.batch.clone(), o.edge_index.clone(), o.x.clone()))))
    res = net(d.to("cuda"), image_feat=[f.cuda() for f in feats] if use_image else None)
    outs = res[0] if use_image else res
    for h in hooks:
        h.remove()
    if "e0" in cap:
        assert cap["e0"].shape[1] == int(z["n_edges0"])
        assert gold.edges_sha(cap["e0"]) == str(z["edges0_sha"])
        for k in range(1, 5):
            pos, batch, ei, x = cap[f"pool{k}"]
            assert np.array_equal(batch.cpu().numpy(), z[f"pool{k}_batch"]), (name, k)
            assert np.array_equal(ei.cpu().numpy(), z[f"pool{k}_edge_index"]), (name, k)
            assert np.array_equal(pos.cpu().numpy(), z[f"pool{k}_pos"]), (name, k)
            if f"pool{k}_x" in z:
                assert_close(x, torch.from_numpy(z[f"pool{k}_x"]), what=f"{name} pool{k} x")
    for nm, o in zip(("out3", "out4"), outs):
        assert np.array_equal(o.edge_index.cpu().numpy(), z[f"{nm}_edge_index"]), (name, nm)
        assert np.array_equal(o.batch.cpu().numpy(), z[f"{nm}_batch"]), (name, nm)
        assert np.array_equal(o.pos.cpu().numpy(), z[f"{nm}_pos"]), (name, nm)
        assert_close(o.edge_attr, torch.from_numpy(z[f"{nm}_edge_attr"]), rtol=1e-6, atol=1e-6, what=f"{name} {nm} edge_attr")
        assert_close(o.x, torch.from_numpy(z[f"{nm}_x"]), what=f"{name} {nm} x")
