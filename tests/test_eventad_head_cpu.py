"""EventAD head on the CPU: the batched, masked formulation of eventad_b200/models/EventAD.py:head against
the per-track python loops of THE REFERENCE'S OWN models/EventAD.py, through the committed logits
tests/golden/ref_eventad_*.npz (generator tests/golden/make_ref_eventad_golden.py runs the reference file unmodified).
The box-pooled features come from the golden file here (the pooling kernel is CUDA only: tests/test_eventad_gpu.py)."""
import importlib.util
import os

import numpy as np
import pytest
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
spec = importlib.util.spec_from_file_location("make_ref_eventad_golden", os.path.join(HERE, "golden", "make_ref_eventad_golden.py"))
gold = importlib.util.module_from_spec(spec)
spec.loader.exec_module(gold)
CASES = {c[0]: c for c in gold.CASES}


@pytest.mark.parametrize("name", sorted(CASES))
def test_batched_head_matches_reference_logits(name):
    z = np.load(os.path.join(HERE, "golden", f"ref_eventad_{name}.npz"))
    ead, args, d, feats, labels = gold.case_inputs(CASES[name])
    assert gold.sd_sha({k: v for k, v in ead.state_dict().items()}) == str(z["sd_sha"])
    bf = torch.from_numpy(z["batch_features"])
    logits, valid, target, loss = ead.head(bf, d, labels)
    per = valid.sum(1)
    assert per[per > 0].tolist() == z["per_frame"].tolist()
    np.testing.assert_allclose(logits[valid].numpy(), z["logits"], rtol=1e-4, atol=1e-5)      # row-major = (frame, track id)
    assert target[valid].tolist() == z["labels"].tolist()
    np.testing.assert_allclose(float(loss), float(z["loss"]), rtol=1e-5)


def test_state_dict_keys_of_the_head():
    """the trainable head of best_rol.pth['model'] (1 115 234 parameters)."""
    ead, *_ = gold.case_inputs(CASES["clip_s"])
    head = {k: v for k, v in ead.state_dict().items() if not k.startswith("dagr_model.")}
    assert sum(v.numel() for v in head.values()) == 1_115_234
    assert {k.split(".")[0] for k in head} == {"gru_net_event", "gru_net_cor", "fusion_module", "soft_attention", "soft_attention_cor"}
    assert all(not p.requires_grad for p in ead.dagr_model.parameters())
    assert all(p.requires_grad for n, p in ead.named_parameters() if not n.startswith("dagr_model."))


def test_head_is_differentiable_and_masks_gradients():
    """Training step of BASELINE configs[3]: gradients reach the head parameters through the valid slots only."""
    z = np.load(os.path.join(HERE, "golden", "ref_eventad_clip_s.npz"))
    ead, args, d, feats, labels = gold.case_inputs(CASES["clip_s"])
    with torch.enable_grad():
        logits, valid, target, loss = ead.head(torch.from_numpy(z["batch_features"]), d, labels)
        loss.backward()
    g = ead.gru_net_event.gru.weight_ih_l0.grad
    assert g is not None and bool(torch.isfinite(g).all()) and float(g.abs().sum()) > 0
    assert ead.fusion_module.attention.in_proj_weight.grad is None          # never called
