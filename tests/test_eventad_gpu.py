"""Full EventAD inference (BASELINE configs[1]: event GNN + CNN feature maps + anomaly head) on the B200 against THE
REFERENCE'S OWN models/EventAD.py + src/dagr Python (tests/golden/ref_eventad_*.npz, generated on the CPU by
tests/golden/make_ref_eventad_golden.py): box-pooled features and per-frame anomaly logits within north_star's
rtol 1e-4 / atol 1e-5; which (frame, track) slots produce a prediction and their labels exactly."""
import importlib.util
import os

import numpy as np
import pytest
import torch

from tests.util import assert_close

pytestmark = pytest.mark.gpu
HERE = os.path.dirname(os.path.abspath(__file__))
spec = importlib.util.spec_from_file_location("make_ref_eventad_golden", os.path.join(HERE, "golden", "make_ref_eventad_golden.py"))
gold = importlib.util.module_from_spec(spec)
spec.loader.exec_module(gold)
CASES = {c[0]: c for c in gold.CASES}


def _run(name, fused=True):
    z = np.load(os.path.join(HERE, "golden", f"ref_eventad_{name}.npz"))
    ead, args, d, feats, labels = gold.case_inputs(CASES[name])
    assert gold.sd_sha({k: v for k, v in ead.state_dict().items()}) == str(z["sd_sha"])
    _, W, H = CASES[name][:3]
    ead.dagr_model.cache_luts(W, H, args.radius)
    ead = ead.cuda()
    ead.dagr_model.backbone.net = gold.FixedFeats([f.cuda() for f in feats])
    ead.dagr_model.backbone.fused = fused
    d = d.to("cuda")
    losses, outputs, out_labels = ead(d, labels.cuda(), testing=True)
    return z, ead, losses, outputs, out_labels


@pytest.mark.parametrize("fused", [True, False])
@pytest.mark.parametrize("name", sorted(CASES))
def test_anomaly_logits_match_reference(name, fused):
    z, ead, losses, outputs, out_labels = _run(name, fused)
    assert_close(ead.batch_feature_cache, torch.from_numpy(z["batch_features"]), what=f"{name} box features")
    lo, la, per = gold.flatten(outputs, out_labels)
    assert per.tolist() == z["per_frame"].tolist()
    assert la.cpu().tolist() == z["labels"].tolist()
    assert_close(lo, torch.from_numpy(z["logits"]), what=f"{name} anomaly logits")
    assert abs(float(losses["cross_entropy"]) - float(z["loss"])) <= 1e-4 * abs(float(z["loss"])) + 1e-5


def test_feature_cache_quirk_is_reproduced():
    """EventAD.py:349-354: a second batch whose polarity sum equals the first one's gets the cached features."""
    z, ead, *_ = _run("clip_s")
    first = ead.batch_feature_cache
    ead2, args, d, feats, labels = gold.case_inputs(CASES["clip_s"])
    d = d.to("cuda")
    d.pos = d.pos.flip(0).contiguous()                           # different events, same sum of polarities
    assert ead.extract_features(d) is first
    ead.emulate_feature_cache = False
    assert ead.extract_features(d) is not first


def test_bbox_pool_edge_cases():
    """Duplicate (sample, track) boxes: the last one wins; track ids outside 1..30 are skipped; a sample without nodes
    gets zeros; a box without nodes inside takes the mean of the 5 nearest nodes (fewer if the sample has fewer)."""
    from eventad_b200.data import Batch
    from eventad_b200.models.EventAD import EventADModel
    ead = EventADModel(torch.nn.Linear(1, 1)).cuda()
    g = torch.Generator().manual_seed(0)
    feat = torch.randn(9, 64, generator=g)
    pos = torch.tensor([[0.1, 0.1], [0.2, 0.1], [0.3, 0.3], [0.8, 0.8], [0.82, 0.8], [0.5, 0.5],      # sample 0
                        [0.4, 0.4], [0.6, 0.6], [0.9, 0.1]])                                           # sample 2 (3 nodes)
    nb = torch.tensor([0, 0, 0, 0, 0, 0, 2, 2, 2])
    out4 = Batch(x=feat.cuda(), pos=torch.cat([pos, torch.ones(9, 1)], 1).cuda(), batch=nb.cuda())
    W, H = 100.0, 100.0
    boxes = torch.tensor([[5., 5., 20., 10., 0., 1.],      # sample 0 track 1: nodes 0, 1 inside
                          [70., 70., 20., 20., 1., 1.],    # sample 0 track 1 again: nodes 3, 4 -> this one wins
                          [40., 10., 5., 5., 0., 2.],      # sample 0 track 2: empty box -> 5 nearest of 6
                          [10., 10., 5., 5., 0., 31.],     # track id out of range: skipped
                          [10., 10., 50., 50., 0., 3.],    # sample 1: no nodes -> zeros
                          [0., 0., 1., 1., 0., 4.]])       # sample 2 track 4: empty box, only 3 nodes -> mean of 3
    bb = torch.tensor([0, 0, 0, 0, 1, 2])
    data = Batch(bbox=boxes.cuda(), bbox_batch=bb.cuda())
    data.width, data.height = torch.tensor([W] * 3), torch.tensor([H] * 3)
    bf = ead.pool_boxes(out4, data, 3).cpu()

    def ref_in_box(ids, box):
        cx, cy = (box[0] + box[2] / 2) / W, (box[1] + box[3] / 2) / H
        dist = torch.norm(pos[ids] - torch.tensor([cx, cy]), dim=1)
        w = 1.0 / (dist + 1e-6)
        w = w / w.sum()
        return (feat[ids] * w[:, None]).sum(0)
    assert_close(bf[0, 1, 1], ref_in_box([3, 4], boxes[1]), what="last duplicate wins")
    c = torch.tensor([0.425, 0.125])
    near = torch.topk(torch.norm(pos[:6] - c, dim=1), 5, largest=False)[1]
    assert_close(bf[0, 1, 2], feat[:6][near].mean(0), what="5 nearest")
    assert float(bf[:, :, 31:].abs().sum()) == 0 and float(bf[1].abs().sum()) == 0
    assert_close(bf[2, 1, 4], feat[6:9].mean(0), what="fewer than 5 nodes")
    assert float(bf[:, 0].abs().sum()) == 0                                  # no bbox0 given
