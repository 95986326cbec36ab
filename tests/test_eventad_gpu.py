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


@pytest.mark.parametrize("B,events,drop,overlap", [(2, 20_000, 0.0, False), (3, 8_000, 0.4, False), (1, 30_000, 0.0, True),
                                                   (3, 8_000, 0.4, True)])
def test_graphed_full_model_equals_eager_forward(B, events, drop, overlap):
    """engine.GraphedEventAD (CNN branch, event branch and box pooling + GRU head as three CUDA-graph replays) returns
    what EventADModel.forward returns on the same batches -- same samples with predictions, same track order, logits
    within the score tolerance (the CNN maps come from cuDNN twice: eager NCHW hooks vs captured channels-last) -- and
    keeps doing so on the next batch of the same shape (static buffers re-loaded, padding rows re-zeroed)."""
    from eventad_b200 import synthetic
    from eventad_b200.engine import GraphedEventAD
    from eventad_b200.models.EventAD import EventADModel
    model, args = synthetic.make_backbone(seed=0, use_image=True)
    model.cache_luts(320, 240, args.radius)
    torch.manual_seed(1)
    ead = EventADModel(model).cuda().eval()
    ead.emulate_feature_cache = False
    runner = None
    for s in range(3):
        d = synthetic.make_data(events, B, seed=40 + s, dist="U", device="cuda")
        for k, v in synthetic.make_boxes(B, seed=60 + s, drop=drop if s else 0.0).items():
            setattr(d, k, v.cuda())
        d.image = torch.rand((B, 3, 240, 320), generator=torch.Generator().manual_seed(s)).cuda()
        labels = d.bbox[:, 4]
        # fp32 convolutions on both sides: with cuDNN's default TF32 the eager and the captured ResNet pick different
        # algorithms and their maps differ by 1e-3 relative, which is the CNN's noise, not the runner's
        with torch.no_grad(), torch.backends.cudnn.flags(enabled=True, allow_tf32=False):
            l0, o0, t0 = ead(d, labels, testing=True)
            if runner is None:
                runner = GraphedEventAD(ead, B, d.pos.shape[0], overlap=overlap)   # overlap: both branches as one graph
            l1, o1, t1 = runner(d, labels)
        assert [len(a) for a in o0] == [len(a) for a in o1] and len(o0) > 0
        for a, b in zip(o0, o1):
            for x, y in zip(a, b):
                assert_close(y, x, what="graphed logits")
        assert [[int(v) for v in a] for a in t0] == [[int(v) for v in a] for a in t1]
        assert abs(float(l0["cross_entropy"]) - float(l1["cross_entropy"])) <= 1e-4 * abs(float(l0["cross_entropy"])) + 1e-4


def test_graphed_features_feed_the_training_step():
    """GraphedEventAD.features == EventADModel.extract_features, and train_step(feature_fn=...) takes the same step as the
    eager one: equal loss and equal head gradients from the same start."""
    import copy
    from eventad_b200 import synthetic, train
    from eventad_b200.engine import GraphedEventAD
    from eventad_b200.models.EventAD import EventADModel
    B, events = 2, 10_000
    model, args = synthetic.make_backbone(seed=0, use_image=True)
    model.cache_luts(320, 240, args.radius)
    torch.manual_seed(1)
    ead = EventADModel(model).cuda()
    ead.emulate_feature_cache = False
    d = synthetic.make_data(events, B, seed=44, dist="U", device="cuda")
    for k, v in synthetic.make_boxes(B, seed=66).items():
        setattr(d, k, v.cuda())
    d.image = torch.rand((B, 3, 240, 320), generator=torch.Generator().manual_seed(3)).cuda()
    labels = d.bbox[:, 4]
    with torch.backends.cudnn.flags(enabled=True, allow_tf32=False):
        runner = GraphedEventAD(ead, B, d.pos.shape[0])
        assert_close(runner.features(d), ead.extract_features(d), what="graphed box features")
        ead.train()                                      # cuDNN's RNN backward needs training mode ...
        ead.gru_net_event.gru.dropout = 0.0              # ... without its dropouts (GRU layers, fusion MLP) the steps compare
        ead.fusion_module.fusion[2].p = 0.0
        ead2 = copy.deepcopy(ead)
        o1, o2 = train.make_optimizer(ead), train.make_optimizer(ead2)
        l1 = train.train_step(ead, d, labels, o1)
        l2 = train.train_step(ead2, d, labels, o2, feature_fn=runner.features)
        assert abs(float(l1) - float(l2)) <= 1e-4 * abs(float(l1)) + 1e-5
        for p, q in zip(train.head_parameters(ead), train.head_parameters(ead2)):
            assert_close(q, p, rtol=1e-3, atol=1e-5, what="head parameter after one step")


def test_pipelined_full_model_equals_eager_forward():
    """engine.PipelinedEventAD (two graphed runners on two streams, results one batch behind): five batches, each equal to
    the eager forward of that batch."""
    from eventad_b200 import synthetic
    from eventad_b200.engine import PipelinedEventAD
    from eventad_b200.models.EventAD import EventADModel
    B, events = 2, 12_000
    model, args = synthetic.make_backbone(seed=0, use_image=True)
    model.cache_luts(320, 240, args.radius)
    torch.manual_seed(1)
    ead = EventADModel(model).cuda().eval()
    ead.emulate_feature_cache = False
    batches = []
    for s in range(5):
        d = synthetic.make_data(events, B, seed=70 + s, dist="U", device="cuda")
        for k, v in synthetic.make_boxes(B, seed=80 + s, drop=0.3 if s % 2 else 0.0).items():
            setattr(d, k, v.cuda())
        d.image = torch.rand((B, 3, 240, 320), generator=torch.Generator().manual_seed(10 + s)).cuda()
        batches.append(d)
    with torch.no_grad(), torch.backends.cudnn.flags(enabled=True, allow_tf32=False):
        want = [ead(d, d.bbox[:, 4], testing=True) for d in batches]
        pipe = PipelinedEventAD(ead, B, batches[0].pos.shape[0])
        got = [pipe.submit(d, d.bbox[:, 4]) for d in batches]
        assert got[0] is None
        got = got[1:] + [pipe.drain()]
    assert pipe.drain() is None
    for (l0, o0, t0), (l1, o1, t1) in zip(want, got):
        assert [len(a) for a in o0] == [len(a) for a in o1] and len(o0) > 0
        for a, b in zip(o0, o1):
            for x, y in zip(a, b):
                assert_close(y, x, what="pipelined logits")
        assert [[int(v) for v in a] for a in t0] == [[int(v) for v in a] for a in t1]
