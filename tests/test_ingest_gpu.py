"""Device-resident event ingest (eventad_b200/ingest.py, csrc/ingest.cu) through the C ABI against the
outputs of the reference's own function bodies (tests/golden/ref_ingest_*.npz): bit-exact (integer work; the filter's
float32 cells with float64 expressions included)."""
import importlib.util
import os

import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu
HERE = os.path.dirname(os.path.abspath(__file__))
spec = importlib.util.spec_from_file_location("make_ref_ingest_golden", os.path.join(HERE, "golden", "make_ref_ingest_golden.py"))
gold = importlib.util.module_from_spec(spec)
spec.loader.exec_module(gold)


def _dev(a):
    return torch.from_numpy(np.ascontiguousarray(a).astype(np.int64)).cuda()


@pytest.mark.parametrize("seed", [0, 1])
def test_ingest_matches_reference_golden(seed):
    from eventad_b200 import ingest
    g = np.load(os.path.join(HERE, "golden", f"ref_ingest_{seed}.npz"))
    rec = gold.recording(seed)
    ms = ingest.create_ms_to_idx(_dev(rec["t"]))
    assert np.array_equal(ms.cpu().numpy(), g["ms_to_idx"])
    ev = dict(x=_dev(rec["x"]), y=_dev(rec["y"]), t=_dev(rec["t"]), p=_dev(2 * rec["p"].astype("int8") - 1))
    cm = None
    for c, (a, b) in enumerate(((0, 35_000), (35_000, 60_000))):
        e, cm = ingest.downsample_events({k: v[a:b] for k, v in ev.items()}, 480, 640, 240, 320, cm)
        for k in "xytp":
            assert np.array_equal(e[k].cpu().numpy().astype(np.int64), g[f"ds{c}_{k}"].astype(np.int64)), (c, k)
        assert np.array_equal(cm.cpu().numpy(), g[f"ds{c}_map"]), c
    store = ingest.EventStore(x=rec["x"], y=rec["y"], p=rec["p"], t=rec["t"], t_offset=rec["t_offset"])
    assert np.array_equal(store.ms_to_idx.cpu().numpy(), g["ms_to_idx"])
    for i, kw in enumerate(gold.LOADS):
        kw = dict(kw)
        dt = kw.pop("dt")
        (xy, t, p), tq = store.load_events(rec["t_offset"] + dt, **kw)
        assert xy.dtype == torch.int16 and t.dtype == torch.int32 and p.dtype == torch.int8
        assert tuple(t.shape) == g[f"load{i}_t"].shape and tuple(p.shape) == g[f"load{i}_p"].shape
        assert np.array_equal(xy.cpu().numpy(), g[f"load{i}_xy"]), i
        assert np.array_equal(t.cpu().numpy(), g[f"load{i}_t"]) and np.array_equal(p.cpu().numpy(), g[f"load{i}_p"]), i
        assert tq == int(g[f"load{i}_tq"]), i


def test_ingest_window_feeds_the_event_branch():
    """load_events -> to_data -> format_data -> the fused event branch: the serving path end to end on device tensors."""
    from eventad_b200 import ingest, synthetic
    from eventad_b200.utils.buffers import format_data
    rec = gold.recording(0, n=200_000, width=320, height=240, span_us=300_000)
    store = ingest.EventStore(x=rec["x"], y=rec["y"], p=rec["p"], t=rec["t"], t_offset=rec["t_offset"])
    (xy, t, p), tq = store.load_events(rec["t_offset"] + 100_000, num_us=50_000, height=240, time_window=50_000)
    assert 20_000 < xy.shape[0] < 50_000 and int(t.min()) > 0 and int(t.max()) <= 50_000
    d = ingest.to_data(x=xy[:, 0], y=xy[:, 1], t=t[:, 0], p=p[:, 0])
    assert d.pos.dtype == torch.int16 and d.x.shape == (xy.shape[0], 1)
    d.batch = torch.zeros(xy.shape[0], dtype=torch.long, device="cuda")
    d.width, d.height, d.time_window = torch.tensor([320]), torch.tensor([240]), torch.tensor([50_000])
    d.num_graphs = 1
    d = format_data(d)
    model, args = synthetic.make_backbone(seed=0, use_image=True)
    model.cache_luts(320, 240, args.radius)
    net = model.cuda().backbone
    feats = [f.cuda() for f in synthetic.make_image_features(1, seed=0)]
    with torch.no_grad():
        out3, out4 = net(d, image_feat=feats)[0]
    assert out4.x.shape[1] == 64 and out4.x.shape[0] > 0 and torch.isfinite(out4.x).all()


def test_ingest_rejects_cpu_tensors():
    from eventad_b200 import ingest
    with pytest.raises(RuntimeError, match="CUDA"):
        ingest.create_ms_to_idx(torch.arange(10))
