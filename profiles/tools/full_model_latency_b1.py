"""Full EventAD inference of ONE 100k-event window (B = 1) through engine.GraphedEventAD: wall-clock p50 / p90 of a call."""
import time, numpy as np, torch
from eventad_b200 import synthetic
from eventad_b200.engine import GraphedEventAD
from eventad_b200.models.EventAD import EventADModel
dev = torch.device("cuda")
model, args = synthetic.make_backbone(seed=0, use_image=True)
model.cache_luts(320, 240, args.radius)
torch.manual_seed(1)
ead = EventADModel(model).to(dev).eval()
sets = []
for s in range(3):
    d = synthetic.make_data(100_000, 1, seed=900 + s, dist="U", device=dev)
    for k, v in synthetic.make_boxes(1, seed=950 + s).items():
        setattr(d, k, v.to(dev))
    d.image = torch.rand((1, 3, 240, 320), device=dev).contiguous(memory_format=torch.channels_last)
    sets.append(d)
r = GraphedEventAD(ead, 1, int(sets[0].pos.shape[0]), device=dev)
ro = GraphedEventAD(ead, 1, int(sets[0].pos.shape[0]), device=dev, overlap=True)
eager, graphed, over = [], [], []
with torch.no_grad():
    for i in range(5):
        r(sets[i % 3], sets[i % 3].bbox[:, 4]); ro(sets[i % 3], sets[i % 3].bbox[:, 4]); ead(sets[i % 3], sets[i % 3].bbox[:, 4], testing=True)
    ead.emulate_feature_cache = False
    for i in range(60):
        d = sets[i % 3]
        torch.cuda.synchronize(); t0 = time.perf_counter(); r(d, d.bbox[:, 4]); torch.cuda.synchronize(); graphed.append((time.perf_counter() - t0) * 1e3)
        torch.cuda.synchronize(); t0 = time.perf_counter(); ro(d, d.bbox[:, 4]); torch.cuda.synchronize(); over.append((time.perf_counter() - t0) * 1e3)
        torch.cuda.synchronize(); t0 = time.perf_counter(); ead(d, d.bbox[:, 4], testing=True); torch.cuda.synchronize(); eager.append((time.perf_counter() - t0) * 1e3)
print(dict(overlap_p50_ms=float(np.percentile(over, 50)), overlap_p90_ms=float(np.percentile(over, 90)), graphed_p50_ms=float(np.percentile(graphed, 50)), graphed_p90_ms=float(np.percentile(graphed, 90)),
           eager_p50_ms=float(np.percentile(eager, 50)), eager_p90_ms=float(np.percentile(eager, 90))))
