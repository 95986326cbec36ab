"""Where a batch of the full EventAD inference (BASELINE configs[1], B = 6) spends its time: wall clock with a
synchronize around each phase (so launch overheads of eager phases count)."""
import time, torch
from eventad_b200 import synthetic
from eventad_b200.models.EventAD import EventADModel
dev = torch.device("cuda")
B, events = 6, 100_000
model, args = synthetic.make_backbone(seed=0, use_image=True)
model.cache_luts(320, 240, args.radius)
torch.manual_seed(1)
ead = EventADModel(model).to(dev).eval()
ead.dagr_model.backbone.net = ead.dagr_model.backbone.net.to(memory_format=torch.channels_last)
sets = []
for s in range(3):
    d = synthetic.make_data(events, B, seed=500 + s, dist="U", device=dev)
    for k, v in synthetic.make_boxes(B, seed=600 + s).items():
        setattr(d, k, v.to(dev))
    d.image = torch.rand((B, 3, 240, 320), device=dev).contiguous(memory_format=torch.channels_last)
    sets.append(d)
bb = ead.dagr_model.backbone
def t(fn, n=12):
    ts = []
    for i in range(n):
        torch.cuda.synchronize(); t0 = time.perf_counter(); r = fn(); torch.cuda.synchronize(); ts.append((time.perf_counter() - t0) * 1e3)
    ts.sort(); return ts[len(ts) // 2], r
with torch.no_grad():
    for i in range(3): ead(sets[i], sets[i].bbox[:, 4], testing=True)
    d = sets[0]
    ms_all, _ = t(lambda: ead(sets[0], sets[0].bbox[:, 4], testing=True))
    ead.emulate_feature_cache = False
    ms_nocache, _ = t(lambda: ead(sets[0], sets[0].bbox[:, 4], testing=True))
    ms_img, (feat, outs) = t(lambda: bb.net(d.image))
    ms_ev, res = t(lambda: bb.forward_fused(d, feat))
    ms_sum, _ = t(lambda: d.x.sum().item())
    ms_pool, bf = t(lambda: ead.pool_boxes(res[1], d, d.num_graphs))
    ms_head, _ = t(lambda: ead.head(bf, d, d.bbox[:, 4]))
print(dict(forward_ms=ms_all, forward_nocache_ms=ms_nocache, image_ms=ms_img, event_fused_ms=ms_ev, x_sum_item_ms=ms_sum, pool_boxes_ms=ms_pool, head_ms=ms_head))
