"""engine.GraphedEventAD per phase (B = 6, 100k events per window): CUDA-event time of each replay, wall clock of a call."""
import time, torch
from eventad_b200 import synthetic
from eventad_b200.engine import GraphedEventAD
from eventad_b200.models.EventAD import EventADModel
dev = torch.device("cuda")
B, events = 6, 100_000
model, args = synthetic.make_backbone(seed=0, use_image=True)
model.cache_luts(320, 240, args.radius)
torch.manual_seed(1)
ead = EventADModel(model).to(dev).eval()
d = synthetic.make_data(events, B, seed=500, dist="U", device=dev)
for k, v in synthetic.make_boxes(B, seed=600).items():
    setattr(d, k, v.to(dev))
d.image = torch.rand((B, 3, 240, 320), device=dev).contiguous(memory_format=torch.channels_last)
r = GraphedEventAD(ead, B, int(d.pos.shape[0]), device=dev)
with torch.no_grad():
    for i in range(3): r(d, d.bbox[:, 4])
    def ev_ms(fn, n=20):
        ts = []
        for i in range(n):
            a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            torch.cuda.synchronize(); a.record(); fn(); b.record(); torch.cuda.synchronize(); ts.append(a.elapsed_time(b))
        ts.sort(); return ts[len(ts) // 2]
    out = dict(image_replay_ms=ev_ms(r.img.graph.replay), event_replay_ms=ev_ms(r.ev.graph.replay), head_replay_ms=ev_ms(r.graph.replay),
               load_events_ms=ev_ms(lambda: r.ev.load(d, None)), load_boxes_ms=ev_ms(lambda: r._load_boxes(d, d.bbox[:, 4])),
               load_image_ms=ev_ms(lambda: r.img.image.copy_(d.image)))
    ts = []
    for i in range(20):
        torch.cuda.synchronize(); t0 = time.perf_counter(); r(d, d.bbox[:, 4]); torch.cuda.synchronize(); ts.append((time.perf_counter() - t0) * 1e3)
    ts.sort(); out["call_wall_ms"] = ts[10]
print(out)
