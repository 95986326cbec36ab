"""B = 6 full-model throughput: GraphedEventAD with the two branches one after the other vs as two arms of one graph."""
import torch
from eventad_b200 import synthetic
from eventad_b200.engine import GraphedEventAD, PipelinedEventAD
from eventad_b200.models.EventAD import EventADModel
dev = torch.device("cuda"); B = 6
model, args = synthetic.make_backbone(seed=0, use_image=True)
model.cache_luts(320, 240, args.radius)
torch.manual_seed(1)
ead = EventADModel(model).to(dev).eval()
sets = []
for s in range(3):
    d = synthetic.make_data(100_000, B, seed=500 + s, dist="U", device=dev)
    for k, v in synthetic.make_boxes(B, seed=600 + s).items(): setattr(d, k, v.to(dev))
    d.image = torch.rand((B, 3, 240, 320), device=dev).contiguous(memory_format=torch.channels_last)
    sets.append(d)
n = int(sets[0].pos.shape[0])
def timed(fn, nb=17):
    with torch.no_grad():
        fn(4); torch.cuda.synchronize()
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record(); fn(nb); b.record(); torch.cuda.synchronize()
    return a.elapsed_time(b) / nb
out = {}
for name, ov in (("sequential", False), ("overlap", True)):
    r = GraphedEventAD(ead, B, n, device=dev, overlap=ov)
    out[name] = timed(lambda k: [r(sets[i % 3], sets[i % 3].bbox[:, 4]) for i in range(k)])
    p = PipelinedEventAD(ead, B, n, device=dev, overlap=ov)
    def run_p(k):
        for i in range(k): p.submit(sets[i % 3], sets[i % 3].bbox[:, 4])
        p.drain()
    out[name + "_pipelined"] = timed(run_p)
print(out)
