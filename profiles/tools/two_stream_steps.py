"""Does the event branch gain from two steps in flight?  Two GraphedForward captures (own static buffers and scratch) on
two streams, steps alternating, against one capture on one stream; 64 x 100k events per step."""
import torch
import bench
from eventad_b200 import synthetic
from eventad_b200.engine import GraphedForward, pin_raw
dev = torch.device("cuda")
runner, args = bench.build_runner(dev)
B, N = 64, 100_000
raws = [pin_raw(synthetic.make_batch(N, B, seed=s, dist="U")) for s in range(4)]
feats = [f.to(dev).contiguous(memory_format=torch.channels_last) for f in synthetic.make_image_features(B, seed=0)]
datas = [runner.to_device(r) for r in raws]
gfs = [GraphedForward(runner.net, N * B, B, feats, device=dev, static_feats=True) for _ in range(2)]
for g in gfs:
    g.capture(datas[0], feats)
streams = [torch.cuda.Stream(device=dev) for _ in range(2)]
def run(n, two):
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for i in range(n):
        k = i % 2 if two else 0
        s = streams[k]
        s.wait_stream(torch.cuda.current_stream())
        with torch.cuda.stream(s):
            gfs[k](datas[i % 4], feats, finalize=False)
    for s in streams:
        torch.cuda.current_stream().wait_stream(s)
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / n
run(4, False); run(4, True)
print("one stream ms/step", run(20, False), "two streams ms/step", run(20, True))
