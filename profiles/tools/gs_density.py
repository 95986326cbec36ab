"""Graph stage (evad_graph_build_rows): per-event choice between row spans and per-cell walk (span limit 32, default) against all events walked (0)."""
import numpy as np, torch
from eventad_b200 import synthetic, _lib
from eventad_b200.graph.ev_graph import SlidingWindowGraph
for n, B, dist in [(50_000, 64, "U"), (50_000, 64, "C"), (100_000, 64, "U"), (100_000, 64, "C"), (300_000, 16, "U"), (300_000, 16, "C"),
                   (1_000_000, 6, "U"), (1_000_000, 6, "C")]:
    raw = synthetic.make_batch(n, B, seed=0, dist=dist)
    pos = np.concatenate([raw["pos"].astype(np.int32), raw["t"].reshape(-1, 1)], 1).astype(np.int32)
    bt, pt = torch.from_numpy(raw["batch"].astype(np.int32)).cuda(), torch.from_numpy(pos).cuda()
    g = SlidingWindowGraph(width=raw["width"], height=raw["height"], batch_size=B, max_num_neighbors=16, max_queue_size=128,
                           radius=4, delta_t_us=10000)
    out = []
    for pm in (32, 0):
        _lib.lib().evad_graph_set_span_limit(pm)
        ts = []
        for it in range(6):
            g.reset()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record(); rows = g.build_rows(bt, pt, want_cls=True); e1.record(); torch.cuda.synchronize()
            ts.append(e0.elapsed_time(e1))
        out.append(round(float(np.median(ts[2:])), 3))
    _lib.lib().evad_graph_set_span_limit(32)
    print(n, B, dist, "mixed / walk ms:", out, flush=True)
