import numpy as np, torch, time
from eventad_b200 import synthetic
from eventad_b200.graph.ev_graph import SlidingWindowGraph
B=64
raw = synthetic.make_batch(100000, B, seed=0, dist="U")
pos = np.concatenate([raw["pos"].astype(np.int32), raw["t"].reshape(-1,1)],1).astype(np.int32)
bt, pt = torch.from_numpy(raw["batch"].astype(np.int32)).cuda(), torch.from_numpy(pos).cuda()
g = SlidingWindowGraph(width=raw["width"], height=raw["height"], batch_size=B, max_num_neighbors=16, max_queue_size=128, radius=4, delta_t_us=10000)
for it in range(3):
    g.reset()
    torch.cuda.synchronize(); t0=time.time()
    rows = g.build_rows(bt, pt, want_cls=True)
    torch.cuda.synchronize(); print("build_rows ms", (time.time()-t0)*1e3)
print(int(rows["count"].sum()))
