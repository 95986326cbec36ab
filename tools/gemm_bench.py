"""Micro-benchmark of the two contraction kernels on the pooled-layer shapes (CUDA events, 20 iterations each)."""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from eventad_b200 import ops

shapes = [(17920, 2132), (17920, 1664), (17920, 82), (4480, 3380), (4480, 1664), (1120, 3380), (280, 3380),
          (2240, 2132), (2240, 1664), (560, 3380), (140, 3380), (35, 3380)]
for m, k in shapes:
    lda = (k + 3) // 4 * 4
    A = torch.randn(m, lda, device="cuda")
    W = torch.randn(k, 64, device="cuda") / k ** 0.5
    wt = ops.transpose_pad_weight(W)
    res = {}
    for name, fn in (("mma.sync", lambda: ops.gemm64_tc(A, W, m, None, dict(), k=k)),
                     ("tcgen05", lambda: ops.gemm64_tc5(A, wt, m, None, dict(), k=k))):
        for _ in range(3):
            fn()
        torch.cuda.synchronize()
        g = torch.cuda.CUDAGraph()                      # 20 calls in one graph: device time without host enqueue cost
        with torch.cuda.graph(g):
            for _ in range(20):
                fn()
        g.replay()
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        g.replay()
        e1.record()
        torch.cuda.synchronize()
        res[name] = e0.elapsed_time(e1) / 20 * 1e3
    gb = (m * lda + k * 64 + m * 64) * 4 / 1e9
    print(f"M={m:6d} K={k:5d}  mma.sync {res['mma.sync']:7.1f} us   tcgen05 {res['tcgen05']:7.1f} us   "
          f"(HBM floor {gb / 6.55e3 * 1e6:5.1f} us)", flush=True)
