"""Generates tests/golden/ref_graph_*.npz from the UNMODIFIED reference CUDA kernels (oracle/_ref/ev_graph_cuda.so =
/root/reference/src/dagr/graph/ev_graph.cu compiled as-is by oracle/build_ref.py).  Needs a GPU:
    gpurun -- 'python tests/golden/make_ref_graph_golden.py && cp tests/golden/*.npz gpurun_out/'
The fixtures pin the CPU oracle (tests/test_oracle_cpu.py, no GPU) and the sm_100a kernels (tests/test_golden_gpu.py)
to outputs of the reference itself."""
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
from eventad_b200 import synthetic  # noqa: E402
from tests.test_ref_kernel_gpu import RefGraph, _ref_module  # noqa: E402

CASES = [  # name, events per window, B, dist, W, H, Q, K, r, dt, micro-batch cuts (None = one reset-mode call)
    ("u3000_b1", 3000, 1, "U", 64, 48, 128, 16, 4, 10000, None),
    ("c4000_b2_q8", 2000, 2, "C", 48, 36, 8, 16, 4, 10000, None),
    ("c2500_k6_r2", 2500, 1, "C", 40, 30, 16, 6, 2, 3000, None),
    ("stream_c3000", 3000, 1, "C", 64, 48, 16, 16, 4, 10000, [0, 1500, 2500, 2990, 2991, 2992, 3000]),
]


def main():
    m = _ref_module()
    out_dir = os.path.dirname(os.path.abspath(__file__))
    for name, n, B, dist, W, H, Q, K, r, dt, cuts in CASES:
        raw = synthetic.make_batch(n, B, seed=11, dist=dist, width=W, height=H)
        pos = np.concatenate([raw["pos"].astype(np.int32), raw["t"].reshape(-1, 1)], 1).astype(np.int32)
        batch = raw["batch"].astype(np.int32)
        g = RefGraph(m, W, H, B, K, Q, r, dt)
        cuts_ = cuts or [0, len(batch)]
        edges = []
        for a, b in zip(cuts_[:-1], cuts_[1:]):
            e = g.forward(torch.from_numpy(batch[a:b]).cuda(), torch.from_numpy(pos[a:b]).cuda())
            edges.append(e.cpu().numpy())
        np.savez_compressed(os.path.join(out_dir, f"ref_graph_{name}.npz"), pos=pos, batch=batch,
                            params=np.array([W, H, B, Q, K, r, dt], np.int64), cuts=np.array(cuts_, np.int64),
                            edge_counts=np.array([e.shape[1] for e in edges], np.int64),
                            edges=np.concatenate(edges, 1).astype(np.int32))
        print(name, [e.shape for e in edges])


if __name__ == "__main__":
    main()
