"""Streaming-latency leg of bench.py on its own (BASELINE configs[4]).  Usage: python tools/bench_stream.py [U|C]"""
import json
import os
import sys
from types import SimpleNamespace

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import bench  # noqa: E402

a = SimpleNamespace(dist=sys.argv[1] if len(sys.argv) > 1 else "U")
print(json.dumps(bench.leg_streaming(a, torch.device("cuda"))))
