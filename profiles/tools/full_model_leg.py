"""bench.py's full_model leg alone (BASELINE configs[1]): eager EventADModel.forward vs engine.GraphedEventAD."""
import json, torch
import bench
res = bench.leg_full_model(torch.device("cuda"), 1, 0)
print(json.dumps({k: res[k] for k in ("ms_per_batch", "windows_per_s", "fps_reference_definition", "cuda_graphs")}))
