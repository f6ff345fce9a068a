"""cfg5 ray-cast microbench alone (for ncu captures of k_trace_batch): python tools/raycast_only.py [steps]"""
import json
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch  # noqa: E402

import bench  # noqa: E402

dev = torch.device("cuda", 0)
torch.cuda.set_device(dev)
stream = torch.cuda.Stream(dev)
torch.cuda.set_stream(stream)
teapot, scene, env, settings = bench.load_workload()
peak, _ = bench.measured_peak_gbs()
print(json.dumps(bench.run_raycast(dev, 0, 1, teapot, peak, int(sys.argv[1]) if len(sys.argv) > 1 else 3, 3, check=False)))
