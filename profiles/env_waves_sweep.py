"""Diagnostics: env kernel duration (20 launches back to back, as bench.py measures it where the state exceeds L2) for one
value of OIS_ENV_WAVES (read once per process):   OIS_ENV_WAVES=8 python profiles/env_waves_sweep.py 1048576"""
import json
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch

import bench

n = int(sys.argv[1]) if len(sys.argv) > 1 else 1 << 20
w = bench.World(bench.WORKLOADS["kamma_4d"], n, "bf16", torch.device("cuda", 0), seed=99)
best = min(bench.kernels_back_to_back(w)[1] for _ in range(3))
peak = json.load(open(os.path.join(os.path.dirname(__file__), "..", "MEASURED_PEAKS.json")))["hbm_gbs"]
print(json.dumps({"waves": os.environ.get("OIS_ENV_WAVES", "default"), "envs": n, "env_us": best * 1e3,
                  "hbm_frac": 335 * n / (best * 1e-3) / 1e9 / peak}))
