"""DDPG._update timing on one B200: python profiles/update_bench.py [--precision bf16|fp32] [--iters N]
Prints ms per update (CUDA events around N back-to-back updates, replay sampling included) and the host-side time of
the same loop; run under `ncu --metrics gpu__time_duration.sum` for the per-kernel list (profiles/gpu_round.sh)."""
import argparse
import json
import os
import sys
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch  # noqa: E402

import bench  # noqa: E402

ap = argparse.ArgumentParser()
ap.add_argument("--precision", default="bf16")
ap.add_argument("--iters", type=int, default=200)
ap.add_argument("--envs", type=int, default=4096)
args = ap.parse_args()
wl = bench.WORKLOADS["kamma_4d"]
w = bench.World(wl, args.envs, args.precision, torch.device("cuda", 0), seed=1234)
for _ in range(5):
    w.step()
for _ in range(5):
    w.agent._update(w.t, w.T)
a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
torch.cuda.synchronize()
l0 = w.launches()
t0 = time.perf_counter()
a.record()
for _ in range(args.iters):
    w.agent._update(w.t, w.T)
b.record()
t1 = time.perf_counter()
torch.cuda.synchronize()
ms = a.elapsed_time(b) / args.iters
flop = 6645248 * 512
print(json.dumps({"precision": args.precision, "ms_per_update": ms, "host_issue_ms_per_update": 1e3 * (t1 - t0) / args.iters,
                  "launches_per_update": (w.launches() - l0) / args.iters, "tflops": flop / (ms * 1e-3) / 1e12,
                  "graph": bool(getattr(w.agent, "_graph_update", False))}))
