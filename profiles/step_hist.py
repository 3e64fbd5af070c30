"""Diagnostics: distribution of the per-step event-pair readings of the one-launch policy step (4096 envs, L2 flushed
between steps, as bench.py times it).    python profiles/step_hist.py"""
import collections
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch

import bench

world = bench.World(bench.WORKLOADS["kamma_4d"], 4096, "bf16", torch.device("cuda", 0), seed=1)
flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")
for fused in (True, False):
    step = world.step_fused if fused else world.step
    for _ in range(20):
        step()
    ev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(400)]
    torch.cuda.synchronize()
    for a, b in ev:
        flush.zero_()
        a.record()
        step()
        b.record()
    torch.cuda.synchronize()
    t = [a.elapsed_time(b) * 1e3 for a, b in ev]
    h = collections.Counter(round(x, 1) for x in t)
    print("fused" if fused else "two launches", f"mean {sum(t) / len(t):.2f} us; histogram (us: count):",
          ", ".join(f"{k}: {v}" for k, v in sorted(h.items())))
