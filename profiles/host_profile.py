"""Diagnostics: where the HOST time of one call-by-call policy step goes (cProfile over 3000 steps, no sync inside).
    python profiles/host_profile.py"""
import cProfile
import os
import pstats
import sys
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch

import bench

wl = bench.WORKLOADS["kamma_4d"]
world = bench.World(wl, 4096, "bf16", torch.device("cuda", 0), seed=1)
for _ in range(50):
    world.step()
torch.cuda.synchronize()
t0 = time.perf_counter()
for _ in range(300):
    world.step()
dt = time.perf_counter() - t0
torch.cuda.synchronize()
print(f"host issue per step: {1e6 * dt / 300:.1f} us")
pr = cProfile.Profile()
pr.enable()
for _ in range(3000):
    world.step()
    if world.t % 250 == 0:
        torch.cuda.synchronize()
pr.disable()
pstats.Stats(pr).sort_stats("tottime").print_stats(22)
