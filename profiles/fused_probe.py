"""Runs only the one-launch policy step (rollout.fused_step, 4096 envs, bf16) so that ncu can be pointed at it:
    ncu --set full --clock-control none --import-source on -k regex:tct_actor_kernel -s 20 -c 2 -o out python profiles/fused_probe.py"""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch

import bench

n = int(sys.argv[1]) if len(sys.argv) > 1 else 4096
world = bench.World(bench.WORKLOADS["kamma_4d"], n, "bf16", torch.device("cuda", 0), seed=1)
for _ in range(40):
    world.step_fused()
torch.cuda.synchronize()
print("fused launches:", __import__("options_in_simulation_b200").lib().ois_rollout_fused_launches())
