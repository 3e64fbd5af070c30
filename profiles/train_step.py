"""One policy step + two DDPG updates (the reference trainer's cadence, base_KA_DDPG.py:265,283) on the bench world,
eager launches (graph_update off) so that a profiler sees every kernel.   python profiles/train_step.py [steps]
Without a profiler it prints the device-timed ms per train step with the update graphed (the product configuration)."""
import json
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch

import bench

steps = int(sys.argv[1]) if len(sys.argv) > 1 else 6
profiled = steps <= 10
wl = bench.WORKLOADS["kamma_4d"]
world = bench.World(wl, 4096, "bf16", torch.device("cuda", 0), seed=1)
if profiled:
    world.agent.cfg["graph_update"] = False
for _ in range(300 // max(1, 4096 // 512) + 3):      # fill enough transitions for a 512 batch
    world.step_fused()


def train_step():
    world.step_fused()
    world.agent._update(world.t, world.T)
    world.agent._update(world.t, world.T)


for _ in range(3):
    train_step()
a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
torch.cuda.synchronize()
a.record()
for _ in range(steps):
    train_step()
b.record()
torch.cuda.synchronize()
ms = a.elapsed_time(b) / steps
print(json.dumps({"train_ms_per_step": ms, "env_steps_per_s": 4096 * 1e3 / ms, "steps": steps, "graph_update": not profiled,
                  "cadence": "1 policy step (act + env step + replay append) + 2 updates of B = 512"}))
