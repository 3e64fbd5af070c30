"""Diagnostics: the fused env kernel alone (fused replay append on), CUDA events, inputs larger than L2 at 1M+ envs.
    python profiles/env_bench.py [variant] [n1,n2,...]"""
import json
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch

import options_in_simulation_b200 as ois
from options_in_simulation_b200.config import script_cfg

variant = sys.argv[1] if len(sys.argv) > 1 else "KAMMA_medium"
peak = json.load(open(os.path.join(os.path.dirname(__file__), "..", "MEASURED_PEAKS.json")))["hbm_gbs"] if os.path.exists(
    os.path.join(os.path.dirname(__file__), "..", "MEASURED_PEAKS.json")) else 6650.0
bytes_per = {"KAMMA_medium": 265, "KA_DDPG_2D": 253, "KA_DDPG_1D": 249, "KAMMA_eval": 289}[variant] + 70
flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")
sizes = [int(x) for x in sys.argv[2].split(",")] if len(sys.argv) > 2 else (4096, 65536, 1 << 20, 1 << 22)
for n in sizes:
    task = ois.TorqueDistributionTask(sim_config=script_cfg(n), variant=variant, seed=1)
    mem = ois.RandomMemory(memory_size=2, num_envs=n, device="cuda:0")
    for name, size, dt in (("states", 6, torch.float32), ("next_states", 6, torch.float32), ("actions", 4, torch.float32),
                           ("rewards", 1, torch.float32), ("terminated", 1, torch.bool), ("truncated", 1, torch.bool)):
        mem.create_tensor(name, size, dt)
    task.bind_replay(mem)
    task.post_reset()
    a = torch.rand(n, task.num_actions, device="cuda")
    for _ in range(5):
        task.step(a)
    ev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(20)]
    torch.cuda.synchronize()
    for s, e in ev:
        if n <= 65536:
            flush.zero_()
        s.record()
        task.step(a)
        e.record()
    torch.cuda.synchronize()
    t = sorted(s.elapsed_time(e) for s, e in ev)
    med = t[len(t) // 2]
    print(f"{variant} n={n:8d} median {med * 1e3:8.1f} us  min {t[0] * 1e3:8.1f} us  -> {bytes_per * n / med / 1e6:7.1f} GB/s "
          f"= {bytes_per * n / med / 1e6 / peak:.3f} of measured HBM peak; {n / med / 1e6:.2f} G env-steps/s", flush=True)
    del task, mem
    torch.cuda.empty_cache()
