"""Diagnostics: timeline (us, %globaltimer, CTA 0) of the one-launch policy step in the host-in-the-loop mode
(rollout.RolloutGraph(host_io=True): observations read from pinned host memory, results stored to it), next to the
host's wall time for the same call.
    python profiles/e2e_trace.py"""
import ctypes as C
import os
import sys
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch

import bench
from options_in_simulation_b200 import _lib as L
from options_in_simulation_b200.rollout import RolloutGraph

lib = L.lib()
world = bench.World(bench.WORKLOADS["kamma_4d"], 4096, "bf16", torch.device("cuda", 0), seed=1)
for host_io in (True, False):
    g = RolloutGraph(world.env, world.agent, world.t, world.T, host_io=host_io)
    if host_io:
        g.h_states.copy_(world.obs)
    for _ in range(50):
        g.step()
    torch.cuda.synchronize()
    rows = []
    for it in range(20):
        lib.ois_debug_tct_trace_reset()
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        g.step()
        if not host_io:
            torch.cuda.synchronize()
        wall = 1e6 * (time.perf_counter() - t0)
        torch.cuda.synchronize()
        buf = (C.c_ulonglong * 64)()
        lib.ois_debug_tct_trace(buf, 64)
        b0 = buf[0]
        rows.append((wall, [(buf[k] - b0) / 1e3 for k in (1, 2, 8, 23, 29, 31, 32, 33, 30, 38, 39, 40, 41, 61)], (buf[61] - buf[60]) / 1e3))
    rows.sort(key=lambda r: r[0])
    wall, st, span = rows[len(rows) // 2]
    names = ("set-up done", "X written", "L2 issued", "actions stored", "actions seen", "sub-steps begin", "sub-steps end",
             "final stores", "envs stepped", "CTA 0 fence done", "CTA 0 ticket taken", "last CTA: ticket taken", "last CTA: block advanced / host flag written",
             "last warp end")
    print(f"host_io={host_io}: host wall {wall:.1f} us (median of 20; device mode includes a device synchronise); launch span {span:.2f} us; CTA 0: "
          + ", ".join(f"{n} {v:.2f}" for n, v in zip(names, st)))
    world.t = g.timestep
