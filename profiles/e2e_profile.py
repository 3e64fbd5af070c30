"""Diagnostics: where one host-in-the-loop step (rollout.RolloutGraph(host_io=True).step()) spends its wall time:
the Python around the one C call, cudaGraphLaunch, and the spin on the completion word.
    python profiles/e2e_profile.py"""
import ctypes as C
import os
import sys
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch

import bench
from options_in_simulation_b200 import _lib as L
from options_in_simulation_b200.rollout import RolloutGraph

wl = bench.WORKLOADS["kamma_4d"]
world = bench.World(wl, 4096, "bf16", torch.device("cuda", 0), seed=1)
g = RolloutGraph(world.env, world.agent, world.t, world.T, host_io=True)
g.h_states.copy_(world.obs)
for _ in range(50):
    g.step()
N = 2000
t0 = time.perf_counter()
for _ in range(N):
    g.step()
whole = (time.perf_counter() - t0) / N
# the same C call alone, from a tight loop (no RolloutGraph bookkeeping): launch + spin
lib = L.lib()
graph, raw = g._graphs[g._cur if g._zero_copy else 0]
stream = L.current_stream(g._dev_index)
done = g._h_done.data_ptr()
src = g._cur
t_c = []
for _ in range(N):
    graph, raw = g._graphs[g._cur]
    t1 = time.perf_counter()
    L.check(lib.ois_graph_launch_wait(raw, stream, done, (g._steps_done + 1) & 0xFFFFFFFF, g._wait_timeout_ns))
    t_c.append(time.perf_counter() - t1)
    g._advance_host()
    g._steps_done += 1
    g._cur = 1 - g._cur
t_c.sort()
# launch only (no wait), then a device sync: the cost of cudaGraphLaunch itself
t_l = []
for _ in range(200):
    graph, raw = g._graphs[g._cur]
    t1 = time.perf_counter()
    L.check(lib.ois_graph_launch(raw, stream))
    t_l.append(time.perf_counter() - t1)
    torch.cuda.synchronize()
    g._advance_host()
    g._steps_done += 1
    g._cur = 1 - g._cur
t_l.sort()
print(f"g.step() wall {1e6 * whole:.1f} us | the C call alone (launch + spin) median {1e6 * t_c[N // 2]:.1f} us, min {1e6 * t_c[0]:.1f} | "
      f"cudaGraphLaunch alone median {1e6 * t_l[100]:.1f} us -> Python around the call {1e6 * (whole - t_c[N // 2]):.1f} us")
