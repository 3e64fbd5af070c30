"""Where one DDPG update spends its GPU time: each C-ABI phase (and the replay gather) captured as its own CUDA graph
and replayed back to back (the numerical results of such replays are meaningless; the kernel sequence is the real one).
    python profiles/update_phases.py [--precision bf16|fp32]"""
import argparse
import ctypes as C
import json
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch  # noqa: E402

import bench  # noqa: E402
from options_in_simulation_b200 import _lib as L  # noqa: E402

ap = argparse.ArgumentParser()
ap.add_argument("--precision", default="bf16")
args = ap.parse_args()
w = bench.World(bench.WORKLOADS["kamma_4d"], 4096, args.precision, torch.device("cuda", 0), seed=1234)
for _ in range(5):
    w.step()
ag = w.agent
for _ in range(3):
    ag._update(w.t, w.T)
graph, idx, io, batch = ag._graph
lib = L.lib()
out = {}


def timed(name, fn, reps=50):
    g = torch.cuda.CUDAGraph()
    torch.cuda.synchronize()
    with torch.cuda.graph(g):
        fn()
    for _ in range(3):
        g.replay()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(reps):
        g.replay()
    e1.record()
    torch.cuda.synchronize()
    out[name] = round(1e3 * e0.elapsed_time(e1) / reps, 1)


timed("empty_graph(add_)", lambda: ag._sample_ctl.add_(0))


def sample_gather():
    ag.memory.sample_gather(ag._tensors_names, idx, ag._sample_ctl)
    ag.memory._draws -= 1


timed("sample+gather", sample_gather)
for ph in range(3):
    timed(f"phase{ph}", lambda ph=ph: L.check(lib.ois_ddpg_update_phase(ag._handle, ag._scaler_ref(), C.byref(io), ph, 1.0, L.current_stream())))
timed("phases 0-2", lambda: [L.check(lib.ois_ddpg_update_phase(ag._handle, ag._scaler_ref(), C.byref(io), ph, 1.0, L.current_stream())) for ph in range(3)])
timed("whole update", lambda: (sample_gather(), [L.check(lib.ois_ddpg_update_phase(ag._handle, ag._scaler_ref(), C.byref(io), ph, 1.0, L.current_stream())) for ph in range(3)]))
print(json.dumps({"precision": args.precision, "us": out}))
# every launch's share of the chain: phase p cut off after its first k launches (launch_budget, csrc/ois_common.cuh)
chain = {}
for ph in range(3):
    prev, k, shares = 0.0, 1, []
    while k <= 20:
        def cut(ph=ph, k=k):
            lib.ois_debug_launch_budget(k)
            L.check(lib.ois_ddpg_update_phase(ag._handle, ag._scaler_ref(), C.byref(io), ph, 1.0, L.current_stream()))
            lib.ois_debug_launch_budget(-1)
        timed("cut", cut)
        shares.append(round(out["cut"] - prev, 1))
        if abs(out["cut"] - out[f"phase{ph}"]) < 0.25 and k > 1 and shares[-1] < 0.3:
            shares.pop()
            break
        prev = out["cut"]
        k += 1
    chain[f"phase{ph}"] = shares
print(json.dumps({"us_added_by_each_launch": chain}))
