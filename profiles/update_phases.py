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


timed("empty_graph(add_)", lambda: ag._step_dev.add_(0))
timed("gather", lambda: ag.memory.sample_by_index(ag._tensors_names, idx))
for ph in range(3):
    timed(f"phase{ph}", lambda ph=ph: L.check(lib.ois_ddpg_update_phase(ag._handle, ag._scaler_ref(), C.byref(io), ph, 1.0, L.current_stream())))
timed("whole", lambda: [L.check(lib.ois_ddpg_update_phase(ag._handle, ag._scaler_ref(), C.byref(io), ph, 1.0, L.current_stream())) for ph in range(3)])
print(json.dumps({"precision": args.precision, "us": out}))
