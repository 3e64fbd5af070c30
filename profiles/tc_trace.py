"""Diagnostics: per-phase timeline (ns, %globaltimer) of CTA 0 of the tcgen05 actor kernel for one launch.
Run on the GPU box:  python profiles/tc_trace.py [n_envs]"""
import ctypes as C
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch

import options_in_simulation_b200 as ois
from options_in_simulation_b200 import agent as A


class Box:
    def __init__(self, lo, hi, n):
        self.shape, self.low, self.high = (n,), np.full(n, lo, np.float32), np.full(n, hi, np.float32)


n = int(sys.argv[1]) if len(sys.argv) > 1 else 4096
osp, asp = Box(-1, 1, 6), Box(0, 1, 4)
models = {k: (A.DeterministicActor if "policy" in k else A.Critic)(osp, asp, "cuda:0") for k in ("policy", "target_policy", "critic", "target_critic")}
agent = A.DDPG(models, None, osp, asp, "cuda:0", {"mixed_precision": True, "state_preprocessor": A.RunningStandardScaler,
                                                  "state_preprocessor_kwargs": {"size": osp, "device": "cuda:0"},
                                                  "exploration": {"noise": A.OrnsteinUhlenbeckNoise(0.15, 0.1, 0.05, device="cuda:0"), "timesteps": 5e5},
                                                  "experiment": {"write_interval": 0, "checkpoint_interval": 0}})
agent.init({"timesteps": 1000})
x = torch.randn(n, 6, device="cuda")
flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")
names = {0: "start", 1: "setup done", 2: "X written", **{3 + q: f"L1 epi {q}" for q in range(8)}, **{11 + q: f"L2 epi {q}" for q in range(8)},
         19: "d3_full seen", 20: "act3 written", 21: "d4 seen", 22: "done",
         32: "MMA: x_ready", 33: "MMA: W1 landed", 34: "MMA: L1 issued", 35: "MMA: act1_ready", 36: "MMA: L2 p0 issued",
         37: "MMA: L2 all issued", 38: "MMA: L3 issued", 39: "MMA: L4 issued"}
for mode in ("warm", "cold"):
    for it in range(3):
        if mode == "cold":
            flush.zero_()
        agent.act(x, 100, 1000)
        torch.cuda.synchronize()
    buf = (C.c_ulonglong * 64)()
    assert ois.lib().ois_debug_tc_trace(buf, 64) == 0
    t0 = buf[0]
    print(f"--- {mode} L2, n={n}")
    for k in sorted(names, key=lambda k: buf[k]):
        print(f"{(buf[k] - t0) / 1e3:8.2f} us  {names[k]}")
    if buf[37] > buf[35] and buf[45]:
        print(f"SM clock during this launch ~ {buf[45] / (buf[37] - buf[35]):.2f} GHz (cycles of issuer A1's layer 2 / its wall time)")
    print(f"issuer A1 cycles in tile 0 (OIS_TC_DIAG): layer 2 total {buf[45]}: wait weights {buf[40]} accumulators {buf[41]} (incl. layer 1) "
          f"MMA issue {buf[43]} commits {buf[44]}")
    print(f"issuer A2: layer 2 total {buf[55]}: wait weights {buf[50]} accumulators {buf[51]} MMA issue {buf[53]} commits {buf[54]}")
    print(f"issuer B: wait act2 slices {buf[42]}; epilogue layer-2 waits: group 0 d_full {buf[46]} slot_free {buf[47]}; group 1 d_full {buf[48]} slot_free {buf[49]}")
