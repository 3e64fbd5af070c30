"""Diagnostics: per-phase timeline (us, %globaltimer) of CTA 0 of the transposed tcgen05 actor kernel (csrc/mlp_tct.cu)
for one launch, warm and cold L2, plus event-timed launch durations of both actor kernels.
Run on the GPU box:  python profiles/tct_trace.py [n_envs]"""
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
agent = A.DDPG(models, None, osp, asp, "cuda:0", {"mixed_precision": True, "batch_size": 512, "state_preprocessor": A.RunningStandardScaler,
                                                  "state_preprocessor_kwargs": {"size": osp, "device": "cuda:0"},
                                                  "exploration": {"noise": A.OrnsteinUhlenbeckNoise(0.15, 0.1, 0.05, device="cuda:0"), "timesteps": 5e5},
                                                  "experiment": {"write_interval": 0, "checkpoint_interval": 0}})
agent.init({"timesteps": 1000})
x = torch.randn(n, 6, device="cuda")
flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")
names = {0: "start", 1: "setup done", 2: "aux: X written", 3: "MMA: x_ready", **{4 + k: f"MMA: h1_ready[{k}]" for k in range(4)},
         8: "MMA: L2 all issued", 9: "MMA: L3 all issued", 10: "MMA: L4 issued", **{12 + g: f"epi: H1 block {g} written" for g in range(4)},
         16: "epi: d2_full seen", **{17 + g: f"epi: H2 block {g} written" for g in range(4)}, 21: "epi: d3_full seen",
         22: "epi: d4_full seen", 23: "done (actions stored)", **{52 + k: f"producer: stage {12 + k} requested" for k in range(8)},
         **{45 + k: f"MMA: L3 tile {k} in shared memory" for k in range(4)}}
lib = ois.lib()
lib.ois_debug_set_tct_max_rows.argtypes = [C.c_longlong]
for mode in ("warm", "cold"):
    for it in range(3):
        if mode == "cold":
            flush.zero_()
        torch.cuda.synchronize()
        lib.ois_debug_tct_trace_reset()
        agent.act(x, 100, 1000)
        torch.cuda.synchronize()
    buf = (C.c_ulonglong * 64)()
    assert lib.ois_debug_tct_trace(buf, 64) == 0
    t0 = buf[0]
    print(f"--- {mode} L2, n={n}")
    for k in sorted(names, key=lambda k: buf[k]):
        print(f"{(buf[k] - t0) / 1e3:8.2f} us  {names[k]}")
    if buf[51]:
        print(f"debug 4: 128 layer-3 MMAs re-issued with all other warps parked: {(buf[51] - buf[50]) / 1e3:.2f} us = {(buf[51] - buf[50]) * 1.965 / 128:.0f} cycles per MMA at 1965 MHz")
    print(f"all CTAs: first start {(buf[60] - t0) / 1e3:.2f} us, last start {(buf[62] - t0) / 1e3:.2f} us, last end {(buf[61] - t0) / 1e3:.2f} us "
          f"(launch span {(buf[61] - buf[60]) / 1e3:.2f} us)")
for label, max_rows in (("transposed 64-env tiles (mlp_tct.cu)", 1 << 30), ("128-env row tiles (mlp_tc.cu)", 0)):
    lib.ois_debug_set_tct_max_rows(max_rows)
    for cold in (False, True):
        ts = []
        for it in range(30):
            if cold:
                flush.zero_()
            a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            a.record()
            agent.act(x, 100, 1000)
            b.record()
            torch.cuda.synchronize()
            ts.append(a.elapsed_time(b) * 1e3)
        ts.sort()
        print(f"{label}: n={n} {'cold' if cold else 'warm'} L2: median {ts[len(ts) // 2]:.2f} us, min {ts[0]:.2f} us per launch (events)")
lib.ois_debug_set_tct_max_rows(-1)

# the same kernel inside the bench's rollout step (bench.World: agent.act -> env.step with the fused replay append)
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import bench

w = bench.World(bench.WORKLOADS["kamma_4d"], n, "bf16", torch.device("cuda", 0), seed=1)
for fused in (True, False):
  step = w.step_fused if fused else w.step
  print("== rollout.fused_step (actor + env step: one kernel)" if fused else "== agent.act + env.step (two kernels; env stamps need make EXTRA=-DOIS_ENV_TRACE)")
  for cold in (False, True):
      for it in range(5):
          torch.cuda.synchronize()
          lib.ois_debug_tct_trace_reset()
          lib.ois_debug_env_trace(None, 1)
          flush.zero_() if cold else flush[: 64 << 20].zero_()     # as in bench.py: the host queues both launches behind it
          step()
          torch.cuda.synchronize()
      buf = (C.c_ulonglong * 64)()
      lib.ois_debug_tct_trace(buf, 64)
      t0 = buf[0]
      eb = (C.c_ulonglong * 4)()
      lib.ois_debug_env_trace(eb, 0)
      print(f"env kernel of the same step (us after the actor's start): first CTA start {(eb[0] - t0) / 1e3:.2f}, griddepcontrol.wait "
            f"returned {(eb[1] - t0) / 1e3:.2f} .. {(eb[2] - t0) / 1e3:.2f}, last CTA end {(eb[3] - t0) / 1e3:.2f}")
      print(f"bench.World step, {'cold' if cold else 'warm'} L2: first start {(buf[60] - t0) / 1e3:.2f} us, last start {(buf[62] - t0) / 1e3:.2f} us, "
            f"last end {(buf[61] - t0) / 1e3:.2f} us (span {(buf[61] - buf[60]) / 1e3:.2f} us); X written {(buf[2] - t0) / 1e3:.2f}, "
            f"L2 issued {(buf[8] - t0) / 1e3:.2f}, done {(buf[23] - t0) / 1e3:.2f}; set-up done {(buf[1] - t0) / 1e3:.2f}, aux before / after "
            f"griddepcontrol.wait {(buf[24] - t0) / 1e3:.2f} / {(buf[25] - t0) / 1e3:.2f}; env part: loads issued {(buf[28] - t0) / 1e3:.2f}, "
            f"actions seen {(buf[29] - t0) / 1e3:.2f}, envs stepped {(buf[30] - t0) / 1e3:.2f} "
            f"[aux after the barrier: enters its loop {(buf[42] - t0) / 1e3:.2f}, copies landed {(buf[43] - t0) / 1e3:.2f}, X computed {(buf[44] - t0) / 1e3:.2f}] [at the set-up barrier: warp 0 {(buf[34] - t0) / 1e3:.2f}, MMA warp {(buf[35] - t0) / 1e3:.2f} (before tmem alloc {(buf[37] - t0) / 1e3:.2f}), aux warp {(buf[36] - t0) / 1e3:.2f}] "
            f"(sub-steps begin {(buf[31] - t0) / 1e3:.2f}, end {(buf[32] - t0) / 1e3:.2f}, final stores {(buf[33] - t0) / 1e3:.2f})")
      ev = [torch.cuda.Event(enable_timing=True) for _ in range(3)]
      if cold:
          flush.zero_()
      ev[0].record()
      actions = w.agent.act(w.obs, w.t, w.T)[0]
      ev[1].record()
      w.obs = w.env.step(actions)[0]
      ev[2].record()
      torch.cuda.synchronize()
      print(f"   events: act {ev[0].elapsed_time(ev[1]) * 1e3:.2f} us, env {ev[1].elapsed_time(ev[2]) * 1e3:.2f} us")
