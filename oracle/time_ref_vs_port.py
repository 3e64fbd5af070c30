"""TEST INFRASTRUCTURE ONLY — how much slower is the reference's OWN env code than the CPU port bench.py times?

Runs the UNMODIFIED reference task (src/my_envs/KAMMA_medium.py through oracle/ref_harness: its ~40 `.item()` logging
syncs per step, the per-env reset loop of :491-496, every torch op of pre_/post_physics_step) and oracle/env_oracle.py
on the same state, the same actions and the same physics stand-in (the float64 skid-steer twin), and prints the ratio
BASELINE.md section 3 quotes.  Needs /root/reference (build container only):
    PYTHONPATH=/root/reference/src python oracle/time_ref_vs_port.py [n_envs] [steps]
"""
import os
import sys
import time

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, "/root/reference/src")

from oracle.env_oracle import VARIANTS, EnvOracle, EnvParams  # noqa: E402
from oracle.gen_golden import action_sample, load_state_into_reference, synth_state  # noqa: E402
from oracle.ref_harness.ref_env import RefRunner  # noqa: E402

n = int(sys.argv[1]) if len(sys.argv) > 1 else 4096
steps = int(sys.argv[2]) if len(sys.argv) > 2 else 20
F = np.float32
name = "KAMMA_medium"
variant = VARIANTS[name]
rng = np.random.default_rng(0)
runner = RefRunner(name, n, physics=True, init_uniforms=rng.random((n, 8)).astype(F))
s0 = synth_state(variant, n, rng, physics=True)
load_state_into_reference(runner, variant, s0)
oracle = EnvOracle(variant, s0.copy(), EnvParams(physics=True))
draws = [(action_sample(variant, n, rng), rng.random(n).astype(F), rng.random((n, 8)).astype(F)) for _ in range(steps + 2)]
for a, rm, ru in draws[:2]:
    runner.step(a, rand_mask=rm, reset_uniforms=ru)
    oracle.step(a, rm, ru)
t0 = time.perf_counter()
for a, rm, ru in draws[2:]:
    runner.step(a, rand_mask=rm, reset_uniforms=ru)
t_ref = (time.perf_counter() - t0) / steps
t0 = time.perf_counter()
for a, rm, ru in draws[2:]:
    oracle.step(a, rm, ru)
t_port = (time.perf_counter() - t0) / steps
import torch  # noqa: E402

print(f"{name} n={n}, {steps} steps, {torch.get_num_threads()} torch threads: reference's own env step {1e3 * t_ref:.1f} ms, "
      f"CPU port {1e3 * t_port:.1f} ms -> the port is {t_ref / t_port:.2f}x faster than the reference's own code")
