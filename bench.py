#!/usr/bin/env python
"""bench.py — env-steps/s of the rollout hot path (act + dyn + reward + obs + fused replay append).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference]
    python -m torch.distributed.run --nnodes=1 --nproc-per-node N ... bench.py --gpus N ...

One "step" = one policy step of ALL envs of a rank: ``agent.act`` (scaler -> actor MLP -> OU noise ->
clamp) followed by the fused env kernel (torque blend, 10 skid-steer sub-steps, done / reward / reset /
observation, replay row append).  Workload = BASELINE.json configs[2]: KAMMA-4d nofifo random RLIL,
4096 envs per GPU; envs and replay shards are rank-local (weak scaling, no data-path collective).

Timing: CUDA events on the launching stream around every step; L2 is flushed (a 256 MiB memset, outside
the event brackets) between timed steps because the 4096-env working set (~1.4 MB) would otherwise sit in
the 126 MB L2; ms_per_step = sum of the step durations / K, max over ranks.

The JSON line also carries: ``e2e`` (same step with the observations arriving from pinned host memory and
obs / reward / reset returned to pinned host memory every step), ``roofline`` of the dominant kernel
(durations measured live with events around that kernel), ``cpu_baseline`` (the CPU port in ``oracle/``
timed on the host cores — a reported baseline, not the target), ``at_scale`` (the same step at env counts
whose state exceeds L2, where the kernels reach their bandwidth / tensor limits) and ``update`` (DDPG._update
at the reference cadence, reported separately: it is not part of the metric).

At N > 1 the line also carries the data-parallel update (BASELINE.json configs[2] "with NCCL gradient all-reduce"):
``update.ms_per_update`` with the gradients exchanged inside the Adam kernels over NVLink peer memory (csrc/comm.cuh),
the same with NCCL all-reduces captured in the update graph, and without any exchange (``collective`` = the difference),
``train_*`` at the reference cadence, and an assertion that all ranks hold bit-identical weights afterwards.

``--workload`` selects the other BASELINE.json configs (kaddpg_1d, kaddpg_2d, kamma_gv_fifo, kamma_eval_sweep); the
default, kamma_4d, is the headline.

``--impl reference`` times the CPU implementation of the same step: the reference is pure Python on Isaac
Sim + skrl and cannot travel to (or run on) the GPU box, so this arm runs the pinned CPU port
(oracle/env_oracle.py + oracle/agent_oracle.py: numpy / torch restatements, verified against the reference's
own code in the build container) with all host threads.
"""
from __future__ import annotations

import argparse
import json
import os
import subprocess
import sys
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)


def host_cores():
    """Host threads the CPU legs may use: the affinity mask, capped by the cgroup CPU quota when there is one
    (more OpenMP threads than the quota allows makes the spin-waiting workers starve each other)."""
    try:
        n = len(os.sched_getaffinity(0)) or 1
    except AttributeError:
        n = os.cpu_count() or 1
    for path in ("/sys/fs/cgroup/cpu.max", "/sys/fs/cgroup/cpu/cpu.cfs_quota_us"):
        try:
            with open(path) as f:
                parts = f.read().split()
            if path.endswith("cpu.max"):
                quota, period = parts[0], float(parts[1])
            else:
                quota = parts[0]
                with open("/sys/fs/cgroup/cpu/cpu.cfs_period_us") as f:
                    period = float(f.read().split()[0])
            if quota not in ("max", "-1") and period > 0:
                n = max(1, min(n, int(float(quota) / period)))
            break
        except (OSError, ValueError, IndexError):
            continue
    return n


# torchrun exports OMP_NUM_THREADS=1 to every rank unless the caller set it; the CPU legs (cpu_baseline, --impl
# reference) run on rank 0 alone and are meant to use all the host cores, so rank 0 undoes that default before torch
# (and its OpenMP runtime) is imported.  The GPU path does no host math and is unaffected.
if "TORCHELASTIC_RUN_ID" in os.environ and os.environ.get("RANK", "0") == "0" and os.environ.get("OMP_NUM_THREADS") == "1":
    os.environ["OMP_NUM_THREADS"] = str(host_cores())

# stdout carries exactly ONE line, the JSON: libraries that print to fd 1 (NCCL's "NCCL version ..." banner under
# NCCL_DEBUG=VERSION, for one) are sent to stderr for the whole run, and emit() writes to the saved descriptor.
_JSON_FD = None


def own_stdout():
    global _JSON_FD
    if _JSON_FD is None:
        sys.stdout.flush()
        _JSON_FD = os.dup(1)
        os.dup2(2, 1)


def emit(obj):
    os.write(1 if _JSON_FD is None else _JSON_FD, (json.dumps(obj) + "\n").encode())


METRIC = "env-steps/s (dyn+obs+reward+act)"
UNIT = "env-steps/s"
ENV_BYTES = {"KAMMA_medium": 265, "KA_DDPG_2D": 253, "KA_DDPG_1D": 249}   # SURVEY.md §8(d), per env-step
REPLAY_BYTES = 70
ACTOR_FLOP = {4: 662528, 2: 662016, 1: 661760}                             # 2 * MACs, SURVEY.md §8(d)
UPDATE_FLOP = 6645248 * 512                                                # SURVEY.md §8 row A11, B = 512
WORKLOADS = {
    "kamma_4d": dict(variant="KAMMA_medium", act_dim=4, envs=4096, ou=0.05, T_e=5e5, memory_total=1_000_000, memory="random",
                     name="KAMMA-4d nofifo random RLIL (BASELINE.json configs[2])"),
    "kaddpg_2d": dict(variant="KA_DDPG_2D", act_dim=2, envs=4096, ou=0.3, T_e=7e5, memory_total=35_000, memory="fifo",
                      name="KA-DDPG-2d nofifo random RLIL (BASELINE.json configs[1])"),
    # train/train_kaddpg_1d_nofifo_random_RLIL.py:209,269 (FIFOMemory(35_000), OU base scale 0.5), 64 envs
    "kaddpg_1d": dict(variant="KA_DDPG_1D", act_dim=1, envs=64, ou=0.5, T_e=7e5, memory_total=35_000, memory="fifo",
                      name="KA-DDPG-1d nofifo random RLIL (BASELINE.json configs[0])"),
    # train/train_kamma_4d_fifo_gv_RLIL.py:208 FIFOMemory(10_000) + the gv command curriculum of KAMMA_curr1.py:577-599.
    # BASELINE.json calls this the "observation-history FIFO path"; the reference has no observation history (obs dim is
    # 6 in every variant) — its FIFO is the replay ring, which is what runs here
    "kamma_gv_fifo": dict(variant="KAMMA_curr1", act_dim=4, envs=16384, ou=0.05, T_e=5e5, memory_total=10_000, memory="fifo",
                          name="KAMMA-4d fifo gv RLIL (BASELINE.json configs[3])"),
    # eval/eval_kamma_4d.py:175-234 + KAMMA_eval.py:545-550: 65 536 envs over 8 GPUs = 8192 per GPU
    "kamma_eval_sweep": dict(variant="KAMMA_eval", act_dim=4, envs=8192, ou=None, T_e=5e5, memory_total=16_384, memory="random",
                             name="KAMMA-4d eval rollout, TE-vs-speed sweep (BASELINE.json configs[4])"),
}
ENV_BYTES.update({"KAMMA_curr1": 265, "KAMMA_eval": 265 + 24})    # eval: three more sums, SURVEY.md 8(d)


def ncu_traffic():
    """{(kernel, envs): dram bytes per launch} from the latest committed `ncu --set full` capture, read from
    profiles/ncu_traffic.json (written by profiles/ncu_summary.py next to the capture it came from).  The capture's L2
    condition is stated there: ncu replays each launch on a cold, flushed cache (--cache-control all, the default)."""
    try:
        d = json.load(open(os.path.join(ROOT, "profiles", "ncu_traffic.json")))
    except (OSError, ValueError):
        return {}, None
    return {(e["kernel"], int(e["envs"])): int(e["bytes"]) for e in d.get("entries", [])}, d.get("source")


def workload_config(wl, n, world_size):
    """The `config` object of the JSON line: identical in the GPU arm and in the reference arm."""
    return {"workload": wl["name"], "envs_per_gpu": n, "variant": wl["variant"], "act_dim": wl["act_dim"],
            "replay": "fused append, rank-local shard",
            "outputs": "obs, clamped obs, rew, reset, wheel_torqs, state, replay row; the task attributes nothing on the path reads "
                       "(pre-clip torques, guiding_reward, timeout_buf) are computed but not stored (GPU arm)", "rng": "in-kernel Philox (GPU arm) / numpy PCG64 (CPU arm)",
            "l2": "GPU arm: flushed between timed steps (256 MiB memset outside the event brackets); CPU arm: n/a",
            "parallelism": f"env/replay shards x{world_size}; rollout: no data-path collective; update: critic + actor "
                           "gradient all-reduce (ddpg.py:409-410,431-432)"}


def pin_to_cores(local_rank, local_world):
    """One slice of the affinity mask per rank (8 ranks on one host otherwise migrate and share cores with each other's
    pollers: SCALE_r01's e2e efficiency of 0.52 at N = 8 was two ranks whose host issue time doubled)."""
    try:
        cores = sorted(os.sched_getaffinity(0))
    except AttributeError:
        return None
    per = len(cores) // max(1, local_world)
    if local_world <= 1 or per < 2:
        return None
    mine = cores[local_rank * per:(local_rank + 1) * per]
    os.sched_setaffinity(0, mine)
    return mine


def parse():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=200)
    ap.add_argument("--warmup", type=int, default=20)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--workload", default="kamma_4d", choices=sorted(WORKLOADS))
    ap.add_argument("--envs", type=int, default=0, help="envs per GPU (default: the workload's 4096)")
    ap.add_argument("--precision", default="bf16", choices=["bf16", "fp32"])
    ap.add_argument("--no-extras", action="store_true", help="skip at_scale / update / cpu_baseline legs")
    ap.add_argument("--sweep-steps", type=int, default=300, help="kamma_eval_sweep: eval steps per timed sweep")
    return ap.parse_args()


# ------------------------------------------------------------------------------------------------ clocks
class ClockSampler:
    """nvidia-smi clocks / throttle reasons sampled DURING the timed region."""

    Q = ("clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index: int):
        self.index, self.rows, self.proc, self.out = index, [], None, None

    def start(self):
        # nvidia-smi writes straight into a temporary file that is parsed after the timed legs: no reader thread in
        # this process competing with the launch loop for the interpreter lock
        import tempfile

        try:
            self.out = tempfile.TemporaryFile(mode="w+")
            self.proc = subprocess.Popen(["nvidia-smi", f"--id={self.index}", f"--query-gpu={self.Q}",
                                          "--format=csv,noheader,nounits", "-lms", "50"],
                                         stdout=self.out, stderr=subprocess.DEVNULL)
        except OSError:
            self.proc = None

    def stop(self):
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.15)
        self.proc.terminate()
        try:
            self.proc.wait(timeout=5)
        except subprocess.TimeoutExpired:
            self.proc.kill()
        self.out.seek(0)
        self.rows = [[c.strip() for c in line.split(",")] for line in self.out if line.strip()]
        self.out.close()
        sm = sorted(int(r[0]) for r in self.rows if r and r[0].isdigit())
        mx = [int(r[1]) for r in self.rows if len(r) > 1 and r[1].isdigit()]
        names = ("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap")
        reasons = [n for i, n in enumerate(names) if any(len(r) > 2 + i and r[2 + i].lower().startswith("active") for r in self.rows)]
        return {"sm_mhz": sm[len(sm) // 2] if sm else None, "sm_max_mhz": max(mx) if mx else None, "reasons": reasons,
                "samples": len(sm)}


# ------------------------------------------------------------------------------------------------ CPU port
class CpuPort:
    """The same step on the host: numpy env oracle (float64 sub-steps) + CPU-torch actor (fp32)."""

    def __init__(self, wl, n, seed=1234):
        import numpy as np
        import torch

        from oracle.agent_oracle import AgentOracle
        from oracle.env_oracle import VARIANTS, EnvOracle, EnvParams
        from oracle.gen_golden import synth_state

        self.np, self.torch = np, torch
        torch.set_num_threads(host_cores())
        torch.manual_seed(seed)
        v = VARIANTS[wl["variant"]]
        self.rng = np.random.default_rng(seed)
        self.env = EnvOracle(v, synth_state(v, n, self.rng, physics=True), EnvParams(physics=True))
        self.agent = AgentOracle(6, wl["act_dim"], "sigmoid", ou=(0.15, 0.1, wl["ou"]) if wl["ou"] else None,
                                 exploration_timesteps=wl["T_e"])
        self.n, self.A = n, wl["act_dim"]
        self.obs = np.zeros((n, 6), np.float32)
        self.t = 0

    def step(self):
        np = self.np
        z = self.rng.standard_normal((self.n, self.A), dtype=np.float32)
        a = self.agent.act(self.obs, 1000 + self.t, 600000, normals=z).numpy()
        rm = self.rng.random(self.n, dtype=np.float32)
        ru = self.rng.random((self.n, 8), dtype=np.float32)
        self.obs, _, _ = self.env.step(a, rm, ru)
        self.t += 1


def time_cpu_port(wl, n, budget_s=12.0, min_steps=3, max_steps=200):
    port = CpuPort(wl, n)
    port.step()
    t0 = time.perf_counter()
    k = 0
    while k < max_steps and (k < min_steps or time.perf_counter() - t0 < budget_s):
        port.step()
        k += 1
    dt = time.perf_counter() - t0
    return n * k / dt, k, dt


def time_cpu_parts(wl, n, reps=20):
    """the CPU port timed piece by piece on the same workload — env step only, actor only, one DDPG
    update (B = 512) — so that each GPU number has its own CPU number beside it (ms per call, all host threads)."""
    import numpy as np
    import torch

    from oracle.agent_oracle import synth_transitions

    port = CpuPort(wl, n)
    port.step()
    A = wl["act_dim"]
    z = port.rng.standard_normal((n, A), dtype=np.float32)
    acts = port.agent.act(port.obs, 1000, 600000, normals=z).numpy()
    rm, ru = port.rng.random(n, dtype=np.float32), port.rng.random((n, 8), dtype=np.float32)

    def timed(fn, k):
        fn()
        t0 = time.perf_counter()
        for _ in range(k):
            fn()
        return 1e3 * (time.perf_counter() - t0) / k

    env_ms = timed(lambda: port.env.step(acts, rm, ru), reps)
    act_ms = timed(lambda: port.agent.act(port.obs, 1000, 600000, normals=z), reps)
    batch = synth_transitions(512, 6, A, torch.Generator().manual_seed(3))
    upd_ms = timed(lambda: port.agent.update(**batch), reps)
    return {"env_step_ms": env_ms, "act_ms": act_ms, "update_ms": upd_ms,
            "train_step_ms": env_ms + act_ms + 2 * upd_ms, "envs": n, "batch": 512,
            "note": "CPU port (oracle/), ms per call; train_step = env + act + 2 updates (reference cadence)"}


def run_reference(args, wl, n):
    """Reference arm: the CPU implementation on the box's host cores (rank 0 only)."""
    if int(os.environ.get("RANK", "0")) != 0:
        return
    port = CpuPort(wl, n)
    for _ in range(args.warmup):          # exactly --warmup untimed and --steps timed steps (~10 ms each at 4096 envs)
        port.step()
    steps = args.steps
    t0 = time.perf_counter()
    for _ in range(steps):
        port.step()
    dt = time.perf_counter() - t0
    v = n * steps / dt
    cores = host_cores()
    sample = f"{steps} steps x {n} envs (single host process; the CPU path does not shard across GPUs)"
    emit({
        "impl": "reference", "metric": METRIC, "value": v, "unit": UNIT, "n_gpus": args.gpus, "steps": steps,
        "warmup": args.warmup, "ms_per_step": 1e3 * dt / steps, "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "f32 (dynamics f64)", "data": "synthetic",
        "config": workload_config(wl, n, args.gpus),
        "cpu_baseline": {"value": v, "unit": UNIT, "cores": cores, "kind": "port", "sample": sample},
        "e2e": {"value": v, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "note": "reference = pure Python on Isaac Sim/PhysX + skrl: not runnable on this box; this is its pinned CPU port",
    })


# ------------------------------------------------------------------------------------------------ our arm
class World:
    """task + memory + agent of one rank, wired like train/train_kamma_4d_nofifo_random_RLIL.py:94-186."""

    def __init__(self, wl, n, precision, device, seed, model_seed=None):
        import torch

        import options_in_simulation_b200 as ois
        from options_in_simulation_b200 import agent as A
        from options_in_simulation_b200.config import script_cfg
        from options_in_simulation_b200.memory import memory_size_for

        self.torch, self.n, self.wl = torch, n, wl
        torch.manual_seed(seed if model_seed is None else model_seed)
        cfg = script_cfg(n, seed=seed, sim_device=str(device))
        self.task = ois.TorqueDistributionTask(sim_config=cfg, variant=wl["variant"], seed=seed)
        self.task.intermediate_outputs = False     # torques / guiding_reward / timeout_buf: computed, not stored (config.outputs)
        self.env = ois.wrap_env(self.task)
        dev = str(device)
        osp, asp = self.task.observation_space, self.task.action_space
        models = {"policy": A.DeterministicActor(osp, asp, dev), "target_policy": A.DeterministicActor(osp, asp, dev),
                  "critic": A.Critic(osp, asp, dev), "target_critic": A.Critic(osp, asp, dev)}
        # capacity rule: the scripts' memory_size counts transitions at numEnvs = 1; keep the total
        mem_cls = ois.FIFOMemory if wl["memory"] == "fifo" else ois.RandomMemory
        steps_cap = max(memory_size_for(wl["memory_total"], n), -(-512 // n))     # never below one batch of 512
        self.memory = mem_cls(memory_size=steps_cap, num_envs=n, device=dev, replacement=False)
        cfg_a = {"batch_size": 512, "discount_factor": 0.999, "polyak": 0.005, "actor_learning_rate": 3e-4,
                 "critic_learning_rate": 1e-3, "random_timesteps": 0, "learning_starts": 0,
                 "mixed_precision": precision == "bf16", "sampler": "device", "graph_update": True,
                 "state_preprocessor": A.RunningStandardScaler, "state_preprocessor_kwargs": {"size": osp, "device": dev},
                 "exploration": {"noise": A.OrnsteinUhlenbeckNoise(0.15, 0.1, wl["ou"], device=dev) if wl["ou"] else None,
                                 "initial_scale": 1.0, "final_scale": 1e-4, "timesteps": wl["T_e"]},
                 "experiment": {"write_interval": 0, "checkpoint_interval": 0}}
        self.agent = A.DDPG(models=models, memory=self.memory, cfg=cfg_a, observation_space=osp, action_space=asp, device=dev)
        self.agent.init({"timesteps": 600000})
        self.agent.fuse_replay_with(self.task)
        self.obs, _ = self.env.reset()
        self.t = 1000
        self.T = 600000

    def launches(self):
        return self.task.kernel_launches + self.agent.kernel_launches + self.memory.kernel_launches

    def step_fused(self):
        """The same step through rollout.fused_step: one kernel launch where the library can serve it."""
        from options_in_simulation_b200.rollout import fused_step

        self.obs, rew, term, _, _ = fused_step(self.env, self.agent, self.obs, self.t, self.T)
        self.t += 1
        return rew, term

    def step(self):
        """One policy step through the public API (what the trainer's loop body does, base_KA_DDPG.py:175-219)."""
        actions = self.agent.act(self.obs, self.t, self.T)[0]
        self.obs, rew, term, trunc, _ = self.env.step(actions)
        self.agent.record_transition(None, None, rew, self.obs, term, trunc, None, self.t, self.T)
        self.t += 1
        return rew, term


def device_timed(world, steps, warmup, flush, fused=False):
    torch = world.torch
    step = world.step_fused if fused else world.step
    for _ in range(warmup):
        step()
    ev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(steps)]
    torch.cuda.synchronize()
    l0 = world.launches()
    for a, b in ev:
        if flush is not None:
            flush.zero_()
        a.record()
        step()
        b.record()
    torch.cuda.synchronize()
    return sum(a.elapsed_time(b) for a, b in ev) / steps, world.launches() - l0


def host_issue_timed(world, steps=300):
    """Host time to ISSUE one call-by-call step (no synchronisation inside the loop; 2 launches per step, so the
    driver's launch queue never fills): when this approaches ms_per_step the loop is host-bound."""
    torch = world.torch
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    for _ in range(steps):
        world.step()
    dt = time.perf_counter() - t0
    torch.cuda.synchronize()
    return 1e3 * dt / steps


def kernel_split(world, steps, flush):
    """Per-kernel-group durations (act vs env) with events between the two launches."""
    torch = world.torch
    acts, envs = [], []
    for _ in range(steps):
        e = [torch.cuda.Event(enable_timing=True) for _ in range(3)]
        if flush is not None:
            flush.zero_()
        e[0].record()
        actions = world.agent.act(world.obs, world.t, world.T)[0]
        e[1].record()
        world.obs = world.env.step(actions)[0]
        e[2].record()
        world.t += 1
        acts.append((e[0], e[1]))
        envs.append((e[1], e[2]))
    torch.cuda.synchronize()
    f = lambda L: sorted(a.elapsed_time(b) for a, b in L)[len(L) // 2]      # median: the first iterations warm up
    return f(acts), f(envs)


def kernels_back_to_back(world, reps=20):
    """Average launch duration of each kernel with `reps` launches of the SAME kernel issued back to back inside one event
    bracket (no per-launch event pair: an event pair around a single kernel adds ~6 us of front-end latency on this
    system, profiles/micro/launch_overhead.cu).  Only used where the state exceeds L2, so no flush is needed between
    launches; the env kernel advances its own state from launch to launch, the actor re-reads the same observations."""
    torch = world.torch
    actions = world.agent.act(world.obs, world.t, world.T)[0].clone()
    for _ in range(2):
        world.env.step(actions)
    a, b, c = (torch.cuda.Event(enable_timing=True) for _ in range(3))
    torch.cuda.synchronize()
    a.record()
    for _ in range(reps):
        world.env.step(actions)
    b.record()
    for _ in range(reps):
        world.agent.act(world.obs, world.t, world.T)
    c.record()
    torch.cuda.synchronize()
    world.obs = world.task.obs_clipped
    return b.elapsed_time(c) / reps, a.elapsed_time(b) / reps


def actor_device_span(world, flush, reps=9, fused=False):
    """Duration of the transposed actor kernel on the device's own clock: %globaltimer of the first CTA's start to the
    last warp's exit (diagnostic stamps of csrc/mlp_tct.cu), median over `reps` flushed steps.  An event pair around ONE
    kernel reads ~6 us more than this (front-end latency of the two records, profiles/micro/launch_overhead.cu)."""
    import ctypes

    import options_in_simulation_b200 as ois

    lib, torch = ois.lib(), world.torch
    spans = []
    for _ in range(reps):
        torch.cuda.synchronize()
        if lib.ois_debug_tct_trace_reset() != 0:
            return None
        flush.zero_()
        world.step_fused() if fused else world.step()
        torch.cuda.synchronize()
        buf = (ctypes.c_ulonglong * 64)()
        lib.ois_debug_tct_trace(buf, 64)
        if buf[61] == 0 or buf[60] == 2 ** 64 - 1:
            return None                      # the 128-env-row-tile kernel served this size: no stamps
        spans.append((buf[61] - buf[60]) * 1e-6)
    spans.sort()
    return spans[len(spans) // 2]


def e2e_timed(world, steps, warmup):
    """Host in the loop: observations come from pinned host memory, results go back to pinned host memory."""
    torch = world.torch
    n = world.n
    dev = world.task.device
    h_obs = torch.empty(n, 6, dtype=torch.float32).pin_memory()
    h_rew = torch.empty(n, 1, dtype=torch.float32).pin_memory()
    h_done = torch.empty(n, 1, dtype=torch.bool).pin_memory()
    h_obs.copy_(world.obs)
    d_obs = torch.empty(n, 6, dtype=torch.float32, device=dev)

    def one():
        d_obs.copy_(h_obs, non_blocking=True)                      # H2D: this step's inputs
        actions = world.agent.act(d_obs, world.t, world.T)[0]
        obs, rew, term, trunc, _ = world.env.step(actions)
        world.agent.record_transition(None, None, rew, obs, term, trunc, None, world.t, world.T)
        h_obs.copy_(obs, non_blocking=True)                         # D2H: the step's results
        h_rew.copy_(rew, non_blocking=True)
        h_done.copy_(term, non_blocking=True)
        torch.cuda.synchronize()                                    # the host consumes them before the next step
        world.t += 1

    for _ in range(warmup):
        one()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    torch.cuda.synchronize()
    a.record()
    for _ in range(steps):
        one()
    b.record()
    torch.cuda.synchronize()
    return a.elapsed_time(b) / steps, n * 6 * 4, n * (6 * 4 + 4 + 1)


def e2e_graph_timed(world, steps, warmup, mode=True):
    """The same host-in-the-loop step through rollout.RolloutGraph: ONE graph launch per step.  mode=True: the actor
    kernel reads the observations from pinned host memory in place and the env kernel stores (obs | rew | reset) to
    pinned host memory in place (zero-copy over PCIe); mode="copy": an H2D copy node, the two kernels and ONE packed
    D2H copy node.  Either way the host synchronises and hands the new observations back every step."""
    from options_in_simulation_b200.rollout import RolloutGraph

    torch = world.torch
    n = world.n
    g = RolloutGraph(world.env, world.agent, world.t, world.T, host_io=mode)
    g.h_states.copy_(world.obs)

    if mode is True:
        # zero-copy: step t reads its observations from the pinned pack step t-1 wrote; step() returns when the env
        # kernel's completion word says the step's results are in host memory (no device synchronise, no host copy)
        def one():
            g.step()
    else:
        def one():
            obs, rew, term = g.step()      # synchronised: results are in pinned host memory
            g.h_states.copy_(obs)          # copy-node mode: the host hands them back (host memcpy, 96 KB at N = 4096)

    for _ in range(warmup + 3):            # +3: the eager warm-up step and the capture(s)
        one()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    a.record()
    for _ in range(steps):
        one()
    b.record()
    torch.cuda.synchronize()
    wall_ms = 1e3 * (time.perf_counter() - t0) / steps
    ms = max(a.elapsed_time(b) / steps, wall_ms) if mode is True else a.elapsed_time(b) / steps
    pipelined_ms = None
    if mode is True:                       # one step in flight while the host consumes the previous one
        g.step(wait=False)
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        for _ in range(steps):
            g.step(wait=False)
            g.wait(lag=1)
        g.wait()
        pipelined_ms = 1e3 * (time.perf_counter() - t0) / steps
    world.t = g.timestep
    world.task._obs_clipped[0].copy_(g.h_next_states)     # zero-copy mode leaves the task's device-side copy untouched
    world.obs = world.task.obs_clipped
    return ms, n * 6 * 4, n * 29, pipelined_ms


def graph_device_timed(world, steps, warmup, flush):
    """Device-resident loop through rollout.RolloutGraph (no host buffers): events around every replay."""
    from options_in_simulation_b200.rollout import RolloutGraph

    torch = world.torch
    g = RolloutGraph(world.env, world.agent, world.t, world.T, host_io=False)
    for _ in range(warmup + 2):
        g.step()
    ev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(steps)]
    torch.cuda.synchronize()
    for a, b in ev:
        if flush is not None:
            flush.zero_()
        a.record()
        g.step()
        b.record()
    torch.cuda.synchronize()
    world.t = g.timestep
    world.obs = world.task.obs_clipped
    return sum(a.elapsed_time(b) for a, b in ev) / steps, sum(g._per_step)


def update_timed(world, iters=30):
    torch = world.torch
    for _ in range(3):
        world.agent._update(world.t, world.T)
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    torch.cuda.synchronize()
    a.record()
    for _ in range(iters):
        world.agent._update(world.t, world.T)
    b.record()
    torch.cuda.synchronize()
    return a.elapsed_time(b) / iters


def train_timed(world, flush, steps=30):
    """One policy step (the one-launch step the trainer issues, rollout.fused_step) followed by two updates (the reference
    trainer's cadence), device-timed."""
    torch = world.torch
    for _ in range(3):
        world.step_fused()
        world.agent._update(world.t, world.T)
        world.agent._update(world.t, world.T)
    ev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(steps)]
    torch.cuda.synchronize()
    for a, b in ev:
        flush.zero_()
        a.record()
        world.step_fused()
        world.agent._update(world.t, world.T)
        world.agent._update(world.t, world.T)
        b.record()
    torch.cuda.synchronize()
    return sum(a.elapsed_time(b) for a, b in ev) / steps


def weights_identical_across_ranks(agent, dist, device):
    """Bit-level check that the data-parallel update left the same parameters on every rank: per-network (sum, xor-free
    integer checksum) gathered and compared.  Returns (ok, detail)."""
    import torch

    sums = []
    for name in ("policy", "critic", "target_policy", "target_critic"):
        flat = getattr(agent, name).flat
        sums.append(flat.view(torch.int32).to(torch.int64).sum())          # exact: any differing bit changes it
        sums.append((flat.view(torch.int32).to(torch.int64) * torch.arange(1, flat.numel() + 1, device=device)).sum())
    mine = torch.stack(sums)
    allr = [torch.empty_like(mine) for _ in range(dist.get_world_size())]
    dist.all_gather(allr, mine)
    ok = all(torch.equal(allr[0], x) for x in allr)
    return ok, [x.tolist() for x in allr] if not ok else None


def update_legs(world, args, flush, dist, device, world_size, rank):
    """DDPG._update (one CUDA graph per update, B = 512) and the reference-cadence train step at this world size.
    N > 1: the gradient exchange runs three ways — inside the Adam kernels over NVLink peer memory ("peer"), as NCCL
    all-reduces captured in the same graph ("nccl"), and not at all ("local", for the collective's cost by difference)."""
    torch = world.torch
    agent = world.agent

    def timed_max(fn):
        if world_size > 1:
            dist.barrier()
        torch.cuda.synchronize()
        v = fn()
        if world_size > 1:
            t = torch.tensor([v], device=device, dtype=torch.float64)
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
            v = float(t.item())
        return v

    def set_mode(mode):
        agent._dp_exchange = mode
        agent._graph = None              # the captured update contains (or not) the exchange: capture again
        torch.cuda.synchronize()

    out = {"batch": 512}
    first = agent._dp_exchange           # "peer" (or "nccl" when CUDA IPC is unavailable), "none" at N = 1
    upd_ms = timed_max(lambda: update_timed(world))
    train_ms = timed_max(lambda: train_timed(world, flush))
    out.update({"ms_per_update": upd_ms, "tflops": UPDATE_FLOP / (upd_ms * 1e-3) / 1e12,
                "path": ("bf16 tcgen05 grouped GEMMs" if args.precision == "bf16" else "fp32 SIMT") + ", one CUDA graph per update",
                "train_ms_per_step": train_ms, "train_env_steps_per_s": world_size * world.n * 1e3 / train_ms,
                "note": "DDPG._update; train_* = one policy step + 2 updates, the reference cadence "
                        "(base_KA_DDPG.py:265,283), max over ranks; not in the metric"})
    if world_size == 1:
        return out
    ok, detail = weights_identical_across_ranks(agent, dist, device)
    if not ok:
        raise SystemExit(f"data-parallel update ({first}): ranks hold different weights: {detail}")
    coll = {"exchange": first, "weights_identical_across_ranks": True, "ms_per_update": {first: upd_ms},
            "bytes_per_update_per_rank": 4 * (agent.critic.num_params + agent.policy.num_params),
            "reductions_per_update": 2, "in_update_graph": True}
    if agent._exchange is not None:
        done_c, done_a, status = agent._exchange.status()
        coll["peer_exchanges_completed"] = [done_c, done_a]
        coll["peer_barrier_timeouts"] = status
        if status:
            raise SystemExit(f"peer exchange: a barrier timed out waiting for rank {status - 1}")
    if first == "peer":
        set_mode("nccl")
        coll["ms_per_update"]["nccl"] = timed_max(lambda: update_timed(world))
        ok, detail = weights_identical_across_ranks(agent, dist, device)
        if not ok:
            raise SystemExit(f"data-parallel update (nccl): ranks hold different weights: {detail}")
    set_mode("local")                    # no exchange at all: every rank applies its own gradient (timing only; last)
    coll["ms_per_update"]["local"] = timed_max(lambda: update_timed(world))
    coll["us_per_update"] = {k: 1e3 * (v - coll["ms_per_update"]["local"]) for k, v in coll["ms_per_update"].items() if k != "local"}
    coll["overlap"] = ("none possible: the critic's reduced gradient feeds the critic step that the policy pass reads "
                       "(ddpg.py:416-423); the actor forward of ddpg.py:420 is already hoisted above the exchange")
    out["collective"] = coll
    set_mode(first)
    return out


def run_eval_sweep(args, wl, n, device, rank, world_size, dist):
    """BASELINE.json configs[4]: every rank rolls its shard of KAMMA_eval envs over the speed ramp; the per-speed-bin
    tracking-error sums are all-reduced ONCE at the end (evaluate.tracking_error_sweep)."""
    import torch

    import options_in_simulation_b200 as ois
    from options_in_simulation_b200 import parallel as P
    from options_in_simulation_b200.evaluate import phase_lock, tracking_error_sweep

    world = World(wl, n, args.precision, device, seed=1234 + rank, model_seed=77)    # ONE checkpoint on every rank
    period = phase_lock(world.task)
    steps = max(args.sweep_steps, 2 * period + 10)

    def barrier():
        if world_size > 1:
            dist.barrier()
        torch.cuda.synchronize()

    sampler = ClockSampler(device.index) if rank == 0 else None
    out = None
    for it in range(2):                  # first sweep warms up (lazy init, capture); the second is timed
        phase_lock(world.task)
        barrier()
        if it == 1 and rank == 0:
            sampler.start()
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        t0 = time.perf_counter()
        a.record()
        out = tracking_error_sweep(world.env, world.agent, steps, skip_steps=period)
        b.record()
        torch.cuda.synchronize()
        wall_ms = 1e3 * (time.perf_counter() - t0)
        dev_ms = a.elapsed_time(b)
    clocks = sampler.stop() if rank == 0 else None
    # the one collective of this workload, timed on its own: two [bins] fp64 vectors
    te = torch.zeros(10, dtype=torch.float64, device=device)
    cnt = torch.zeros(10, dtype=torch.float64, device=device)
    for _ in range(3):
        P.reduce_sums(te, cnt)
    barrier()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(20):
        P.reduce_sums(te, cnt)
    b.record()
    torch.cuda.synchronize()
    reduce_us = 1e3 * a.elapsed_time(b) / 20
    t = torch.tensor([dev_ms, wall_ms, reduce_us], device=device, dtype=torch.float64)
    if world_size > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    dev_ms, wall_ms, reduce_us = t.tolist()
    if rank != 0:
        return
    total = world_size * n * steps
    log_bytes = (steps + 1) * 24 * 4
    line = {"metric": METRIC, "value": total / (dev_ms * 1e-3), "unit": UNIT, "n_gpus": world_size, "steps": steps, "warmup": steps,
            "ms_per_step": dev_ms / steps, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
            "dtype": "bf16 (actor MLP, fp32 accumulate) + f32 (env)" if args.precision == "bf16" else "f32", "data": "synthetic",
            "config": workload_config(wl, n, world_size),
            "e2e": {"value": total / (wall_ms * 1e-3), "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": log_bytes / steps,
                    "ms_per_step": wall_ms / steps,
                    "path": "evaluate.tracking_error_sweep (the public call): device-resident rollout, episode statistics reduced "
                            "into a device log by the env kernel, ONE device->host read of the log and ONE all-reduce at the end; "
                            "wall clock around the call"},
            "gpu_launches": world.launches(), "clocks": clocks,
            "sweep": {"speed": [round(float(x), 3) for x in out["speed"]],
                      "tracking_error": [round(float(x), 5) for x in out["tracking_error"]],
                      "episodes": [int(x) for x in out["episodes"]], "episode_period_steps": period,
                      "all_reduce_us": reduce_us, "all_reduce_bytes": 2 * 10 * 8,
                      "note": "random-init policy: the curve is a parity artefact (tests/test_eval_sweep_gpu.py), not a control result"}}
    # the two kernels of one eval step on their own (events between the calls, L2 flushed), as in the headline workload
    peaks = {}
    try:
        peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
    except OSError:
        pass
    hbm_peak, tc_peak = float(peaks.get("hbm_gbs", 6650.0)), float(peaks.get("bf16_tflops", 1590.0))
    flush = torch.empty(256 << 20, dtype=torch.uint8, device=device)
    act_ms, env_ms = kernel_split(world, 50, flush)
    env_bytes = ENV_BYTES[wl["variant"]] * n            # eval: no replay append
    roofs = [{"kernel": "env_step_kernel", "bound": "hbm", "achieved": env_bytes / (env_ms * 1e-3) / 1e9, "peak": hbm_peak,
              "unit": "GB/s", "traffic": None, "ms": env_ms},
             {"kernel": "actor MLP (ois_agent_act)", "bound": "tensor", "achieved": ACTOR_FLOP[wl["act_dim"]] * n / (act_ms * 1e-3) / 1e12,
              "peak": tc_peak, "unit": "TFLOP/s", "traffic": None, "ms": act_ms}]
    for r in roofs:
        r["frac"] = r["achieved"] / r["peak"]
        r["peak_source"] = "measured (MEASURED_PEAKS.json, burst)" if peaks else "fallback (B200_PROFILING.md)"
    roofs.sort(key=lambda r: -r["ms"])
    line["roofline"], line["roofline_other"] = roofs
    if world_size == 1 and not args.no_extras:
        v, k, dt = time_cpu_port(wl, n)
        line["cpu_baseline"] = {"value": v, "unit": UNIT, "cores": host_cores(), "kind": "port",
                                "sample": f"{k} steps x {n} envs in {dt:.1f} s (numpy env oracle + CPU-torch actor, oracle/)"}
    emit(line)


def main():
    args = parse()
    own_stdout()
    wl = WORKLOADS[args.workload]
    n = args.envs or wl["envs"]
    if args.impl == "reference":
        run_reference(args, wl, n)
        return

    rank, local, world_size = int(os.environ.get("RANK", 0)), int(os.environ.get("LOCAL_RANK", 0)), int(os.environ.get("WORLD_SIZE", 1))
    pinned = pin_to_cores(local, int(os.environ.get("LOCAL_WORLD_SIZE", world_size)))     # before any thread is created

    import torch
    import torch.distributed as dist

    import options_in_simulation_b200 as ois

    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a CUDA device (B200); there is no CPU fallback for the product path")
    if not os.path.exists(ois.LIB_PATH):
        raise SystemExit(f"{ois.LIB_PATH} missing: run `python -c 'import __graft_entry__ as g; g.build()'` first")
    torch.cuda.set_device(local)
    device = torch.device("cuda", local)
    if world_size > 1:
        dist.init_process_group("nccl", device_id=device)
    if args.workload == "kamma_eval_sweep":
        run_eval_sweep(args, wl, n, device, rank, world_size, dist)
        if world_size > 1:
            dist.barrier()
            dist.destroy_process_group()
        return
    peaks = {}
    try:
        peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
    except OSError:
        pass
    hbm_peak = float(peaks.get("hbm_gbs", 6650.0))
    tc_peak = float(peaks.get("bf16_tflops", 1590.0))
    peak_src = "measured (MEASURED_PEAKS.json, burst)" if peaks else "fallback (B200_PROFILING.md)"

    world = World(wl, n, args.precision, device, seed=1234 + rank)
    flush = torch.empty(256 << 20, dtype=torch.uint8, device=device)

    def barrier():
        if world_size > 1:
            dist.barrier()
        torch.cuda.synchronize()

    sampler = ClockSampler(local) if rank == 0 else None     # one nvidia-smi poller per job, not one per rank
    barrier()
    if rank == 0:
        sampler.start()
    ms, launches = device_timed(world, args.steps, args.warmup, flush)
    barrier()
    e2e_eager_ms, _, _ = e2e_timed(world, min(args.steps, 200), 5)
    barrier()
    e2e_ms, h2d, d2h, e2e_pipe_ms = e2e_graph_timed(world, min(args.steps, 200), 5)
    barrier()
    e2e_copy_ms, _, _, _ = e2e_graph_timed(world, min(args.steps, 200), 5, mode="copy")
    barrier()
    graph_ms, graph_kernels_per_step = graph_device_timed(world, args.steps, 5, flush)
    barrier()
    fused_ms, fused_launches = device_timed(world, args.steps, 5, flush, fused=True)
    barrier()
    host_ms = host_issue_timed(world)
    barrier()
    clocks = sampler.stop() if rank == 0 else None      # sampled across ALL the timed legs above
    per_rank = [[ms, graph_ms, host_ms, e2e_ms, fused_ms]]
    if world_size > 1:
        mine = torch.tensor(per_rank[0], device=device, dtype=torch.float64)
        allr = [torch.empty_like(mine) for _ in range(world_size)]
        dist.all_gather(allr, mine)
        per_rank = [a.tolist() for a in allr]
        t = torch.tensor([ms, e2e_ms, e2e_eager_ms, graph_ms, e2e_copy_ms, e2e_pipe_ms, fused_ms], device=device, dtype=torch.float64)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        ms, e2e_ms, e2e_eager_ms, graph_ms, e2e_copy_ms, e2e_pipe_ms, fused_ms = t.tolist()
    # the same device-resident step issued two ways, both timed over EXACTLY args.steps steps, max over ranks: call by
    # call (2 launches per step from Python) and as one graph launch per step; the headline is the faster of the two
    # (with several ranks sharing a host the call-by-call loop becomes host-bound on the slowest rank)
    call_ms = ms
    issue = "call-by-call (agent.act + env.step)"
    if graph_ms < ms:
        ms, issue, launches = graph_ms, "rollout.RolloutGraph replay (one graph launch per step)", graph_kernels_per_step * args.steps
    if fused_ms < ms:
        ms, issue, launches = fused_ms, "rollout.fused_step (agent.act + env.step + record_transition as one kernel launch)", fused_launches
    value = world_size * n * 1e3 / ms
    e2e_value = world_size * n * 1e3 / e2e_ms
    e2e_sorted = sorted(r[3] for r in per_rank)

    out = {
        "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world_size, "steps": args.steps, "warmup": args.warmup,
        "ms_per_step": ms, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
        "dtype": "bf16 (actor MLP, fp32 accumulate) + f32 (env)" if args.precision == "bf16" else "f32",
        "data": "synthetic",
        "config": workload_config(wl, n, world_size),
        "issue": issue,
        "e2e": {"value": e2e_value, "unit": UNIT, "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h,
                "ms_per_step": e2e_ms,
                "path": "rollout.RolloutGraph(host_io=True): one C call per step (ois_graph_launch_wait) = CUDA-graph launch of "
                        "the actor kernel reading the observations from pinned host memory in place + the env kernel (step, "
                        "replay append, step-control advance) storing obs|rew|reset to pinned host memory in place, then a "
                        "spin on the pinned completion word the env kernel's last CTA writes; the bytes below cross PCIe "
                        "every step as zero-copy loads / stores; max(device events, host wall clock) per step, max over ranks",
                "per_rank_ms": {"median": e2e_sorted[len(e2e_sorted) // 2], "max": e2e_sorted[-1], "min": e2e_sorted[0]},
                "pipelined_ms_per_step": e2e_pipe_ms,
                "pipelined_value": world_size * n * 1e3 / e2e_pipe_ms if e2e_pipe_ms else None,
                "pipelined_note": "same path, one step in flight while the host waits for the previous one (step(wait=False) + "
                                  "wait(lag=1)); not the headline",
                "copy_nodes_ms_per_step": e2e_copy_ms,
                "call_by_call_ms_per_step": e2e_eager_ms,
                "call_by_call_value": world_size * n * 1e3 / e2e_eager_ms,
                "host_cores_per_rank": len(pinned) if pinned else None},
        "issue_paths": {"call_by_call_ms_per_step": call_ms, "graph_replay_ms_per_step": graph_ms, "fused_step_ms_per_step": fused_ms,
                        "host_issue_ms_per_step": max(r[2] for r in per_rank),
                        "per_rank_ms": [[round(x, 5) for x in r] for r in per_rank],
                        "note": "max over ranks; per_rank_ms = [call-by-call, graph replay, host issue time, e2e, fused step] per rank"},
        "gpu_launches": launches, "clocks": clocks,
    }

    traffic, traffic_src = ncu_traffic()
    if not args.no_extras:                       # every rank takes part: at N > 1 the update contains the exchange
        out["update"] = update_legs(world, args, flush, dist, device, world_size, rank)
    if rank == 0:
        act_ms, env_ms = kernel_split(world, min(args.steps, 100), flush)
        env_bytes = (ENV_BYTES[wl["variant"]] + REPLAY_BYTES) * n
        act_flop = ACTOR_FLOP[wl["act_dim"]] * n
        tr = (lambda k: traffic.get((k, n)) if args.workload == "kamma_4d" else None)
        env_roof = {"kernel": "env_step_kernel", "bound": "hbm", "achieved": env_bytes / (env_ms * 1e-3) / 1e9,
                    "peak": hbm_peak, "unit": "GB/s", "traffic": tr("env"), "ms": env_ms}
        act_roof = {"kernel": "actor MLP (ois_agent_act)", "bound": "tensor", "achieved": act_flop / (act_ms * 1e-3) / 1e12,
                    "peak": tc_peak, "unit": "TFLOP/s", "traffic": tr("actor"), "ms": act_ms}
        span = actor_device_span(world, flush)
        if span:
            act_roof["device_span_ms"] = span
            act_roof["achieved_device_span"] = act_flop / (span * 1e-3) / 1e12
            act_roof["frac_device_span"] = act_roof["achieved_device_span"] / tc_peak
            act_roof["ms_note"] = ("ms / achieved / frac: CUDA events around the one kernel (they include ~6 us of front-end latency "
                                   "of the two event records); *_device_span: first CTA start to last warp exit on %globaltimer")
        for r in (env_roof, act_roof):
            r["frac"] = r["achieved"] / r["peak"]
            r["peak_source"] = peak_src
            r["traffic_source"] = traffic_src
        dom, other = (act_roof, env_roof) if act_ms >= env_ms else (env_roof, act_roof)
        out["roofline"] = dom
        out["roofline_other"] = other
        if issue.startswith("rollout.fused_step") or graph_kernels_per_step == 1:
            # the timed step is ONE kernel (actor MLP + env step + replay append): that launch is the dominant kernel, its
            # duration between events is the step itself; the two kernels it replaces stay in the line for comparison
            fspan = actor_device_span(world, flush, fused=True)
            fused_roof = {"kernel": "tct_actor_kernel<32, kEnv> (csrc/mlp_tct.cu): actor MLP + env step + replay append, the ONE launch "
                                    "of the timed step",
                          "bound": "tensor", "achieved": act_flop / (ms * 1e-3) / 1e12, "peak": tc_peak, "unit": "TFLOP/s",
                          "traffic": tr("fused"), "ms": ms, "also_hbm_bytes": env_bytes,
                          "ms_note": "ms: CUDA events around the one launch = the timed step (max over ranks); the env step's "
                                     "algorithmic bytes ride in the same launch (also_hbm_bytes)",
                          "peak_source": peak_src, "traffic_source": traffic_src}
            fused_roof["frac"] = fused_roof["achieved"] / tc_peak
            if fspan:
                fused_roof["device_span_ms"] = fspan
                fused_roof["achieved_device_span"] = act_flop / (fspan * 1e-3) / 1e12
                fused_roof["frac_device_span"] = fused_roof["achieved_device_span"] / tc_peak
            out["roofline"] = fused_roof
            out["roofline_other"] = act_roof
            out["roofline_env"] = env_roof
        if not args.no_extras and world_size == 1:
            scale = []
            del world
            torch.cuda.empty_cache()
            # 16 384 = envs per GPU of BASELINE.json configs[3] (128 row tiles: still one wave of the 148 SMs)
            for big in (16384, 65536, 1048576, 4194304):
                w = World(wl, big, args.precision, device, seed=99)
                small = big * 335 < (96 << 20)          # state + replay rows fit in L2: flush between steps, as at 4096
                a_ms, e_ms = kernel_split(w, 20, flush if small else None)
                how = "events around every launch, L2 flushed between steps"
                if not small:
                    a_ms, e_ms = kernels_back_to_back(w)
                    how = "20 launches of the same kernel back to back inside one event bracket (state exceeds L2)"
                for _ in range(3):
                    w.step()
                m, _ = device_timed(w, 20, 3, flush if small else None)
                scale.append({"envs": big, "value": big * 1e3 / m, "ms_per_step": m, "act_ms": a_ms, "env_ms": e_ms,
                              "kernel_ms_measured": how,
                              "env_traffic": traffic.get(("env", big)), "act_traffic": traffic.get(("actor", big)),
                              "env_hbm_frac": (ENV_BYTES[wl["variant"]] + REPLAY_BYTES) * big / (e_ms * 1e-3) / 1e9 / hbm_peak,
                              "act_tensor_frac": ACTOR_FLOP[wl["act_dim"]] * big / (a_ms * 1e-3) / 1e12 / tc_peak,
                              "l2": "flushed between steps" if small else "state exceeds L2 (no flush needed)"})
                del w
                torch.cuda.empty_cache()
            out["at_scale"] = scale
        if not args.no_extras and world_size == 1:   # the CPU baseline is a 1-GPU-run item: rank 0 at N = 1 only
            v, k, dt = time_cpu_port(wl, n)
            out["cpu_baseline"] = {"value": v, "unit": UNIT, "cores": host_cores(), "kind": "port",
                                   "sample": f"{k} steps x {n} envs in {dt:.1f} s (numpy env oracle + CPU-torch actor, oracle/)",
                                   "parts": time_cpu_parts(wl, n)}
        emit(out)
    if world_size > 1:
        dist.barrier()
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
