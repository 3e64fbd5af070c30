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

METRIC = "env-steps/s (dyn+obs+reward+act)"
UNIT = "env-steps/s"
ENV_BYTES = {"KAMMA_medium": 265, "KA_DDPG_2D": 253, "KA_DDPG_1D": 249}   # SURVEY.md §8(d), per env-step
REPLAY_BYTES = 70
ACTOR_FLOP = {4: 662528, 2: 662016, 1: 661760}                             # 2 * MACs, SURVEY.md §8(d)
UPDATE_FLOP = 6645248 * 512                                                # SURVEY.md §8 row A11, B = 512
# dram__bytes_read.sum + dram__bytes_write.sum per launch, from the committed `ncu --set full` captures
# (profiles/r01g_ncu_full_summary.txt; kamma_4d): {(kernel, envs): bytes}
NCU_TRAFFIC = {("env", 4096): 659968, ("actor", 4096): 923648, ("env", 1048576): 159438592 + 233006336,
               ("actor", 1048576): 42703616 + 468480}
WORKLOADS = {
    "kamma_4d": dict(variant="KAMMA_medium", act_dim=4, envs=4096, ou=0.05, T_e=5e5, memory_total=1_000_000,
                     name="KAMMA-4d nofifo random RLIL (BASELINE.json configs[2])"),
    "kaddpg_2d": dict(variant="KA_DDPG_2D", act_dim=2, envs=4096, ou=0.3, T_e=7e5, memory_total=35_000,
                      name="KA-DDPG-2d nofifo random RLIL (BASELINE.json configs[1])"),
}


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
        self.agent = AgentOracle(6, wl["act_dim"], "sigmoid", ou=(0.15, 0.1, wl["ou"]), exploration_timesteps=wl["T_e"])
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
    for _ in range(max(args.warmup, 1) if args.warmup <= 3 else 3):
        port.step()
    steps = min(args.steps, 50)
    t0 = time.perf_counter()
    for _ in range(steps):
        port.step()
    dt = time.perf_counter() - t0
    v = n * steps / dt
    cores = host_cores()
    sample = f"{steps} steps x {n} envs (single host process; the CPU path does not shard across GPUs)"
    print(json.dumps({
        "impl": "reference", "metric": METRIC, "value": v, "unit": UNIT, "n_gpus": args.gpus, "steps": steps,
        "warmup": min(args.warmup, 3), "ms_per_step": 1e3 * dt / steps, "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "f32 (dynamics f64)", "data": "synthetic",
        "config": {"workload": wl["name"], "envs_per_gpu": n, "l2": "n/a (host)"},
        "cpu_baseline": {"value": v, "unit": UNIT, "cores": cores, "kind": "port", "sample": sample},
        "e2e": {"value": v, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "note": "reference = pure Python on Isaac Sim/PhysX + skrl: not runnable on this box; this is its pinned CPU port",
    }), flush=True)


# ------------------------------------------------------------------------------------------------ our arm
class World:
    """task + memory + agent of one rank, wired like train/train_kamma_4d_nofifo_random_RLIL.py:94-186."""

    def __init__(self, wl, n, precision, device, seed):
        import torch

        import options_in_simulation_b200 as ois
        from options_in_simulation_b200 import agent as A
        from options_in_simulation_b200.config import script_cfg
        from options_in_simulation_b200.memory import memory_size_for

        self.torch, self.n, self.wl = torch, n, wl
        torch.manual_seed(seed)
        cfg = script_cfg(n, seed=seed, sim_device=str(device))
        self.task = ois.TorqueDistributionTask(sim_config=cfg, variant=wl["variant"], seed=seed)
        self.env = ois.wrap_env(self.task)
        dev = str(device)
        osp, asp = self.task.observation_space, self.task.action_space
        models = {"policy": A.DeterministicActor(osp, asp, dev), "target_policy": A.DeterministicActor(osp, asp, dev),
                  "critic": A.Critic(osp, asp, dev), "target_critic": A.Critic(osp, asp, dev)}
        self.memory = ois.RandomMemory(memory_size=memory_size_for(wl["memory_total"], n), num_envs=n, device=dev,
                                       replacement=False)
        cfg_a = {"batch_size": 512, "discount_factor": 0.999, "polyak": 0.005, "actor_learning_rate": 3e-4,
                 "critic_learning_rate": 1e-3, "random_timesteps": 0, "learning_starts": 0,
                 "mixed_precision": precision == "bf16", "sampler": "device", "graph_update": True,
                 "state_preprocessor": A.RunningStandardScaler, "state_preprocessor_kwargs": {"size": osp, "device": dev},
                 "exploration": {"noise": A.OrnsteinUhlenbeckNoise(0.15, 0.1, wl["ou"], device=dev),
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

    def step(self):
        """One policy step through the public API (what the trainer's loop body does, base_KA_DDPG.py:175-219)."""
        actions = self.agent.act(self.obs, self.t, self.T)[0]
        self.obs, rew, term, trunc, _ = self.env.step(actions)
        self.agent.record_transition(None, None, rew, self.obs, term, trunc, None, self.t, self.T)
        self.t += 1
        return rew, term


def device_timed(world, steps, warmup, flush):
    torch = world.torch
    for _ in range(warmup):
        world.step()
    ev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(steps)]
    torch.cuda.synchronize()
    l0 = world.launches()
    for a, b in ev:
        if flush is not None:
            flush.zero_()
        a.record()
        world.step()
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

    def one():
        obs, rew, term = g.step()          # synchronised: results are in pinned host memory
        g.h_states.copy_(obs)              # the host consumes them: next step's input (host memcpy, 96 KB at N = 4096)

    for _ in range(warmup + 2):            # +2: the eager warm-up step and the capture
        one()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    torch.cuda.synchronize()
    a.record()
    for _ in range(steps):
        one()
    b.record()
    torch.cuda.synchronize()
    world.t = g.timestep
    world.task._obs_clipped[0].copy_(g.h_next_states)     # zero-copy mode leaves the task's device-side copy untouched
    world.obs = world.task.obs_clipped
    return a.elapsed_time(b) / steps, n * 6 * 4, n * 29


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
    return sum(a.elapsed_time(b) for a, b in ev) / steps


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
    """One policy step followed by two updates (the reference trainer's cadence), device-timed."""
    torch = world.torch
    for _ in range(3):
        world.step()
        world.agent._update(world.t, world.T)
        world.agent._update(world.t, world.T)
    ev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(steps)]
    torch.cuda.synchronize()
    for a, b in ev:
        flush.zero_()
        a.record()
        world.step()
        world.agent._update(world.t, world.T)
        world.agent._update(world.t, world.T)
        b.record()
    torch.cuda.synchronize()
    return sum(a.elapsed_time(b) for a, b in ev) / steps


def main():
    args = parse()
    wl = WORKLOADS[args.workload]
    n = args.envs or wl["envs"]
    if args.impl == "reference":
        run_reference(args, wl, n)
        return

    import torch
    import torch.distributed as dist

    import options_in_simulation_b200 as ois

    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a CUDA device (B200); there is no CPU fallback for the product path")
    if not os.path.exists(ois.LIB_PATH):
        raise SystemExit(f"{ois.LIB_PATH} missing: run `python -c 'import __graft_entry__ as g; g.build()'` first")
    rank, local, world_size = int(os.environ.get("RANK", 0)), int(os.environ.get("LOCAL_RANK", 0)), int(os.environ.get("WORLD_SIZE", 1))
    torch.cuda.set_device(local)
    device = torch.device("cuda", local)
    if world_size > 1:
        dist.init_process_group("nccl", device_id=device)
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
    e2e_ms, h2d, d2h = e2e_graph_timed(world, min(args.steps, 200), 5)
    barrier()
    e2e_copy_ms, _, _ = e2e_graph_timed(world, min(args.steps, 200), 5, mode="copy")
    barrier()
    graph_ms = graph_device_timed(world, args.steps, 5, flush)
    barrier()
    host_ms = host_issue_timed(world)
    barrier()
    clocks = sampler.stop() if rank == 0 else None      # sampled across ALL the timed legs above
    per_rank = [[ms, graph_ms, host_ms]]
    if world_size > 1:
        mine = torch.tensor(per_rank[0], device=device, dtype=torch.float64)
        allr = [torch.empty_like(mine) for _ in range(world_size)]
        dist.all_gather(allr, mine)
        per_rank = [a.tolist() for a in allr]
        t = torch.tensor([ms, e2e_ms, e2e_eager_ms, graph_ms, e2e_copy_ms], device=device, dtype=torch.float64)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        ms, e2e_ms, e2e_eager_ms, graph_ms, e2e_copy_ms = t.tolist()
    # the same device-resident step issued two ways, both timed over EXACTLY args.steps steps, max over ranks: call by
    # call (2 launches per step from Python) and as one graph launch per step; the headline is the faster of the two
    # (with several ranks sharing a host the call-by-call loop becomes host-bound on the slowest rank)
    call_ms = ms
    issue = "call-by-call (agent.act + env.step)"
    if graph_ms < ms:
        ms, issue, launches = graph_ms, "rollout.RolloutGraph replay (one graph launch per step)", 2 * args.steps
    value = world_size * n * 1e3 / ms
    e2e_value = world_size * n * 1e3 / e2e_ms

    out = {
        "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world_size, "steps": args.steps, "warmup": args.warmup,
        "ms_per_step": ms, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
        "dtype": "bf16 (actor MLP, fp32 accumulate) + f32 (env)" if args.precision == "bf16" else "f32",
        "data": "synthetic",
        "config": {"workload": wl["name"], "envs_per_gpu": n, "variant": wl["variant"], "act_dim": wl["act_dim"],
                   "replay": "fused append, rank-local shard", "rng": "in-kernel Philox",
                   "l2": "flushed between timed steps (256 MiB memset outside the event brackets)",
                   "parallelism": f"env/replay shards x{world_size}, no data-path collective", "issue": issue},
        "e2e": {"value": e2e_value, "unit": UNIT, "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h,
                "ms_per_step": e2e_ms,
                "path": "rollout.RolloutGraph(host_io=True): one CUDA-graph launch per step = actor kernel reading the "
                        "observations from pinned host memory in place + env kernel (step, replay append, step-control "
                        "advance) storing obs|rew|reset to pinned host memory in place; host sync every step; the bytes "
                        "below cross PCIe every step as zero-copy loads / stores",
                "copy_nodes_ms_per_step": e2e_copy_ms,
                "call_by_call_ms_per_step": e2e_eager_ms,
                "call_by_call_value": world_size * n * 1e3 / e2e_eager_ms},
        "issue_paths": {"call_by_call_ms_per_step": call_ms, "graph_replay_ms_per_step": graph_ms,
                        "host_issue_ms_per_step": max(r[2] for r in per_rank),
                        "per_rank_ms": [[round(x, 5) for x in r] for r in per_rank],
                        "note": "max over ranks; per_rank_ms = [call-by-call, graph replay, host issue time] per rank"},
        "gpu_launches": launches, "clocks": clocks,
    }

    if rank == 0:
        act_ms, env_ms = kernel_split(world, min(args.steps, 100), flush)
        env_bytes = (ENV_BYTES[wl["variant"]] + REPLAY_BYTES) * n
        act_flop = ACTOR_FLOP[wl["act_dim"]] * n
        tr = (lambda k: NCU_TRAFFIC.get((k, n)) if args.workload == "kamma_4d" else None)
        env_roof = {"kernel": "env_step_kernel", "bound": "hbm", "achieved": env_bytes / (env_ms * 1e-3) / 1e9,
                    "peak": hbm_peak, "unit": "GB/s", "traffic": tr("env"), "ms": env_ms}
        act_roof = {"kernel": "actor MLP (ois_agent_act)", "bound": "tensor", "achieved": act_flop / (act_ms * 1e-3) / 1e12,
                    "peak": tc_peak, "unit": "TFLOP/s", "traffic": tr("actor"), "ms": act_ms}
        for r in (env_roof, act_roof):
            r["frac"] = r["achieved"] / r["peak"]
            r["peak_source"] = peak_src
        dom, other = (act_roof, env_roof) if act_ms >= env_ms else (env_roof, act_roof)
        out["roofline"] = dom
        out["roofline_other"] = other
        if not args.no_extras and world_size == 1:
            upd_ms = update_timed(world)
            train_ms = train_timed(world, flush)
            out["update"] = {"ms_per_update": upd_ms, "batch": 512, "tflops": UPDATE_FLOP / (upd_ms * 1e-3) / 1e12,
                             "path": "bf16 tcgen05 grouped GEMMs, one CUDA graph per update" if args.precision == "bf16" else "fp32 SIMT",
                             "train_ms_per_step": train_ms, "train_env_steps_per_s": n * 1e3 / train_ms,
                             "note": "DDPG._update; train_* = one policy step + 2 updates, the reference cadence "
                                     "(base_KA_DDPG.py:265,283); not in the metric"}
            scale = []
            del world
            torch.cuda.empty_cache()
            # 16 384 = envs per GPU of BASELINE.json configs[3] (128 row tiles: still one wave of the 148 SMs)
            for big in (16384, 65536, 1048576):
                w = World(wl, big, args.precision, device, seed=99)
                small = big * 335 < (96 << 20)          # state + replay rows fit in L2: flush between steps, as at 4096
                a_ms, e_ms = kernel_split(w, 20, flush if small else None)
                for _ in range(3):
                    w.step()
                m, _ = device_timed(w, 20, 3, flush if small else None)
                scale.append({"envs": big, "value": big * 1e3 / m, "ms_per_step": m, "act_ms": a_ms, "env_ms": e_ms,
                              "env_traffic": NCU_TRAFFIC.get(("env", big)), "act_traffic": NCU_TRAFFIC.get(("actor", big)),
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
        print(json.dumps(out), flush=True)
    if world_size > 1:
        dist.barrier()
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
