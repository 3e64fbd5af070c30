"""``options-sim-train`` / ``options-sim-eval`` on the B200 path.

The reference's CLIs are routers: ``options-sim-train --algorithm kamma --action-dim 4d --fifo nofifo --curriculum random
--learning-strategy rlil [-- legacy script flags]`` picks ``train/train_<...>.py`` and runs it in a subprocess
(src/options_sim/cli/train.py:33-160); each of those scripts is the same 200 lines with different constants
(SURVEY.md Appendix B).  Here the constants are a table (``SCRIPTS``) and ONE function builds what the script builds —
task, wrapper, memory, four models, DDPG cfg, SequentialTrainer — on this package's mirrors, with the same flags after
``--`` (``--stiffness --damping --static_friction --dynamic_friction --yaw_constant --linear_x --seed``,
train/train_kamma_4d_nofifo_random_RLIL.py:63-72).  Additional flags (not in the reference) size the run for a GPU:
``--num-envs`` (the scripts hard-code numEnvs = 1), ``--timesteps``, ``--precision``, ``--graph-rollout``.

    python -m options_in_simulation_b200.cli train --algorithm kamma --action-dim 4d --fifo nofifo --curriculum random \\
        --learning-strategy rlil --num-envs 4096 --timesteps 2000 -- --seed 1
    python -m options_in_simulation_b200.cli eval --algorithm kamma --action-dim 4d --num-envs 8192 --timesteps 200 \\
        --checkpoint-path runs/.../checkpoints/agent_500000.pt

Out of scope (not on the hot path): wandb, the TensorBoard -> CSV artefact tools, the ``scripts/legacy_map.json`` override.
"""
from __future__ import annotations

import argparse
import json
import os
import sys

# (algorithm, action_dim, fifo, curriculum, strategy) -> constants of the reference script
SCRIPTS = {
    ("kaddpg", "1d", "nofifo", "random", "rlil"): dict(
        script="train/train_kaddpg_1d_nofifo_random_RLIL.py", variant="KA_DDPG_1D", memory=("fifo", 35_000), ou=0.5,
        T_e=7e5, random_timesteps=80, timesteps=700_000, checkpoint_interval=700_000),
    ("kaddpg", "2d", "nofifo", "random", "rlil"): dict(
        script="train/train_kaddpg_2d_nofifo_random_RLIL.py", variant="KA_DDPG_2D", memory=("fifo", 35_000), ou=0.3,
        T_e=7e5, random_timesteps=80, timesteps=700_000, checkpoint_interval=700_000),
    ("kaddpg", "4d", "nofifo", "random", "rlil"): dict(
        script="train/train_kaddpg_4d_nofifo_random_RLIL.py", variant="KA_DDPG_4D", memory=("fifo", 35_000), ou=0.15,
        T_e=7e5, random_timesteps=80, timesteps=1_200_000, checkpoint_interval=500_000),
    ("kamma", "4d", "nofifo", "random", "rlil"): dict(
        script="train/train_kamma_4d_nofifo_random_RLIL.py", variant="KAMMA_medium", memory=("random", 1_000_000), ou=0.05,
        T_e=5e5, random_timesteps=1000, timesteps=600_000, checkpoint_interval=500_000),
    ("kamma", "4d", "fifo", "random", "rlil"): dict(
        script="train/train_kamma_4d_fifo_random_RLIL.py", variant="KAMMA", memory=("fifo", 10_000), ou=0.05,
        T_e=4e5, random_timesteps=80, timesteps=600_000, checkpoint_interval=400_000),
    ("kamma", "4d", "fifo", "gv", "rlil"): dict(
        script="train/train_kamma_4d_fifo_gv_RLIL.py", variant="KAMMA_curr1", memory=("fifo", 10_000), ou=0.05,
        T_e=5e5, random_timesteps=80, timesteps=600_000, checkpoint_interval=500_000),
    ("kamma", "4d", "fifo", "bd", "rlil"): dict(
        script="train/train_kamma_4d_fifo_bd_RLIL.py", variant="KAMMA_curr2", memory=("fifo", 10_000), ou=0.05,
        T_e=5e5, random_timesteps=80, timesteps=600_000, checkpoint_interval=500_000),
    ("kamma", "4d", "nofifo", "bd", "rlil"): dict(
        script="train/train_kamma_4d_nofifo_bd_RLIL.py", variant="KAMMA_curr2", memory=("random", 1_000_000), ou=0.1,
        T_e=5e5, random_timesteps=80, timesteps=500_000, checkpoint_interval=500_000),
    ("kamma", "4d", "nofifo", "gv", "rlil"): dict(
        script="train/train_kamma_4d_nofifo_gv_RLIL.py", variant="KAMMA_curr1", memory=("random", 1_000_000), ou=0.1,
        T_e=5e5, random_timesteps=80, timesteps=500_000, checkpoint_interval=500_000),
    ("kamma", "4d", "nofifo", "random", "il"): dict(
        script="train/train_kamma_4d_nofifo_random_IL.py", variant="KAMMA_g1d_g21", memory=("random", 1_000_000), ou=0.05,
        T_e=4e5, random_timesteps=80, timesteps=600_000, checkpoint_interval=400_000),
}
# src/options_sim/cli/eval.py:10-14 (EVAL_MAP); the actors of the 1d / 2d scripts end in a sigmoid
# (eval/eval_kaddpg_1d.py:29-36), the 4d one in a tanh (eval/eval_kamma_4d.py:33-36)
EVAL = {
    ("kaddpg", "1d"): dict(script="eval/eval_kaddpg_1d.py", variant="KA_DDPG_1D_eval", timesteps=50_000, output="sigmoid"),
    ("kaddpg", "2d"): dict(script="eval/eval_kaddpg_2d.py", variant="KA_DDPG_2D_eval", timesteps=50_000, output="sigmoid"),
    ("kamma", "4d"): dict(script="eval/eval_kamma_4d.py", variant="KAMMA_eval", timesteps=50_000, output="tanh"),
}
# The two scripts no router entry of the reference reaches (they are run by path); ``--script`` selects them here.
UNROUTED_TRAIN = {
    "train/train_kamma_controller.py": dict(
        script="train/train_kamma_controller.py", variant="KAMMA_g11_g21", memory=("random", 1_000_000), ou=0.05,
        T_e=4e5, random_timesteps=1000, timesteps=400_000, checkpoint_interval=200_000),     # :107,144,167,174-175,180,193
}
UNROUTED_EVAL = {
    "eval/eval_kamma_controller.py": dict(script="eval/eval_kamma_controller.py", variant="KAMMA_eval", timesteps=50_000,
                                          output="sigmoid"),                                  # :37,64,192
}


def _legacy_parser() -> argparse.ArgumentParser:
    """The flags every reference train / eval script parses (train/train_kamma_4d_nofifo_random_RLIL.py:63-72)."""
    p = argparse.ArgumentParser(add_help=False)
    p.add_argument("--stiffness", type=float, default=0.035)
    p.add_argument("--damping", type=float, default=0.005)
    p.add_argument("--static_friction", type=float, default=0.85)
    p.add_argument("--dynamic_friction", type=float, default=0.85)
    p.add_argument("--yaw_constant", type=float, default=0.5)
    p.add_argument("--linear_x", type=float, nargs=2, default=[1.0, 2.0])
    p.add_argument("--seed", type=int, default=42)
    return p


def _split(argv):
    if "--" in argv:
        i = argv.index("--")
        return argv[:i], argv[i + 1:]
    return argv, []


def build(variant: str, spec: dict, legacy, *, num_envs: int, precision: str, device: str, timesteps: int,
          directory: str | None, write_interval, checkpoint_interval, output: str = "sigmoid", ou_scale=None,
          random_timesteps=None):
    """What a reference script builds between its argparse block and ``trainer.train()``
    (train/train_kamma_4d_nofifo_random_RLIL.py:74-186), on this package's mirrors."""
    import torch

    from . import agent as A
    from . import envs, memory
    from .config import script_cfg

    torch.manual_seed(legacy.seed)                       # skrl set_seed (:57)
    cfg = script_cfg(num_envs, stiffness=legacy.stiffness, damping=legacy.damping, static_friction=legacy.static_friction,
                     dynamic_friction=legacy.dynamic_friction, yaw_constant=legacy.yaw_constant,
                     linear_x=tuple(legacy.linear_x), seed=legacy.seed, sim_device=device)
    task = envs.TorqueDistributionTask(sim_config=cfg, variant=variant, seed=legacy.seed)
    env = envs.wrap_env(task, "omniverse-isaacgym")
    kind, total = spec.get("memory", ("random", 1_000_000))
    cls = memory.FIFOMemory if kind == "fifo" else memory.RandomMemory
    mem = cls(memory_size=memory.memory_size_for(total, num_envs), num_envs=num_envs, device=device, replacement=False)
    osp, asp = env.observation_space, env.action_space
    models = {"policy": A.DeterministicActor(osp, asp, device, output=output),
              "target_policy": A.DeterministicActor(osp, asp, device, output=output),
              "critic": A.Critic(osp, asp, device), "target_critic": A.Critic(osp, asp, device)}
    rt = spec.get("random_timesteps", 0) if random_timesteps is None else random_timesteps
    ou = spec.get("ou", 0.0) if ou_scale is None else ou_scale
    acfg = {"gradient_steps": 1, "batch_size": 512, "discount_factor": 0.999, "polyak": 0.005,
            "actor_learning_rate": 3e-4, "critic_learning_rate": 1e-3, "random_timesteps": rt, "learning_starts": rt,
            "state_preprocessor": A.RunningStandardScaler, "state_preprocessor_kwargs": {"size": osp, "device": device},
            "mixed_precision": precision == "bf16", "sampler": "device", "graph_update": precision == "bf16",
            "exploration": {"noise": A.OrnsteinUhlenbeckNoise(theta=0.15, sigma=0.1, base_scale=ou, device=device),
                            "initial_scale": 1.0, "final_scale": 1e-4, "timesteps": spec.get("T_e", timesteps)},
            "experiment": {"directory": directory, "experiment_name": None, "write_interval": write_interval,
                           "checkpoint_interval": checkpoint_interval}}
    agent = A.DDPG(models=models, memory=mem, cfg=acfg, observation_space=osp, action_space=asp, device=device)
    return task, env, mem, agent


def main_train(argv=None) -> int:
    router_argv, legacy_argv = _split(list(sys.argv[1:] if argv is None else argv))
    p = argparse.ArgumentParser(prog="options-sim-train", description="KA-DDPG / KAMMA training on the B200 path")
    by_path = any(x in ("--script", "--list") or x.startswith("--script=") for x in router_argv)
    p.add_argument("--algorithm", required=not by_path, choices=["kamma", "kaddpg"])  # src/options_sim/cli/train.py:51-55
    p.add_argument("--action-dim", required=not by_path, choices=["1d", "2d", "4d"])
    p.add_argument("--fifo", required=not by_path, choices=["fifo", "nofifo"])
    p.add_argument("--curriculum", required=not by_path, choices=["random", "gv", "bd"])
    p.add_argument("--learning-strategy", required=not by_path, choices=["rlil", "il"])
    p.add_argument("--script", default=None, choices=sorted(UNROUTED_TRAIN),
                   help="a reference script no router key reaches, by its path (instead of the five selectors)")
    p.add_argument("--dry-run", action="store_true", help="print the resolved configuration and exit")
    p.add_argument("--list", action="store_true", help="list the script configurations this build knows")
    # not in the reference: sizing for a GPU
    p.add_argument("--num-envs", type=int, default=4096)
    p.add_argument("--timesteps", type=int, default=None, help="default: the reference script's own")
    p.add_argument("--precision", choices=["bf16", "fp32"], default="bf16")
    p.add_argument("--device", default="cuda:0")
    p.add_argument("--graph-rollout", action="store_true", help="act + env step as one CUDA-graph launch per timestep")
    p.add_argument("--directory", default=None, help="experiment directory (default ./runs)")
    p.add_argument("--write-interval", type=int, default=800)                         # train/...RLIL.py:176
    p.add_argument("--checkpoint-interval", type=int, default=None)
    a = p.parse_args(router_argv)
    legacy = _legacy_parser().parse_args(legacy_argv)
    if a.list:
        for k, v in SCRIPTS.items():
            print("::".join(k), "->", v["script"], f"({v['variant']})")
        for k, v in UNROUTED_TRAIN.items():
            print("--script", k, "->", v["script"], f"({v['variant']})")
        return 0
    if a.script is not None:
        spec = UNROUTED_TRAIN[a.script]
    else:
        key = (a.algorithm, a.action_dim, a.fifo, a.curriculum, a.learning_strategy)
        if key not in SCRIPTS:
            print(f"[ERROR] no B200 configuration for {'::'.join(key)}; known: {[ '::'.join(k) for k in SCRIPTS ]}",
                  file=sys.stderr)
            return 2                                                                  # train.py's "no script found" exit
        spec = SCRIPTS[key]
    timesteps = spec["timesteps"] if a.timesteps is None else a.timesteps
    resolved = {"reference_script": spec["script"], "variant": spec["variant"], "memory": list(spec["memory"]),
                "num_envs": a.num_envs, "timesteps": timesteps, "precision": a.precision, "legacy_args": vars(legacy)}
    if a.dry_run:
        print(json.dumps(resolved))
        return 0
    from .trainer import SequentialTrainer

    ckpt = spec["checkpoint_interval"] if a.checkpoint_interval is None else a.checkpoint_interval
    graph = a.graph_rollout
    task, env, mem, agent = build(spec["variant"], spec, legacy, num_envs=a.num_envs, precision=a.precision, device=a.device,
                                  timesteps=timesteps, directory=a.directory,
                                  write_interval=0 if graph else a.write_interval, checkpoint_interval=ckpt)
    if graph or a.write_interval == 0:
        agent.fuse_replay_with(task)
    trainer = SequentialTrainer(cfg={"timesteps": timesteps, "headless": True, "graph_rollout": graph,
                                     "disable_progressbar": not sys.stdout.isatty()}, env=env, agents=agent)
    trainer.train()
    import torch

    torch.cuda.synchronize()
    if getattr(agent, "experiment_dir", None):
        agent.write_checkpoint(timesteps, timesteps)
    print(json.dumps({**resolved, "updates": agent._updates, "transitions": len(mem),
                      "experiment_dir": getattr(agent, "experiment_dir", None)}))
    return 0


def main_eval(argv=None) -> int:
    router_argv, legacy_argv = _split(list(sys.argv[1:] if argv is None else argv))
    p = argparse.ArgumentParser(prog="options-sim-eval", description="tracking-error-vs-speed eval sweep on the B200 path")
    by_path = any(x == "--script" or x.startswith("--script=") for x in router_argv)
    p.add_argument("--algorithm", "-a", required=not by_path, choices=["kaddpg", "kamma"])   # src/options_sim/cli/eval.py:47-48
    p.add_argument("--action-dim", "-d", required=not by_path, choices=["1d", "2d", "4d"])
    p.add_argument("--script", default=None, choices=sorted(UNROUTED_EVAL),
                   help="a reference eval script no router key reaches, by its path")
    p.add_argument("--dry-run", action="store_true", help="print the resolved configuration and exit")
    p.add_argument("--checkpoint-path", default=None, help="agent_<step>.pt written by the reference or by this package")
    p.add_argument("--num-envs", type=int, default=8192)
    p.add_argument("--timesteps", type=int, default=None, help="default: two episode lengths past the skipped first one")
    p.add_argument("--precision", choices=["bf16", "fp32"], default="bf16")
    p.add_argument("--device", default="cuda:0")
    p.add_argument("--bins", type=int, default=10)
    p.add_argument("--out", default=None, help="write tracking_error_vs_speed.csv here (speed, tracking_error, episodes)")
    a = p.parse_args(router_argv)
    legacy = _legacy_parser().parse_args(legacy_argv)
    if a.script is not None:
        spec = UNROUTED_EVAL[a.script]
    else:
        key = (a.algorithm, a.action_dim)
        if key not in EVAL:
            print(f"[ERROR] no B200 eval configuration for {key}; known: {list(EVAL)}", file=sys.stderr)
            return 2
        spec = EVAL[key]
    if a.dry_run:
        print(json.dumps({"reference_script": spec["script"], "variant": spec["variant"], "output": spec["output"],
                          "num_envs": a.num_envs, "precision": a.precision, "legacy_args": vars(legacy)}))
        return 0
    from .evaluate import phase_lock, tracking_error_sweep

    task, env, mem, agent = build(spec["variant"], {"T_e": 5e4}, legacy, num_envs=a.num_envs, precision=a.precision,
                                  device=a.device, timesteps=spec["timesteps"], directory=None, write_interval=0,
                                  checkpoint_interval=0, output=spec["output"], ou_scale=0.0, random_timesteps=0)
    if a.checkpoint_path:
        agent.load(a.checkpoint_path)                                               # eval/eval_kamma_4d.py:228
    env.reset()
    period = phase_lock(task)
    steps = 3 * period + 1 if a.timesteps is None else a.timesteps
    out = tracking_error_sweep(env, agent, steps, lo=float(legacy.linear_x[0]), hi=float(legacy.linear_x[1]), bins=a.bins,
                               skip_steps=period)
    rows = [(float(s), float(t), int(e)) for s, t, e in zip(out["speed"], out["tracking_error"], out["episodes"])]
    if a.out:
        os.makedirs(os.path.dirname(os.path.abspath(a.out)), exist_ok=True)
        with open(a.out, "w") as f:
            f.write("speed_mps,tracking_error_mps,episodes\n")
            f.writelines(f"{s:.6f},{t:.6f},{e}\n" for s, t, e in rows)
    print(json.dumps({"reference_script": spec["script"], "num_envs": a.num_envs, "timesteps": steps, "curve": rows}))
    return 0


def main(argv=None) -> int:
    argv = list(sys.argv[1:] if argv is None else argv)
    if not argv or argv[0] not in ("train", "eval"):
        print("usage: python -m options_in_simulation_b200.cli {train|eval} ...", file=sys.stderr)
        return 2
    return main_train(argv[1:]) if argv[0] == "train" else main_eval(argv[1:])


if __name__ == "__main__":
    raise SystemExit(main())
