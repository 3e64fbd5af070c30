"""TEST INFRASTRUCTURE ONLY — runs the UNMODIFIED reference training loop
(/root/reference/src/my_trainers/sequential.py + base_KA_DDPG.py, and sequential_KA.py + base_KA.py) over the
UNMODIFIED reference task and agent, through ``oracle/ref_harness`` (import stubs, fake articulation view,
float64 skid-steer twin as the physics callback, skrl shim), with every random draw supplied, and writes
``tests/golden/trainer_*.npz``.  On the way it checks the restated loop (oracle/trainer_oracle.py) against it.
Build container only:  python -m oracle.gen_golden_trainer

What the vectors pin: act -> env.step -> relabel ``wheel_torqs / action_scale``
(base_KA_DDPG.py:212; absent in base_KA.py) -> record_transition -> post_interaction TWICE (two updates per step,
base_KA_DDPG.py:265,283) -> ``states = next_states``; the wrapper's single reset step with zero actions; ring-buffer
wrap of the replay memory.
"""
from __future__ import annotations

import contextlib
import io
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
GOLDEN = os.path.join(ROOT, "tests", "golden")
F = np.float32

from oracle.env_oracle import VARIANTS, EnvOracle, EnvParams  # noqa: E402
from oracle.gen_golden import load_state_into_reference, synth_state  # noqa: E402
from oracle.gen_golden_agent import STRIDE, build_reference_agent, subsample  # noqa: E402
from oracle.trainer_oracle import TrainerOracle  # noqa: E402

CASES = [
    # tag, env module, trainer module, N, timesteps, batch, memory steps, random_timesteps, learning_starts, ou scale, T_e
    ("kamma_medium", "KAMMA_medium", "sequential", 16, 10, 32, 8, 2, 3, 0.05, 5e5),
    ("kaddpg_2d", "KA_DDPG_2D", "sequential_KA", 16, 8, 32, 16, 1, 2, 0.3, 7e5),
    ("kaddpg_1d", "KA_DDPG_1D", "sequential_KA", 64, 8, 32, 16, 1, 2, 0.5, 7e5),     # BASELINE.json configs[0]: KA-DDPG-1d, N = 64
    ("kaddpg_4d", "KA_DDPG_4D", "sequential", 16, 8, 32, 16, 1, 2, 0.15, 7e5),       # linear blend with the decaying gamma + relabel
]


class RefVecEnv:
    """skrl ``wrap_env(env, "omniverse-isaacgym")`` over OmniIsaacGymEnvs ``VecEnvRLGames`` (third-party, restated:
    SURVEY.md §8(c)): reset() runs once and steps with zero actions; step() clamps actions / observations and
    returns ``obs, rew[N,1], reset[N,1], truncated = 0, extras``."""

    def __init__(self, runner, variant, draws):
        self.runner, self.task, self.variant, self.draws = runner, runner.task, variant, draws
        self.num_envs, self.num_agents, self.device = runner.n, 1, torch.device("cpu")
        self.observation_space, self.action_space = self.task.observation_space, self.task.action_space
        self.k = 0
        self._reset_once = True
        self._obs = None
        self.log = []

    def _step(self, actions):
        d = self.draws
        rm = d["rand_mask"][self.k] if self.variant.blend == "switch" else None
        obs, rew, reset, extras = self.runner.step(actions.numpy(), rand_mask=rm, reset_uniforms=d["reset_uniforms"][self.k])
        self.k += 1
        return torch.clamp(obs, -self.task.clip_obs, self.task.clip_obs), rew, reset, extras

    def reset(self):
        if self._reset_once:
            self.task.reset()
            self._obs, _, _, _ = self._step(torch.zeros(self.num_envs, self.task.num_actions))
            self._reset_once = False
        return self._obs, {}

    def step(self, actions):
        a = torch.clamp(actions, -self.task.clip_actions, self.task.clip_actions)
        obs, rew, reset, extras = self._step(a)
        self.log.append(dict(obs=obs.numpy().copy(), rew=rew.numpy().copy(), reset=reset.numpy().astype(np.uint8),
                             wheel_torqs=self.task.wheel_torqs.numpy().copy()))
        return obs, rew.view(-1, 1), reset.view(-1, 1), torch.zeros_like(reset.view(-1, 1)), extras

    def render(self, *a, **k):
        pass

    def close(self):
        pass


def run_case(idx, tag, env_mod, trainer_mod, n, T, batch, mem_steps, random_ts, learn_starts, ou_scale, T_e):
    import importlib

    from oracle.ref_harness import stubs
    from oracle.ref_harness.ref_env import RefRunner

    stubs.install()
    stubs._register("skrl.envs.wrappers.torch", Wrapper=object)
    variant = VARIANTS[env_mod]
    A = variant.act_dim
    seed = 9100 + idx
    rng = np.random.default_rng(seed)
    draws = dict(rand_mask=rng.random((T + 1, n)).astype(F), reset_uniforms=rng.random((T + 1, n, 8)).astype(F),
                 normals=rng.standard_normal((T, n, A)).astype(F), uniforms=rng.random((T, n, A)).astype(F))
    runner = RefRunner(env_mod, n, physics=True, init_uniforms=rng.random((n, 8)).astype(F))
    s0 = synth_state(variant, n, rng, physics=True)
    s0.episode_buf[:] = rng.integers(88, 99, n)        # a good share of the envs time out within the run
    load_state_into_reference(runner, variant, s0)
    env = RefVecEnv(runner, variant, draws)

    torch.manual_seed(seed)
    low = float(env.action_space.low[0])
    agent, memory, _ = build_reference_agent(6, A, "sigmoid", batch, ou_scale, low, T_e, 3e-4, 1e-3, 0.999, 0.005,
                                             mem_steps * n, n_envs=n, random_timesteps=random_ts,
                                             learning_starts=learn_starts, timesteps=T)
    # supplied draws: OU normals, random-phase uniforms; record the host randperm indices of every update
    step_of = {"t": 0}
    acts, indexes = [], []

    class FakeNormal:
        def sample(self, size):
            return torch.as_tensor(draws["normals"][step_of["t"]]).clone()

    class FakeUniform:
        def sample(self, sample_shape):
            u = torch.as_tensor(draws["uniforms"][step_of["t"]])
            return u * (torch.tensor(1.0) - torch.tensor(low)) + torch.tensor(low)

    agent._exploration_noise.distribution = FakeNormal()
    agent.policy._random_distribution = FakeUniform()
    orig_act, orig_sample = agent.act, memory.sample

    def act(states, timestep, timesteps):
        step_of["t"] = timestep
        out = orig_act(states, timestep=timestep, timesteps=timesteps)
        acts.append(out[0].detach().numpy().copy())
        return out

    def sample(names, batch_size, mini_batches=1, sequence_length=1):
        out = orig_sample(names=names, batch_size=batch_size, mini_batches=mini_batches)
        indexes.append(memory.sampling_indexes.numpy().copy())
        return out

    agent.act, memory.sample = act, sample
    trainer_cls = importlib.import_module(f"my_trainers.{trainer_mod}").SequentialTrainer
    trainer = trainer_cls(cfg={"timesteps": T, "headless": True, "disable_progressbar": True,
                               "close_environment_at_exit": False}, env=env, agents=agent)
    torch.manual_seed(seed + 1)     # the randperm stream of the updates
    with contextlib.redirect_stdout(io.StringIO()):
        trainer.train()
    n_updates = 2 * sum(1 for t in range(T) if t >= learn_starts)
    assert len(indexes) == n_updates, (len(indexes), n_updates)

    # ---- the restated loop on the same draws
    from oracle.agent_oracle import flat_params

    torch.manual_seed(seed)
    orc = TrainerOracle(variant, s0.copy(), n=n, batch=batch, mem_steps=mem_steps, random_timesteps=random_ts,
                        learning_starts=learn_starts, ou_scale=ou_scale, T_e=T_e, clip_low=low,
                        relabel=trainer_mod == "sequential", clip_obs=float(runner.task.clip_obs))
    orc.run(T, draws, indexes)
    worst = {}
    for t in range(T):
        for key in ("obs", "rew", "wheel_torqs"):
            d = np.abs(env.log[t][key].astype(np.float64) - orc.log[t][key].astype(np.float64)).max()
            worst[key] = max(worst.get(key, 0.0), float(d))
            assert d <= 2e-5 * max(1.0, np.abs(env.log[t][key]).max()), (tag, t, key, d)
        assert np.array_equal(env.log[t]["reset"].astype(bool), orc.log[t]["reset"]), (tag, t)
        d = np.abs(acts[t] - orc.log[t]["actions"]).max()
        worst["actions"] = max(worst.get("actions", 0.0), float(d))
        assert d <= 1e-5, (tag, t, d)
    for name in ("states", "actions", "rewards", "next_states"):
        a, b = memory.tensors_view[name].numpy(), orc.memory[name]
        assert np.array_equal(np.isnan(a), np.isnan(b)), (tag, name)
        d = np.nanmax(np.abs(a.astype(np.float64) - b.astype(np.float64)))
        worst["mem_" + name] = float(d)
        assert d <= 2e-5 * max(1.0, np.nanmax(np.abs(a))), (tag, name, d)
    assert np.array_equal(memory.tensors_view["terminated"].numpy(), orc.memory["terminated"])
    assert memory.memory_index == orc.memory_index and memory.filled == orc.filled
    for name, net_r, net_o in (("policy", agent.policy, orc.agent.policy), ("critic", agent.critic, orc.agent.critic),
                               ("target_policy", agent.target_policy, orc.agent.target_policy),
                               ("target_critic", agent.target_critic, orc.agent.target_critic)):
        d = (flat_params(net_r) - flat_params(net_o)).abs().max().item()
        worst[name] = d
        assert d <= 2e-5, (tag, name, d)

    rec = dict(tag=tag, variant=env_mod, trainer=trainer_mod, seed=seed, n=n, timesteps=T, batch=batch, mem_steps=mem_steps,
               random_timesteps=random_ts, learning_starts=learn_starts, ou_scale=ou_scale, T_e=T_e, clip_low=low,
               relabel=int(trainer_mod == "sequential"), stride=STRIDE, n_updates=n_updates,
               clip_obs=float(runner.task.clip_obs))
    for k, a in s0.__dict__.items():
        rec[f"s0_{k}"] = a
    for k, a in draws.items():
        rec[f"draw_{k}"] = a
    rec["indexes"] = np.stack(indexes)
    rec["actions"] = np.stack(acts)
    for key in ("obs", "rew", "reset", "wheel_torqs"):
        rec[key] = np.stack([env.log[t][key] for t in range(T)])
    for name in ("states", "actions", "rewards", "next_states", "terminated"):
        rec[f"mem_{name}"] = memory.tensors_view[name].numpy().astype(np.uint8 if name == "terminated" else F)
    rec["memory_index"], rec["filled"] = memory.memory_index, int(memory.filled)
    for name, net in (("policy", agent.policy), ("critic", agent.critic), ("target_policy", agent.target_policy),
                      ("target_critic", agent.target_critic)):
        fr = flat_params(net)
        rec[f"{name}_sub"], rec[f"{name}_sum"] = subsample(fr), fr.double().sum().item()
    sc = agent._state_preprocessor
    rec["scaler_mean"], rec["scaler_var"] = sc.running_mean.numpy().copy(), sc.running_variance.numpy().copy()
    rec["scaler_count"] = float(sc.current_count)
    resets = int(rec["reset"].sum())
    print(f"  trainer {tag:14s} N={n} T={T} updates={n_updates} resets={resets} wrap={bool(memory.filled)} worst: "
          + " ".join(f"{k}={v:.1e}" for k, v in worst.items()))
    return rec


def main():
    os.makedirs(GOLDEN, exist_ok=True)
    print("trainer cases (reference my_trainers/* over reference env + agent vs oracle/trainer_oracle.py):")
    only = set(sys.argv[1:])            # optional: case tags to (re)generate; default all (the seed of a case is its index)
    for i, case in enumerate(CASES):
        if only and case[0] not in only:
            continue
        np.savez_compressed(os.path.join(GOLDEN, f"trainer_{case[0]}.npz"), **run_case(i, *case))


if __name__ == "__main__":
    main()
