"""TEST INFRASTRUCTURE ONLY — runs the UNMODIFIED reference agent (/root/reference/src/my_agents/ddpg.py,
on the skrl shim of oracle/ref_harness) on seeded synthetic inputs, checks oracle/agent_oracle.py against
it and writes ``tests/golden/agent_*.npz``.  Build container only (the GPU box has no /root/reference).

Weights are not stored in full (1.3 MB per net): the initial weights are regenerated from the torch seed
(default nn.Linear init, construction order policy, target_policy, critic, target_critic) and pinned by a
float64 checksum; final weights are stored as a strided subsample + per-tensor float64 sums.
"""
from __future__ import annotations

import os
import sys

import numpy as np
import torch
import torch.nn as nn

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
GOLDEN = os.path.join(ROOT, "tests", "golden")
STRIDE = 97

CASES = [
    # tag, obs, act, out_act, batch, ou base_scale, clip low, T_e, lr_a, lr_c, discount, polyak, rows, n_act
    ("kamma_4d", 6, 4, "sigmoid", 512, 0.05, 0.0, 5e5, 3e-4, 1e-3, 0.999, 0.005, 4096, 256),
    ("kaddpg_1d", 6, 1, "sigmoid", 512, 0.5, 0.0, 7e5, 3e-4, 1e-3, 0.999, 0.005, 2048, 64),
    ("kaddpg_2d", 6, 2, "sigmoid", 512, 0.3, 0.0, 7e5, 3e-4, 1e-3, 0.999, 0.005, 2048, 64),
    ("kamma_eval_tanh", 6, 4, "tanh", 64, 0.0, 0.0, 5e4, 3e-4, 1e-3, 0.999, 0.005, 512, 128),
]


def subsample(flat: torch.Tensor) -> np.ndarray:
    return flat[::STRIDE].numpy().copy()


def build_reference_agent(obs_dim, act_dim, out_act, batch, base_scale, clip_low, T_e, lr_a, lr_c, discount, polyak, rows,
                          *, n_envs=64, random_timesteps=10, learning_starts=10, timesteps=1000):
    from oracle.ref_harness import stubs

    stubs.install()
    from my_agents.ddpg import DDPG  # noqa: the unmodified reference class
    from oracle.ref_harness import skrl_shim as sk

    obs_space = stubs.Box(-1.0, 1.0, (obs_dim,))
    act_space = stubs.Box(clip_low, 1.0, (act_dim,))

    # model classes exactly as the train scripts define them (train/train_kamma_4d_nofifo_random_RLIL.py:21-52)
    class DeterministicActor(sk.DeterministicMixin, sk.Model):
        def __init__(self, observation_space, action_space, device, clip_actions=False):
            sk.Model.__init__(self, observation_space, action_space, device)
            sk.DeterministicMixin.__init__(self, clip_actions)
            self.net = nn.Sequential(nn.Linear(self.num_observations, 512), nn.ELU(), nn.Linear(512, 512), nn.ELU(),
                                     nn.Linear(512, 128), nn.ELU(), nn.Linear(128, self.num_actions),
                                     nn.Sigmoid() if out_act == "sigmoid" else nn.Tanh())

        def compute(self, inputs, role):
            return self.net(inputs["states"]), {}

    class Critic(sk.DeterministicMixin, sk.Model):
        def __init__(self, observation_space, action_space, device, clip_actions=False):
            sk.Model.__init__(self, observation_space, action_space, device)
            sk.DeterministicMixin.__init__(self, clip_actions)
            self.net = nn.Sequential(nn.Linear(self.num_observations + self.num_actions, 512), nn.ELU(),
                                     nn.Linear(512, 512), nn.ELU(), nn.Linear(512, 128), nn.ELU(), nn.Linear(128, 1))

        def compute(self, inputs, role):
            return self.net(torch.cat([inputs["states"], inputs["taken_actions"]], dim=1)), {}

    models = {"policy": DeterministicActor(obs_space, act_space, "cpu"),
              "target_policy": DeterministicActor(obs_space, act_space, "cpu"),
              "critic": Critic(obs_space, act_space, "cpu"), "target_critic": Critic(obs_space, act_space, "cpu")}
    memory = sk.RandomMemory(memory_size=rows // n_envs, num_envs=n_envs, device="cpu", replacement=False)
    from my_agents.ddpg import DDPG_DEFAULT_CONFIG
    import copy

    cfg = copy.deepcopy(DDPG_DEFAULT_CONFIG)
    cfg["exploration"]["noise"] = sk.OrnsteinUhlenbeckNoise(theta=0.15, sigma=0.1, base_scale=base_scale, device="cpu")
    cfg["exploration"]["final_scale"] = 1e-4
    cfg["exploration"]["timesteps"] = T_e
    cfg.update(gradient_steps=1, batch_size=batch, discount_factor=discount, polyak=polyak, actor_learning_rate=lr_a,
               critic_learning_rate=lr_c, random_timesteps=random_timesteps, learning_starts=learning_starts,
               state_preprocessor=sk.RunningStandardScaler,
               state_preprocessor_kwargs={"size": obs_space, "device": "cpu"})
    cfg["experiment"]["write_interval"] = 0
    cfg["experiment"]["checkpoint_interval"] = 0
    agent = DDPG(models=models, memory=memory, cfg=cfg, observation_space=obs_space, action_space=act_space, device="cpu")
    agent.init({"timesteps": timesteps})
    return agent, memory, n_envs


def run_case(idx, tag, obs, act, out_act, batch, base_scale, clip_low, T_e, lr_a, lr_c, discount, polyak, rows, n_act):
    from oracle.agent_oracle import AgentOracle, flat_params, synth_transitions

    seed = 4321 + idx
    torch.manual_seed(seed)
    agent, memory, n_envs = build_reference_agent(obs, act, out_act, batch, base_scale, clip_low, T_e, lr_a, lr_c,
                                                  discount, polyak, rows)
    torch.manual_seed(seed)
    oracle = AgentOracle(obs, act, out_act, batch_size=batch, discount=discount, polyak=polyak, actor_lr=lr_a,
                         critic_lr=lr_c, ou=(0.15, 0.1, base_scale), final_scale=1e-4, exploration_timesteps=T_e,
                         random_timesteps=10, clip=(clip_low, 1.0))
    assert torch.equal(flat_params(agent.policy), flat_params(oracle.policy)), "seeded init differs"
    assert torch.equal(flat_params(agent.critic), flat_params(oracle.critic))
    rec = dict(seed=seed, obs_dim=obs, act_dim=act, out_act=out_act, batch=batch, base_scale=base_scale,
               clip_low=clip_low, T_e=T_e, lr_a=lr_a, lr_c=lr_c, discount=discount, polyak=polyak, rows=rows,
               n_envs=n_envs, stride=STRIDE,
               init_policy_sum=flat_params(agent.policy).double().sum().item(),
               init_critic_sum=flat_params(agent.critic).double().sum().item())

    gen = torch.Generator().manual_seed(seed)
    data = synth_transitions(rows, obs, act, gen)
    for t in range(rows // n_envs):
        sl = slice(t * n_envs, (t + 1) * n_envs)
        memory.add_samples(**{k: v[sl] for k, v in data.items()})
    assert memory.filled and len(memory) == rows
    for k, v in data.items():
        rec[f"mem_{k}"] = v.numpy().astype(np.uint8) if v.dtype == torch.bool else v.numpy()

    # ---- random phase (timestep < random_timesteps): supplied uniforms
    states = torch.randn(n_act, obs, generator=gen)
    u = torch.rand(n_act, act, generator=gen)

    class FakeUniform:
        def sample(self, sample_shape):
            assert tuple(sample_shape) == (n_act, act)
            return u * (torch.tensor(1.0) - torch.tensor(clip_low)) + torch.tensor(clip_low)

    agent.policy._random_distribution = FakeUniform()
    ra = agent.act(states, 3, 1000)[0]
    ro = oracle.act(states, 3, 1000, uniforms=u)
    assert torch.equal(ra, ro)
    rec.update(rand_states=states.numpy(), rand_uniforms=u.numpy(), rand_actions=ra.numpy())

    # ---- interleaved act (OU noise, supplied normals) and _update (supplied permutation)
    n_rounds = 3
    orig_randperm = torch.randperm
    worst = {}
    for k in range(n_rounds):
        timestep = 100 + 7 * k
        states = torch.randn(n_act, obs, generator=gen) * 1.5
        z = torch.randn(n_act, act, generator=gen)

        class FakeNormal:
            def sample(self, size):
                assert tuple(size) == (n_act, act)
                return z.clone()

        agent._exploration_noise.distribution = FakeNormal()
        with torch.no_grad():
            a_ref = agent.act(states, timestep, 1000)[0].clone()
        a_or = oracle.act(states, timestep, 1000, normals=z)
        d = (a_ref - a_or).abs().max().item()
        worst["act"] = max(worst.get("act", 0), d)
        assert d <= 1e-6, d
        rec[f"r{k}_timestep"] = timestep
        rec[f"r{k}_states"] = states.numpy()
        rec[f"r{k}_normals"] = z.numpy()
        rec[f"r{k}_actions"] = a_ref.numpy()
        rec[f"r{k}_ou_state"] = agent._exploration_noise.state.numpy().copy()

        perm = orig_randperm(rows, generator=gen)
        torch.randperm = lambda n, **kw: perm.clone()
        try:
            agent.tracking_data.clear()
            agent._update(timestep, 1000)
        finally:
            torch.randperm = orig_randperm
        idx_ = perm[:batch]
        st = oracle.update(*[data[n][idx_] for n in ("states", "actions", "rewards", "next_states", "terminated", "truncated")])
        td = agent.tracking_data
        ref_stats = dict(critic_loss=td["Loss / Critic loss"][-1], policy_loss=td["Loss / Policy loss"][-1],
                         q_max=td["Q-network / Q1 (max)"][-1], q_min=td["Q-network / Q1 (min)"][-1],
                         q_mean=td["Q-network / Q1 (mean)"][-1], y_max=td["Target / Target (max)"][-1],
                         y_min=td["Target / Target (min)"][-1], y_mean=td["Target / Target (mean)"][-1])
        for name, v in ref_stats.items():
            assert abs(v - st[name]) <= 1e-6 * max(1.0, abs(v)), (name, v, st[name])
        rec[f"r{k}_indices"] = idx_.numpy()
        rec[f"r{k}_stats"] = np.array([ref_stats[n] for n in ("critic_loss", "policy_loss", "q_max", "q_min", "q_mean",
                                                              "y_max", "y_min", "y_mean")], dtype=np.float64)
        for name, net_r, net_o in (("policy", agent.policy, oracle.policy), ("critic", agent.critic, oracle.critic),
                                   ("target_policy", agent.target_policy, oracle.target_policy),
                                   ("target_critic", agent.target_critic, oracle.target_critic)):
            fr, fo = flat_params(net_r), flat_params(net_o)
            dd = (fr - fo).abs().max().item()
            worst[name] = max(worst.get(name, 0), dd)
            assert dd <= 1e-6, (name, dd)
            rec[f"r{k}_{name}_sub"] = subsample(fr)
            rec[f"r{k}_{name}_sum"] = fr.double().sum().item()
            rec[f"r{k}_{name}_abs_sum"] = fr.double().abs().sum().item()
        sc_r, sc_o = agent._state_preprocessor, oracle.scaler
        assert torch.allclose(sc_r.running_mean, sc_o.running_mean, rtol=0, atol=1e-12)
        rec[f"r{k}_scaler_mean"] = sc_r.running_mean.numpy().copy()
        rec[f"r{k}_scaler_var"] = sc_r.running_variance.numpy().copy()
        rec[f"r{k}_scaler_count"] = float(sc_r.current_count)
        # Adam moments (subsampled) of the policy optimizer
        m = torch.cat([agent.policy_optimizer.state[p]["exp_avg"].reshape(-1) for p in agent.policy.parameters()])
        v = torch.cat([agent.policy_optimizer.state[p]["exp_avg_sq"].reshape(-1) for p in agent.policy.parameters()])
        rec[f"r{k}_policy_m_sub"], rec[f"r{k}_policy_v_sub"] = subsample(m), subsample(v)
    rec["rounds"] = n_rounds
    print(f"  agent {tag:16s} worst |ref - restatement|: " + " ".join(f"{k}={v:.1e}" for k, v in worst.items()))
    return rec


def main():
    os.makedirs(GOLDEN, exist_ok=True)
    print("agent cases (reference my_agents/ddpg.py vs oracle/agent_oracle.py):")
    for i, case in enumerate(CASES):
        rec = run_case(i, *case)
        np.savez_compressed(os.path.join(GOLDEN, f"agent_{case[0]}.npz"), **rec)


if __name__ == "__main__":
    main()
