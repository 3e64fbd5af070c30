"""TEST INFRASTRUCTURE ONLY — CPU torch (fp32) restatement of the reference's DDPG agent arithmetic:
``DDPG.act`` (src/my_agents/ddpg.py:215-268) and ``DDPG._update`` (:357-444) with the models of
train/train_kamma_4d_nofifo_random_RLIL.py:21-52 and the skrl pieces restated in
oracle/ref_harness/skrl_shim.py (skrl itself is unpinned and absent: "restated, unverifiable offline").

Only ``tests/``, ``__graft_entry__.smoke()`` and ``bench.py``'s cpu_baseline / ``--impl reference`` legs may
import this module.  Pinning: ``oracle/gen_golden_agent.py`` runs the UNMODIFIED ``my_agents/ddpg.py``
through the harness on the same inputs, asserts this restatement agrees, and writes
``tests/golden/agent_*.npz``; ``tests/test_agent_oracle_golden.py`` replays them.
"""
from __future__ import annotations

import numpy as np
import torch
import torch.nn as nn
import torch.nn.functional as Fn

from .ref_harness.skrl_shim import OrnsteinUhlenbeckNoise, RunningStandardScaler


def make_net(in_dim: int, out_dim: int, out_act: str) -> nn.Sequential:
    """nn.Sequential of train/train_kamma_4d_nofifo_random_RLIL.py:26-33 (actor) / :43-49 (critic)."""
    layers = [nn.Linear(in_dim, 512), nn.ELU(), nn.Linear(512, 512), nn.ELU(), nn.Linear(512, 128), nn.ELU(),
              nn.Linear(128, out_dim)]
    if out_act == "sigmoid":
        layers.append(nn.Sigmoid())
    elif out_act == "tanh":
        layers.append(nn.Tanh())
    return nn.Sequential(*layers)


def flat_params(net: nn.Module) -> torch.Tensor:
    return torch.cat([p.detach().reshape(-1) for p in net.parameters()])


def load_flat(net: nn.Module, flat) -> None:
    flat = torch.as_tensor(np.asarray(flat), dtype=torch.float32)
    off = 0
    with torch.no_grad():
        for p in net.parameters():
            p.copy_(flat[off:off + p.numel()].view_as(p))
            off += p.numel()


class AgentOracle:
    def __init__(self, obs_dim=6, act_dim=4, out_act="sigmoid", *, batch_size=512, discount=0.999, polyak=0.005,
                 actor_lr=3e-4, critic_lr=1e-3, ou=(0.15, 0.1, 0.05), initial_scale=1.0, final_scale=1e-4,
                 exploration_timesteps=5e5, random_timesteps=0, clip=(0.0, 1.0), use_scaler=True):
        self.obs_dim, self.act_dim = obs_dim, act_dim
        # construction order of the train scripts: policy, target_policy, critic, target_critic (…RLIL.py:110-113)
        self.policy = make_net(obs_dim, act_dim, out_act)
        self.target_policy = make_net(obs_dim, act_dim, out_act)
        self.critic = make_net(obs_dim + act_dim, 1, "none")
        self.target_critic = make_net(obs_dim + act_dim, 1, "none")
        self.target_policy.load_state_dict(self.policy.state_dict())      # ddpg.py:135-136
        self.target_critic.load_state_dict(self.critic.state_dict())
        for p in list(self.target_policy.parameters()) + list(self.target_critic.parameters()):
            p.requires_grad = False
        self.policy_optimizer = torch.optim.Adam(self.policy.parameters(), lr=actor_lr)     # :174-175
        self.critic_optimizer = torch.optim.Adam(self.critic.parameters(), lr=critic_lr)
        self.scaler = RunningStandardScaler(obs_dim) if use_scaler else None
        self.noise = OrnsteinUhlenbeckNoise(*ou) if ou is not None else None
        self.batch_size, self.discount, self.polyak = batch_size, discount, polyak
        self.initial_scale, self.final_scale = initial_scale, final_scale
        self.exploration_timesteps, self.random_timesteps = exploration_timesteps, random_timesteps
        self.clip = clip

    def _pre(self, x, train=False):
        return self.scaler(x, train=train) if self.scaler is not None else x

    # ddpg.py:215-268
    def act(self, states, timestep, timesteps, normals=None, uniforms=None):
        states = torch.as_tensor(np.asarray(states), dtype=torch.float32)
        lo, hi = self.clip
        if timestep < self.random_timesteps:   # Model.random_act: Uniform(low, high).sample()
            u = torch.as_tensor(np.asarray(uniforms), dtype=torch.float32)
            return u * (torch.tensor(hi) - torch.tensor(lo)) + torch.tensor(lo)
        with torch.no_grad():
            actions = self.policy(self._pre(states))
            if self.noise is not None:
                z = torch.as_tensor(np.asarray(normals), dtype=torch.float32)
                n = self.noise
                if hasattr(n.state, "shape") and n.state.shape != z.shape:
                    n.state = 0
                n.state = n.state + (-n.state * n.theta + n.sigma * z)
                noises = n.base_scale * n.state
                if self.exploration_timesteps is None:
                    self.exploration_timesteps = timesteps
                if timestep <= self.exploration_timesteps:
                    scale = (1 - timestep / self.exploration_timesteps) * (self.initial_scale - self.final_scale) + self.final_scale
                    noises = noises * scale
                    actions = (actions + noises).clamp(min=lo, max=hi)
                    self.last_noises = noises
        return actions

    # ddpg.py:357-444 for one gradient step on an explicit batch
    def update(self, states, actions, rewards, next_states, terminated, truncated):
        t = lambda a, dt=torch.float32: torch.as_tensor(np.asarray(a)).to(dt)
        s, a, r, s2 = t(states), t(actions), t(rewards).view(-1, 1), t(next_states)
        te, tr = t(terminated, torch.bool).view(-1, 1), t(truncated, torch.bool).view(-1, 1)
        s = self._pre(s, train=True)
        s2 = self._pre(s2, train=True)
        with torch.no_grad():
            a2 = self.target_policy(s2)
            q2 = self.target_critic(torch.cat([s2, a2], dim=1))
            y = r + self.discount * (te | tr).logical_not() * q2
        q = self.critic(torch.cat([s, a], dim=1))
        critic_loss = Fn.mse_loss(q, y)
        self.critic_optimizer.zero_grad()
        critic_loss.backward()
        # what critic.reduce_parameters() would all-reduce (ddpg.py:409-410): kept for the gradient-parity tests
        self.last_critic_grad = torch.cat([p.grad.detach().reshape(-1).clone() for p in self.critic.parameters()])
        self.critic_optimizer.step()
        pa = self.policy(s)
        qp = self.critic(torch.cat([s, pa], dim=1))
        policy_loss = -qp.mean()
        self.policy_optimizer.zero_grad()
        policy_loss.backward()
        self.last_actor_grad = torch.cat([p.grad.detach().reshape(-1).clone() for p in self.policy.parameters()])   # ddpg.py:431-432
        self.policy_optimizer.step()
        with torch.no_grad():   # Model.update_parameters(polyak)
            for net, tgt in ((self.policy, self.target_policy), (self.critic, self.target_critic)):
                for p, q_ in zip(tgt.parameters(), net.parameters()):
                    p.mul_(1 - self.polyak)
                    p.add_(self.polyak * q_)
        return dict(critic_loss=critic_loss.item(), policy_loss=policy_loss.item(), q_max=qp.max().item(),
                    q_min=qp.min().item(), q_mean=qp.mean().item(), y_max=y.max().item(), y_min=y.min().item(),
                    y_mean=y.mean().item())


def adam_moments(opt: torch.optim.Adam, net: nn.Module):
    """(exp_avg, exp_avg_sq, step) of a torch Adam flattened in parameter order (zeros before the first step)."""
    m, v, step = [], [], 0
    for p in net.parameters():
        st = opt.state.get(p, {})
        m.append(st["exp_avg"].reshape(-1) if "exp_avg" in st else torch.zeros(p.numel()))
        v.append(st["exp_avg_sq"].reshape(-1) if "exp_avg_sq" in st else torch.zeros(p.numel()))
        step = int(st["step"]) if "step" in st else step
    return torch.cat(m), torch.cat(v), step


def synth_transitions(rows: int, obs_dim: int, act_dim: int, gen: torch.Generator):
    """SURVEY.md §8(d) replay pre-fill: N(0,1) states, U actions, N(-5,5) rewards, 1 % terminated."""
    return dict(
        states=torch.randn(rows, obs_dim, generator=gen) * torch.tensor([1.5, 0.1, 0.3, 0.1, 2.0, 1.0][:obs_dim]),
        actions=torch.rand(rows, act_dim, generator=gen),
        rewards=torch.randn(rows, 1, generator=gen) * 5.0 - 5.0,
        next_states=torch.randn(rows, obs_dim, generator=gen) * torch.tensor([1.5, 0.1, 0.3, 0.1, 2.0, 1.0][:obs_dim]),
        terminated=torch.rand(rows, 1, generator=gen) < 0.01,
        truncated=torch.zeros(rows, 1, dtype=torch.bool),
    )
