"""TEST INFRASTRUCTURE ONLY — restatement of the reference's training loop body
(src/my_trainers/base_KA_DDPG.py:153-293 / base_KA.py) over ``EnvOracle`` + ``AgentOracle`` + a numpy ring
buffer with skrl ``Memory`` index semantics (train/train_kaddpg_1d_nofifo_random_RLIL.py:147-205).

Pinned by ``oracle/gen_golden_trainer.py`` against the UNMODIFIED reference trainer + task + agent
(tests/golden/trainer_*.npz); replayed by tests/test_trainer_oracle_golden.py.  Only tests, smoke() and bench's
CPU legs may import this module.
"""
from __future__ import annotations

import numpy as np

from .agent_oracle import AgentOracle
from .env_oracle import EnvOracle, EnvParams

F = np.float32
NAMES = ("states", "actions", "rewards", "next_states", "terminated", "truncated")


class TrainerOracle:
    def __init__(self, variant, state, *, n, batch, mem_steps, random_timesteps, learning_starts, ou_scale, T_e, clip_low,
                 relabel: bool, clip_obs: float = 1000.0):
        self.variant, self.n = variant, n
        self.env = EnvOracle(variant, state, EnvParams(physics=True))
        self.agent = AgentOracle(6, variant.act_dim, "sigmoid", batch_size=batch, ou=(0.15, 0.1, ou_scale),
                                 final_scale=1e-4, exploration_timesteps=T_e, random_timesteps=random_timesteps,
                                 clip=(clip_low, 1.0))
        self.learning_starts, self.relabel, self.clip_obs = learning_starts, relabel, F(clip_obs)
        # the trainer stores what it passes: relabelled actions are 4 wide, policy actions act_dim wide — but the
        # memory tensor is created act_dim wide (ddpg.py:203), which only matches because every relabelling script
        # uses a 4-d action space
        dims = dict(states=6, actions=variant.act_dim, rewards=1, next_states=6, terminated=1, truncated=1)
        self.memory = {k: (np.full((mem_steps * n, d), np.nan, F) if k not in ("terminated", "truncated")
                           else np.zeros((mem_steps * n, d), bool)) for k, d in dims.items()}
        self.mem_steps, self.memory_index, self.filled = mem_steps, 0, False
        self.log = []

    def _add(self, **rows):   # Memory.add_samples, multi-env branch
        sl = slice(self.memory_index * self.n, (self.memory_index + 1) * self.n)
        for k, v in rows.items():
            self.memory[k][sl] = np.asarray(v).reshape(self.n, -1)
        self.memory_index += 1
        if self.memory_index >= self.mem_steps:
            self.memory_index, self.filled = 0, True

    def run(self, timesteps: int, draws: dict, indexes):
        env, agent, n = self.env, self.agent, self.n
        switch = self.variant.blend == "switch"
        clip = lambda o: np.clip(o, -self.clip_obs, self.clip_obs)
        k = 0       # env-step counter (0 = the wrapper's reset step with zero actions)
        obs, _, _ = env.step(np.zeros((n, self.variant.act_dim), F), draws["rand_mask"][k] if switch else None,
                             draws["reset_uniforms"][k])
        states = clip(obs)
        k += 1
        u = 0       # update counter
        for t in range(timesteps):
            a = agent.act(states, t, timesteps, normals=draws["normals"][t], uniforms=draws["uniforms"][t]).numpy()
            obs, rew, reset = env.step(a, draws["rand_mask"][k] if switch else None, draws["reset_uniforms"][k])
            k += 1
            next_states = clip(obs)
            wt = env.last["wheel_torqs"]
            stored = (wt / F(env.p.action_scale)).astype(F) if self.relabel else a      # base_KA_DDPG.py:212
            self._add(states=states, actions=stored, rewards=rew, next_states=next_states, terminated=reset,
                      truncated=np.zeros(n, bool))
            self.log.append(dict(actions=a, obs=next_states, rew=rew, reset=reset, wheel_torqs=wt.copy()))
            for _ in range(2):                                                          # :265 and :283
                if t >= self.learning_starts:
                    idx = np.asarray(indexes[u])
                    u += 1
                    agent.update(*[self.memory[name][idx] for name in NAMES])
            states = next_states
        self.updates = u
