"""``SequentialTrainer`` — the caller of the hot path.

Mirrors src/my_trainers/sequential.py:25-165 on the loop bodies of src/my_trainers/base_KA_DDPG.py:153-350
(``kind="KA_DDPG"``: KAMMA scripts and KA-DDPG-4d) and src/my_trainers/base_KA.py (``kind="KA"``: KA-DDPG
1d / 2d — identical except that it does not relabel the stored action, base_KA_DDPG.py:212).  Kept quirks:

  * the transition stores the EXECUTED action ``wheel_torqs / action_scale`` (KA_DDPG kind), not the policy's;
  * ``agent.post_interaction`` runs TWICE per timestep (base_KA_DDPG.py:265,283) — two gradient updates and
    two checkpoint / write checks at the same ``timestep``;
  * for ``num_envs > 1`` the env is never reset by the trainer (``states = next_states``);
  * eval calls the BASE-class ``post_interaction`` only (no ``_update``), base_KA_DDPG.py:334.

What is different by design: the reference reads 26 env-0 scalars (27 ``.item()`` host syncs) and every
``extras["episode"]`` value on every step; here that logging is strided (``log_interval``, default: the agent's
``write_interval``; 0 = off) because one host sync per step would cap the loop near 1e4 steps/s.
"""
from __future__ import annotations

import copy
import sys

import torch

from . import _lib as L
from . import parallel as P

SEQUENTIAL_TRAINER_DEFAULT_CONFIG = {   # src/my_trainers/sequential.py:14-20
    "timesteps": 100000,
    "headless": False,
    "disable_progressbar": False,
    "close_environment_at_exit": True,
    "environment_info": "episode",
    # extensions
    "kind": "auto",          # "KA_DDPG" (relabel) | "KA" (no relabel) | "auto": the env variant's pairing
    "log_interval": None,    # steps between env-0 / episode logging reads; None = agent.write_interval
    "pre_step_hook": None,   # callable(trainer, timestep) run at the top of every timestep (parity tests supply draws)
    "fuse_policy_step": True,  # eager loop: issue act + env.step as ONE kernel where the library can (rollout.fused_step;
                               # bit-identical to the two launches; needs agent.fuse_replay_with(task), headless)
    "graph_rollout": False,  # act + env.step (+ fused replay append) as ONE CUDA-graph launch per timestep once past
                             # random_timesteps (rollout.RolloutGraph; needs agent.fuse_replay_with(task), write_interval 0)
}

# observed_components key -> TensorBoard tag suffix (base_KA_DDPG.py:229-254)
_COMP_TAGS = (
    ("env0_desired_v", "desired_v"), ("env0_current_v", "current_v"), ("env0_desired_omega", "desired_omega"),
    ("env0_current_omega", "current_omega"), ("env0_v_delta", "v_delta"), ("env0_omega_delta", "omega_delta"),
    ("env0_linear_acc", "linear_acc"), ("env0_angular_acc", "angular_acc"), ("env0_episode_count", "episode_count"),
    ("env0_torque_apl_fl", "torque_fl"), ("env0_torque_apl_rl", "torque_rl"), ("env0_torque_apl_fr", "torque_fr"),
    ("env0_torque_apl_rr", "torque_rr"), ("env0_exp_left", "exp_torque_l"), ("env0_exp_right", "exp_torque_r"),
    ("env0_policy_torque_fl", "policy_torqs_fl"), ("env0_policy_torque_rl", "policy_torqs_rl"),
    ("env0_policy_torque_fr", "policy_torqs_fr"), ("env0_policy_torque_rr", "policy_torqs_rr"),
    ("env0_perc_r1", "env0_perc_r1"), ("env0_perc_r2", "env0_perc_r2"), ("env0_perc_r3", "env0_perc_r3"),
    ("env0_perc_dense", "env0_perc_dense"), ("env0_perc_sparse", "env0_perc_sparse"),
    ("env0_perc_observed", "env0_perc_observed"), ("env0_perc_guiding", "env0_perc_guiding"),
)


class SequentialTrainer:
    def __init__(self, env, agents, agents_scope=None, cfg=None):
        _cfg = copy.deepcopy(SEQUENTIAL_TRAINER_DEFAULT_CONFIG)
        _cfg.update(cfg if cfg is not None else {})
        self.cfg = _cfg
        self.env = env
        if isinstance(agents, (tuple, list)):
            if len(agents) != 1:
                raise L.OisError("simultaneous agents are not used by any reference script and are not implemented")
            agents = agents[0]
        self.agents = agents
        self.num_simultaneous_agents = 1
        self.timesteps = int(_cfg["timesteps"])
        self.headless = _cfg["headless"]
        self.disable_progressbar = _cfg["disable_progressbar"] or P.distributed_info().rank != 0
        self.environment_info = _cfg["environment_info"]
        self.initial_timestep = 0
        kind = _cfg["kind"]
        if kind == "auto":
            kind = "KA_DDPG" if env.task.variant.relabel == L.RELABEL_EXECUTED else "KA"
        if kind not in ("KA_DDPG", "KA"):
            raise L.OisError(f"unknown trainer kind {kind!r}")
        self.kind = kind
        self.agents.init(trainer_cfg=self.cfg)   # sequential.py:52-55
        li = _cfg["log_interval"]
        self.log_interval = int(self.agents.write_interval if li is None else li)
        if self.log_interval > 0:
            env.task.stats_interval = self.log_interval

    # ------------------------------------------------------------------ helpers
    def _progress(self):
        rng = range(self.initial_timestep, self.timesteps)
        if self.disable_progressbar:
            return rng
        try:
            import tqdm
            return tqdm.tqdm(rng, file=sys.stdout)
        except ImportError:
            return rng

    def _log_env(self, infos):
        task, agent = self.env.task, self.agents
        comp = task.observed_components
        for key, tag in _COMP_TAGS:
            agent.track_data(f"Reward_comp_env0 / {tag}", comp[key])
        if self.environment_info in infos:
            for k, v in infos[self.environment_info].items():
                if isinstance(v, torch.Tensor) and v.numel() == 1:
                    agent.track_data(f"Info / {k}", v.item())

    # ------------------------------------------------------------------ sequential.py:57-85 -> base_KA_DDPG.py:153-293
    def train(self) -> None:
        agent, env = self.agents, self.env
        agent.set_running_mode("train")
        states, infos = env.reset()
        hook = self.cfg["pre_step_hook"]
        graph = None
        if self.cfg["graph_rollout"]:
            from .rollout import RolloutGraph
            graph = RolloutGraph(env, agent, max(self.initial_timestep, agent._random_timesteps), self.timesteps)
        from .rollout import fused_step
        fuse_step = bool(self.cfg["fuse_policy_step"] and agent._fused_replay and not agent.secondary_memories and self.headless)
        for timestep in self._progress():
            if hook is not None:
                hook(self, timestep)
            agent.pre_interaction(timestep=timestep, timesteps=self.timesteps)
            if graph is not None and timestep >= agent._random_timesteps:
                graph.timestep = timestep       # act + step + record_transition: one graph launch
                next_states, rewards, terminated = graph.step()
                truncated, infos = env._truncated, env.task.extras
                agent.post_interaction(timestep=timestep, timesteps=self.timesteps)     # base_KA_DDPG.py:265
                states = self._next_states(next_states, terminated, truncated)
                agent.post_interaction(timestep=timestep, timesteps=self.timesteps)     # base_KA_DDPG.py:283 (again)
                continue
            with torch.no_grad():
                if fuse_step:    # act + env.step (+ the fused replay append) as one kernel launch; same results
                    next_states, rewards, terminated, truncated, infos = fused_step(env, agent, states, timestep, self.timesteps)
                else:
                    actions = agent.act(states, timestep=timestep, timesteps=self.timesteps)[0]
                    next_states, rewards, terminated, truncated, infos = env.step(actions)
                    if self.kind == "KA_DDPG":   # base_KA_DDPG.py:212
                        actions = None if agent._fused_replay else env.task.wheel_torqs / env.task.action_scale
                    if not self.headless:
                        env.render()
                    agent.record_transition(states=states, actions=actions, rewards=rewards, next_states=next_states,
                                            terminated=terminated, truncated=truncated, infos=infos, timestep=timestep,
                                            timesteps=self.timesteps)
                if self.log_interval > 0 and timestep % self.log_interval == 0:
                    self._log_env(infos)
            agent.post_interaction(timestep=timestep, timesteps=self.timesteps)     # base_KA_DDPG.py:265
            states = self._next_states(next_states, terminated, truncated)
            agent.post_interaction(timestep=timestep, timesteps=self.timesteps)     # base_KA_DDPG.py:283 (again)
            states = self._next_states(states, terminated, truncated, already=True)

    def _next_states(self, next_states, terminated, truncated, already=False):
        """base_KA_DDPG.py:267-275: vectorised envs reset themselves; a single env is reset by the trainer."""
        if self.env.num_envs > 1 or already:
            return next_states
        if bool(terminated.any()) or bool(truncated.any()):
            with torch.no_grad():
                return self.env.reset()[0]
        return next_states

    # ------------------------------------------------------------------ sequential.py:140-165 -> base_KA_DDPG.py:295-350
    def eval(self) -> None:
        agent, env = self.agents, self.env
        agent.set_running_mode("eval")
        states, infos = env.reset()
        hook = self.cfg["pre_step_hook"]
        for timestep in self._progress():
            if hook is not None:
                hook(self, timestep)
            with torch.no_grad():
                actions = agent.act(states, timestep=timestep, timesteps=self.timesteps)[0]
                next_states, rewards, terminated, truncated, infos = env.step(actions)
                if not self.headless:
                    env.render()
                agent.record_transition(states=states, actions=actions, rewards=rewards, next_states=next_states,
                                        terminated=terminated, truncated=truncated, infos=infos, timestep=timestep,
                                        timesteps=self.timesteps)
                agent.base_post_interaction(timestep=timestep, timesteps=self.timesteps)   # no _update
                if self.log_interval > 0 and timestep % self.log_interval == 0 and self.environment_info in infos:
                    for k, v in infos[self.environment_info].items():
                        if isinstance(v, torch.Tensor) and v.numel() == 1:
                            agent.track_data(f"Info / {k}", v.item())
            states = self._next_states(next_states, terminated, truncated)
