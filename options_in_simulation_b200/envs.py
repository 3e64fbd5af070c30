"""``TorqueDistributionTask`` — B200-native mirror of the reference's RLTask subclass
(src/my_envs/KAMMA_medium.py:148 and its variants), same method names and buffers:

    pre_physics_step(actions) / post_physics_step() -> (obs_buf, rew_buf, reset_buf, extras)
    get_observations() / calculate_metrics() / is_done() / reset_idx(env_ids) / post_reset()

All per-env state lives in HBM as [N,4] float tensors + int64 counters (``OisEnvState``) and one fused
CUDA kernel (csrc/env_step.cu) does the whole policy step; PhysX is replaced by the skid-steer model.
In the reference the physics already runs inside ``pre_physics_step`` (KAMMA_medium.py:646-653); here the
fused launch also happens there and ``post_physics_step`` returns the buffers it produced, so the
sequence VecEnvRLGames.step drives (pre -> post) is observationally the same.
"""
from __future__ import annotations

import ctypes as C
import math

import numpy as np
import torch

from . import _lib as L
from .config import VARIANTS, EnvVariant, env_params, task_cfg

SUM_KEYS = (  # KAMMA_medium.py:205-213
    "r1: Tracking error reward (squared errors)",
    "r2: Convergence reward (squared accelerations)",
    "r3: Torque penalty (sum of squared torques)",
    "Dense reward",
    "Sparse reward",
    "Guiding reward",
    "Observed reward",
    "Final reward",
)
EVAL_SUM_KEYS = ("Smoothness", "Tracking error", "Desired velocity")   # KAMMA_eval.py:224-226
RATIO_KEYS = ("r1/Final reward", "r2/Final reward", "r3/Final reward", "Dense/Final reward",
              "Sparse/Final reward", "Guiding/Final reward", "Observed/Final reward")   # KAMMA_medium.py:470-476


class Box:
    """Minimal gym.spaces.Box stand-in (the reference only reads shape / low / high)."""

    def __init__(self, low, high, shape, dtype=np.float32):
        self.shape = tuple(shape)
        self.dtype = dtype
        self.low = np.full(self.shape, low, dtype=dtype)
        self.high = np.full(self.shape, high, dtype=dtype)


class _SimConfig:
    def __init__(self, cfg):
        self.config = cfg
        self.task_config = cfg["task"]


class TorqueDistributionTask:
    def __init__(self, name="TorqueDistribution", sim_config=None, env=None, offset=None, *,
                 variant: str | EnvVariant = "KAMMA_medium", physics: bool = True, seed: int | None = None,
                 gamma_const: float | None = None):
        L.lib()   # fail loudly, before any state is built, if the CUDA library is absent
        if not torch.cuda.is_available():
            raise L.OisError("TorqueDistributionTask needs a CUDA device (B200, sm_100a); there is no CPU path")
        self.variant = VARIANTS[variant] if isinstance(variant, str) else variant
        if sim_config is None:
            sim_config = _SimConfig(task_cfg())
        elif isinstance(sim_config, dict):
            sim_config = _SimConfig(sim_config)
        self._name = name
        self._env = env
        self._sim_config = sim_config
        self._cfg = sim_config.config
        self._task_cfg = sim_config.task_config
        dev = self._cfg.get("sim_device", "cuda:0")
        self._device = torch.device("cuda:0" if dev in ("gpu", "cuda") else dev)
        self._num_envs = int(self._task_cfg["env"]["numEnvs"])
        self._num_observations = 6
        self._num_actions = self.variant.act_dim
        self._num_states = 0
        self._num_agents = 1
        self.observation_space = Box(-1.0, 1.0, (self._num_observations,))          # KAMMA_medium.py:158-163
        self.action_space = Box(self.variant.action_low, 1.0, (self._num_actions,))  # :164-169
        self.clip_obs = float(self._task_cfg["env"].get("clipObservations", math.inf))   # rl_task.py:99
        self.clip_actions = float(self._task_cfg["env"].get("clipActions", math.inf))
        self._params = env_params(self.variant, self._cfg, physics=physics, gamma_const=gamma_const)
        self.action_scale = self._params.action_scale
        self.max_episode_length = self._params.max_episode_length
        self.max_episode_length_s = self._params.max_episode_length_s
        self.max_sim_steps = self._params.max_sim_steps
        self.dt = self._params.dt
        self.decimation = self._params.decimation
        self._seed = int(self._cfg.get("seed", 42) if seed is None else seed)
        self._launches = 0          # policy steps issued: Philox offset and eval flag slot
        self.kernel_launches = 0    # CUDA kernels launched by this object (bench "gpu_launches")
        self.stats_interval = 0     # > 0: refresh extras["episode"] every that many steps (one small D2H)
        # device-side statistics log (evaluate.tracking_error_sweep): a zeroed [T, OIS_EP_STATS_LEN] float tensor; step k
        # reduces its episode statistics into row k and the host reads the whole log ONCE at the end (no per-step sync)
        self.stats_log = None
        self._stats_row = 0
        self._io, self._io_key, self._io_static = None, None, None      # cached OisEnvIO block (pre_physics_step)
        # ``torques`` (pre-clip), ``guiding_reward`` and ``timeout_buf`` are attributes of the reference task that nothing on
        # the rollout path reads (the trainer takes ``wheel_torqs``; the wrapper reports truncated = 0).  True keeps them
        # refreshed every step like the reference; False leaves them stale and saves their 28 B of writes per env-step.
        self.intermediate_outputs = True
        self._replay_key, self._replay_rows = None, None
        self.extras: dict = {}
        self._supplied_mask = None
        self._supplied_reset = None
        self._replay = None
        self._alloc()
        self.init_done = False

    # ------------------------------------------------------------------ RLTaskInterface properties
    @property
    def num_envs(self):
        return self._num_envs

    @property
    def device(self):
        return self._device

    @property
    def num_observations(self):
        return self._num_observations

    @property
    def num_actions(self):
        return self._num_actions

    @property
    def num_states(self):
        return self._num_states

    @property
    def name(self):
        return self._name

    # ------------------------------------------------------------------ buffers (RLTask.cleanup, rl_task.py:145-154)
    def _alloc(self):
        n, dev = self._num_envs, self._device
        z4 = lambda: torch.zeros(n, 4, dtype=torch.float32, device=dev)
        zl = lambda: torch.zeros(n, dtype=torch.int64, device=dev)
        self._pose_vx, self._twist_prev = z4(), z4()
        self.dof_vel = z4()
        self.commands = z4()                                  # x vel, y vel, yaw vel, heading  :378-380
        self._sums_a, self._sums_b = z4(), z4()
        ev = self.variant.eval_sums
        self._sums_eval = z4() if ev else None
        self.old_torques = z4() if ev else None
        self._any_reset_flags = torch.zeros(3, dtype=torch.int32, device=dev) if ev else None
        self.sim_steps, self.episode_buf, self.episode_count = zl(), zl(), zl()
        self.obs_buf = torch.zeros(n, 6, dtype=torch.float32, device=dev)
        # clamp(obs_buf, +-clipObservations) as the env wrapper returns it; two buffers alternate so the previous
        # step's observations (the trainer's `states`) stay intact while the next step is written
        # what a step hands back — clamped observations (buffer 0), rewards, reset flags — is ONE allocation
        # [obs N*24 B | rew N*4 B | reset N B], so that a host in the loop gets it with a single device->host copy
        # (rollout.RolloutGraph); the tensors below are views into it
        self._out_pack = torch.zeros(n * 29, dtype=torch.uint8, device=dev)
        self._obs_clipped = [self._out_pack[: n * 24].view(torch.float32).view(n, 6),
                             torch.zeros(n, 6, dtype=torch.float32, device=dev)]
        self.obs_clipped = self._obs_clipped[0]
        self.rew_buf = self._out_pack[n * 24: n * 28].view(torch.float32)
        self.reset_buf = self._out_pack[n * 28:].view(torch.bool)
        self.reset_buf.fill_(True)
        self.timeout_buf = zl()
        self.progress_buf = zl()
        self.states_buf = torch.zeros(n, 0, dtype=torch.float32, device=dev)
        self.torques, self.wheel_torqs = z4(), z4()
        self.actions = torch.zeros(n, self._num_actions, dtype=torch.float32, device=dev)
        self.guiding_reward = torch.zeros(n, dtype=torch.float32, device=dev)
        self._ep_stats = torch.zeros(L.OIS_EP_STATS_LEN, dtype=torch.float32, device=dev)
        self._comp0 = torch.zeros(L.OIS_COMP0_LEN, dtype=torch.float32, device=dev)
        keys = SUM_KEYS + (EVAL_SUM_KEYS if ev else ())
        views = [self._sums_a[:, k] for k in range(4)] + [self._sums_b[:, k] for k in range(4)]
        if ev:
            views += [self._sums_eval[:, k] for k in range(3)]
        self.episode_sums = dict(zip(keys, views))            # strided views of the [N,4] blocks
        st = L.EnvState()
        st.pose_vx, st.twist_prev = self._pose_vx.data_ptr(), self._twist_prev.data_ptr()
        st.wheel_speed, st.commands = self.dof_vel.data_ptr(), self.commands.data_ptr()
        st.sums_a, st.sums_b = self._sums_a.data_ptr(), self._sums_b.data_ptr()
        st.sums_eval = self._sums_eval.data_ptr() if ev else None
        st.old_torques = self.old_torques.data_ptr() if ev else None
        st.sim_steps, st.episode_buf = self.sim_steps.data_ptr(), self.episode_buf.data_ptr()
        st.episode_count = self.episode_count.data_ptr()
        self._state = st

    # ------------------------------------------------------------------ reference-named views of the packed state
    @property
    def base_pos(self):
        z = torch.full((self._num_envs, 1), float(self._task_cfg["env"]["baseInitState"]["pos"][2]), device=self._device)
        return torch.cat([self._pose_vx[:, 0:2], z], dim=1)

    @property
    def base_quat(self):   # wxyz, rotation about z
        h = self._pose_vx[:, 2] * 0.5
        zero = torch.zeros_like(h)
        return torch.stack([torch.cos(h), zero, zero, torch.sin(h)], dim=1)

    @property
    def base_velocities(self):   # world-frame (lin xyz, ang xyz), as ArticulationView.get_velocities serves them
        yaw, vx, vy, wz = self._pose_vx[:, 2], self._pose_vx[:, 3], self._twist_prev[:, 0], self._twist_prev[:, 1]
        c, s = torch.cos(yaw), torch.sin(yaw)
        zero = torch.zeros_like(vx)
        return torch.stack([c * vx - s * vy, s * vx + c * vy, zero, zero, zero, wz], dim=1)

    @property
    def v_forward_projected(self):
        return self._pose_vx[:, 3]

    @property
    def v_lateral_projected(self):
        return self._twist_prev[:, 0]

    @property
    def previous_linear_velocity(self):
        z = torch.zeros(self._num_envs, 3, device=self._device)
        z[:, 0] = self._twist_prev[:, 2]
        return z

    @property
    def previous_angular_velocity(self):
        z = torch.zeros(self._num_envs, 3, device=self._device)
        z[:, 2] = self._twist_prev[:, 3]
        return z

    desired_v = property(lambda self: self.commands[:, 0])
    desired_omega = property(lambda self: self.commands[:, 2])
    v_delta = property(lambda self: self.obs_buf[:, 2])
    omega_delta = property(lambda self: self.obs_buf[:, 3])
    linear_acc = property(lambda self: self.obs_buf[:, 4])
    angular_acc = property(lambda self: self.obs_buf[:, 5])

    def _gamma(self, mode, const):
        if mode == L.GAMMA_DECAY:   # KAMMA_medium.py:627-628
            return torch.clamp(1.0 - (self.sim_steps.float() / self.max_sim_steps), min=0.0)
        return torch.full((self._num_envs,), const, device=self._device)

    gamma_assist1 = property(lambda self: self._gamma(self._params.gamma1_mode, self._params.gamma1_const))
    gamma_assist2 = property(lambda self: self._gamma(self._params.gamma2_mode, self._params.gamma2_const))
    gamma_assist = gamma_assist1

    # ------------------------------------------------------------------ state load / dump (tests, checkpoints)
    def load_state(self, *, px=None, py=None, yaw=None, vx=None, vy=None, wz=None, w=None, prev_lin=None,
                   prev_ang=None, commands=None, sim_steps=None, episode_buf=None, episode_count=None, sums=None,
                   old_torques=None):
        def put(dst, src):
            if src is not None:
                dst.copy_(torch.as_tensor(np.asarray(src)).to(dst.dtype))

        put(self._pose_vx[:, 0], px); put(self._pose_vx[:, 1], py); put(self._pose_vx[:, 2], yaw)
        put(self._pose_vx[:, 3], vx); put(self._twist_prev[:, 0], vy); put(self._twist_prev[:, 1], wz)
        put(self._twist_prev[:, 2], prev_lin); put(self._twist_prev[:, 3], prev_ang)
        put(self.dof_vel, w); put(self.commands, commands)
        put(self.sim_steps, sim_steps); put(self.episode_buf, episode_buf); put(self.episode_count, episode_count)
        if sums is not None:
            sums = np.asarray(sums)
            put(self._sums_a, sums[:, 0:4]); put(self._sums_b, sums[:, 4:8])
            if self.variant.eval_sums:
                self._sums_eval[:, 0:3].copy_(torch.as_tensor(sums[:, 8:11]))
        if old_torques is not None and self.variant.eval_sums:
            put(self.old_torques, old_torques)

    def dump_state(self) -> dict:
        c = lambda t: t.detach().cpu().numpy().copy()
        sums = [self._sums_a, self._sums_b] + ([self._sums_eval[:, 0:3]] if self.variant.eval_sums else [])
        return dict(px=c(self._pose_vx[:, 0]), py=c(self._pose_vx[:, 1]), yaw=c(self._pose_vx[:, 2]),
                    vx=c(self._pose_vx[:, 3]), vy=c(self._twist_prev[:, 0]), wz=c(self._twist_prev[:, 1]),
                    w=c(self.dof_vel), prev_lin=c(self._twist_prev[:, 2]), prev_ang=c(self._twist_prev[:, 3]),
                    commands=c(self.commands), sim_steps=c(self.sim_steps), episode_buf=c(self.episode_buf),
                    episode_count=c(self.episode_count), sums=c(torch.cat(sums, dim=1)))

    # ------------------------------------------------------------------ RNG: supplied draws (parity) or Philox
    def supply_draws(self, rand_mask=None, reset_uniforms=None):
        """Parity mode: the uniforms the reference would draw (KAMMA_medium.py:633 and reset_idx) are
        inputs.  ``None`` returns that stream to the in-kernel Philox generator."""
        dev = self._device
        as_dev = lambda a: None if a is None else torch.as_tensor(np.asarray(a, dtype=np.float32)).to(dev).contiguous()
        self._supplied_mask = as_dev(rand_mask)
        self._supplied_reset = as_dev(reset_uniforms)
        if self._supplied_reset is not None:
            assert tuple(self._supplied_reset.shape) == (self._num_envs, 8)

    def bind_replay(self, memory, relabel: int | None = None):
        """Fuse ``DDPG.record_transition``'s memory.add_samples (ddpg.py:313) into the env kernel."""
        self._replay = memory
        self._relabel = self.variant.relabel if relabel is None else relabel

    # ------------------------------------------------------------------ RLTask API
    def post_reset(self):
        """KAMMA_medium.py:366-414: reset every env once."""
        ids = torch.arange(self._num_envs, dtype=torch.int64, device=self._device)
        self.reset_idx(ids)
        self.init_done = True

    def reset_idx(self, env_ids):
        """KAMMA_medium.py:417-498 for an explicit index list."""
        env_ids = env_ids.to(device=self._device, dtype=torch.int64).contiguous()
        L.check(L.lib().ois_env_reset_idx(
            C.byref(self._params), C.byref(self._state), L.ptr(env_ids), env_ids.numel(),
            L.ptr(self._supplied_reset), self._seed, (1 << 40) + self._launches, self._num_envs, L.current_stream()))
        self.kernel_launches += 1

    def reset(self):
        """RLTask.reset (rl_task.py:251-253): flag every env for reset."""
        self.reset_buf.fill_(True)   # in place: the buffer is a view of _out_pack (the reference allocates a new one)

    def pre_physics_step(self, actions, record: bool = True, _ctl=None, _host_out=None, _done_flag=None):
        """KAMMA_medium.py:596-656 — and, fused behind it, everything post_physics_step computes.
        ``record=False`` suppresses the fused replay append for this step (the wrapper's reset step).
        ``_ctl`` (rollout.RolloutGraph): device OisStepCtl the kernel takes its per-step scalars and replay row from.
        ``_host_out``: device-visible addresses (obs_clipped, rew, reset) of pinned HOST memory the kernel stores this
        step's results to directly, in place of the task's own device buffers.  ``_done_flag``: device-visible address of a
        pinned host word that receives the completed-step count (OisEnvIO.done_flag)."""
        if self._env is not None and hasattr(self._env, "world") and not self._env.world.is_playing():
            return
        if not (actions.is_cuda and actions.dtype == torch.float32 and actions.is_contiguous()
                and actions.numel() == self._num_envs * self._num_actions):
            actions = actions.to(device=self._device, dtype=torch.float32).reshape(self._num_envs, self._num_actions).contiguous()
        self.actions = actions      # :600 keeps a clone; the kernel only reads it, so a reference is enough
        # the io block is kept between steps: only what changes per step is rewritten (this call sits on the host's
        # critical path: at 4096 envs the GPU step is ~30 us, so every microsecond of Python here is visible)
        bufs = (self.obs_buf, self.rew_buf, self.reset_buf, self.wheel_torqs, self.timeout_buf, self.torques,
                self.guiding_reward)
        key = tuple(map(id, bufs)) + (bool(self.intermediate_outputs),)
        if self._io_key != key:
            io = L.EnvIO()
            io.philox_seed = self._seed
            io.obs, io.rew, io.reset = self.obs_buf.data_ptr(), self.rew_buf.data_ptr(), self.reset_buf.data_ptr()
            io.wheel_torqs = self.wheel_torqs.data_ptr()
            if self.intermediate_outputs:
                io.timeout = self.timeout_buf.data_ptr()
                io.torques, io.guiding = self.torques.data_ptr(), self.guiding_reward.data_ptr()
            if self.variant.eval_sums:
                io.any_reset_flags = self._any_reset_flags.data_ptr()
            self._io, self._io_key = io, key
            self._io_static = (io.rew, io.reset, [t.data_ptr() for t in self._obs_clipped])
        io = self._io
        io.actions = actions.data_ptr()
        io.clip_actions = self.clip_actions if math.isfinite(self.clip_actions) else -1.0
        io.rand_mask = L.ptr(self._supplied_mask)
        io.reset_uniforms = L.ptr(self._supplied_reset)
        io.philox_offset = self._launches
        io.rew, io.reset = self._io_static[0], self._io_static[1]
        io.ep_stats = io.comp0 = None
        io.ctl = _ctl
        io.done_flag = _done_flag
        want_stats = _ctl is None and self.stats_interval > 0 and (self._launches % self.stats_interval == 0)
        if self.stats_log is not None and _ctl is None and self._stats_row < self.stats_log.shape[0]:
            want_stats = False
            io.ep_stats = self.stats_log[self._stats_row].data_ptr()
            self._stats_row += 1
        if want_stats:
            self._ep_stats.zero_()
            io.ep_stats = self._ep_stats.data_ptr()
            io.comp0 = self._comp0.data_ptr()
        io.flag_slot = self._launches % 3
        io.clip_obs = self.clip_obs if math.isfinite(self.clip_obs) else -1.0
        which = 0 if _ctl is not None else self._launches & 1
        self.obs_clipped = self._obs_clipped[which]
        io.obs_clipped = self._io_static[2][which]
        if _host_out is not None:
            io.obs_clipped, io.rew, io.reset = _host_out
        if self._replay is not None and record:
            m = self._replay
            r = m.memory_index if _ctl is None else 0   # with a ctl block the kernel offsets the base by ctl->replay_row
            T = m.tensors
            mkey = (id(m), m.generation)     # create_tensor bumps the generation (agent.init re-creates the tensors)
            if self._replay_key != mkey:     # base address and bytes per ring slot of each (memory_size, N, dim) tensor
                self._replay_rows = [(T[k].data_ptr(), T[k].stride(0) * T[k].element_size())
                                     for k in ("states", "actions", "rewards", "next_states", "terminated", "truncated")]
                self._replay_key = mkey
            rr = self._replay_rows
            io.replay_states = rr[0][0] + r * rr[0][1]
            io.replay_actions = rr[1][0] + r * rr[1][1]
            io.replay_rewards = rr[2][0] + r * rr[2][1]
            io.replay_next_states = rr[3][0] + r * rr[3][1]
            io.replay_terminated = rr[4][0] + r * rr[4][1]
            io.replay_truncated = rr[5][0] + r * rr[5][1]
            io.relabel = self._relabel
            m.advance()
        else:
            io.replay_states = io.replay_actions = io.replay_rewards = io.replay_next_states = None
            io.replay_terminated = io.replay_truncated = None
        L.check(L.lib().ois_env_step(C.byref(self._params), C.byref(self._state), C.byref(io), self._num_envs,
                                     L.current_stream(self._device.index)))
        self.kernel_launches += 1
        self._launches += 1
        if want_stats:
            self._refresh_extras()

    def post_physics_step(self):
        """KAMMA_medium.py:674-727: returns the buffers the fused step produced."""
        self.progress_buf += 1
        return self.obs_buf, self.rew_buf, self.reset_buf, self.extras

    def step(self, actions, record: bool = True, _ctl=None, _host_out=None, _done_flag=None):
        self.pre_physics_step(actions, record, _ctl, _host_out, _done_flag)
        return self.obs_buf, self.rew_buf, self.reset_buf, self.extras

    def get_observations(self):   # :825-869
        return {"robot_view": {"obs_buf": self.obs_buf}}

    def calculate_metrics(self):  # :767-822
        return self.rew_buf

    def is_done(self):            # :736-765 (reset_buf / timeout_buf are already written by the fused step)
        return None

    def get_states(self):
        return self.states_buf

    def get_extras(self):
        return self.extras

    # ------------------------------------------------------------------ strided logging (one small D2H, opt-in)
    def _refresh_extras(self):
        st = self._ep_stats.cpu().numpy()
        n_reset = float(st[18])
        if n_reset > 0:   # KAMMA_medium.py:480-488
            T = self.max_episode_length_s
            keys = SUM_KEYS + (EVAL_SUM_KEYS if self.variant.eval_sums else ())
            ep = {"rew_" + k: torch.tensor(st[i] / n_reset / T) for i, k in enumerate(keys)}
            for i, k in enumerate(RATIO_KEYS):
                ep["rew_" + k] = torch.tensor(st[11 + i] / n_reset / T)
            if self.variant.blend == L.BLEND_LINEAR:
                ep["gamma assist"] = torch.tensor(st[19] / self._num_envs)
            else:
                ep["gamma_1 action assist"] = torch.tensor(st[19] / self._num_envs)
                ep["gamma_2 reward assist"] = torch.tensor(st[20] / self._num_envs)
            ep["terrain_level"] = torch.tensor(0.0)
            self.extras["episode"] = ep

    @property
    def observed_components(self):
        """env-0 scalars (KAMMA_medium.py:840-867); needs ``stats_interval > 0`` (they ride on the stats step)."""
        c = self._comp0.cpu().numpy()
        s = self.action_scale
        g2, rew = float(c[19]), float(c[27])
        pct = lambda v, w: (100.0 * w * float(v) / rew) if rew != 0.0 else float("nan")
        return {
            "env0_desired_v": float(c[0]), "env0_current_v": float(c[1]), "env0_desired_omega": float(c[2]),
            "env0_current_omega": float(c[3]), "env0_v_delta": float(c[4]), "env0_omega_delta": float(c[5]),
            "env0_linear_acc": float(c[6]), "env0_angular_acc": float(c[7]), "env0_episode_count": float(c[8]),
            "env0_torque_apl_fl": float(c[9]), "env0_torque_apl_rl": float(c[10]), "env0_torque_apl_fr": float(c[11]),
            "env0_torque_apl_rr": float(c[12]), "env0_exp_left": float(c[13]), "env0_exp_right": float(c[14]),
            "env0_policy_torque_fl": s * float(c[15]), "env0_policy_torque_rl": s * float(c[16]),
            "env0_policy_torque_fr": s * float(c[17]), "env0_policy_torque_rr": s * float(c[18]),
            "env0_perc_r1": pct(c[20], 1 - g2), "env0_perc_r2": pct(c[21], 1 - g2), "env0_perc_r3": pct(c[22], 1 - g2),
            "env0_perc_dense": pct(c[23], 1 - g2), "env0_perc_sparse": pct(c[24], 1 - g2),
            "env0_perc_observed": pct(c[25], 1 - g2), "env0_perc_guiding": pct(c[26], g2),
        }

    # ------------------------------------------------------------------ dynamics only (float64-twin test)
    def dynamics_step(self, wheel_torqs):
        """``decimation`` sub-steps under constant torques: the stand-in for SimulationContext.step x10."""
        t = wheel_torqs.to(device=self._device, dtype=torch.float32).contiguous()
        L.check(L.lib().ois_dynamics_step(C.byref(self._params), C.byref(self._state), L.ptr(t), self._num_envs,
                                          L.current_stream()))
        self.kernel_launches += 1


class VecEnv:
    """What the trainers see: skrl's ``wrap_env(env, "omniverse-isaacgym")`` over VecEnvRLGames
    (third-party, restated — SURVEY.md §8(c)): ``step`` returns ``(obs, rew[N,1], terminated[N,1],
    truncated[N,1] = 0, extras)``, observations clamped to +-clipObservations, ``reset`` runs once."""

    def __init__(self, task: TorqueDistributionTask):
        self.task = self._task = task
        self.num_envs = task.num_envs
        self.num_agents = 1
        self.device = task.device
        self.observation_space = task.observation_space
        self.action_space = task.action_space
        self._reset_once = True
        self._obs = None
        self._truncated = torch.zeros(task.num_envs, 1, dtype=torch.bool, device=task.device)

    def reset(self):
        if self._reset_once:
            if not self.task.init_done:
                self.task.post_reset()
            # VecEnvRLGames.reset steps once with zero actions to produce the first observation
            zero = torch.zeros(self.num_envs, self.task.num_actions, device=self.device)
            self.task.step(zero, record=False)
            self._obs = self.task.obs_clipped
            self._reset_once = False
        return self._obs, {}

    def step(self, actions):
        """Both clamps of VecEnvRLGames.step (actions to +-clipActions, observations to +-clipObservations) are
        done inside the fused kernel; the returned observation tensor alternates between two buffers."""
        _, rew, reset, extras = self.task.step(actions)
        self._obs = self.task.obs_clipped
        return self._obs, rew.view(-1, 1), reset.view(-1, 1), self._truncated, extras

    def render(self, *a, **k):
        return None

    def close(self):
        pass


def wrap_env(task, wrapper: str = "omniverse-isaacgym") -> VecEnv:
    return VecEnv(task)
