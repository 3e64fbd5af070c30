"""TEST INFRASTRUCTURE ONLY — numpy restatement of the reference's env tensor math
(``TorqueDistributionTask.pre_physics_step`` / ``post_physics_step`` and everything they call).

Only ``tests/``, ``__graft_entry__.smoke()`` and ``bench.py``'s cpu_baseline / ``--impl reference``
legs may import this module; the product (``options_in_simulation_b200``) never does.

Pinning: ``oracle/gen_golden.py`` runs the UNMODIFIED reference classes
(/root/reference/src/my_envs/{KAMMA_medium,KAMMA_eval,KAMMA_curr1,KA_DDPG_1D,KA_DDPG_2D,KA_DDPG_4D}.py)
through ``oracle/ref_harness`` on the same inputs and writes ``tests/golden/env_*.npz``;
``tests/test_oracle_golden.py`` replays those vectors through this file.  The reference itself holds no
golden vectors for this path.

Every block cites the reference lines it restates (file = src/my_envs/KAMMA_medium.py unless noted).
Arithmetic is float32, op by op in the reference's order (python-double scalars are cast to float32
exactly where torch casts them); integer/bool results are therefore bit-exact, floats agree to 1 ulp
of the few library calls involved (norm).
"""
from __future__ import annotations

from dataclasses import dataclass, field

import numpy as np

from .dynamics_twin import SkidSteerParams, skid_substep

F = np.float32

SUM_KEYS = (  # live accumulators, KAMMA_medium.py:803-810
    "r1: Tracking error reward (squared errors)",
    "r2: Convergence reward (squared accelerations)",
    "r3: Torque penalty (sum of squared torques)",
    "Dense reward",
    "Sparse reward",
    "Guiding reward",
    "Observed reward",
    "Final reward",
)
EVAL_SUM_KEYS = ("Smoothness", "Tracking error", "Desired velocity")  # KAMMA_eval.py:224-226
RATIO_KEYS = (  # KAMMA_medium.py:470-476 (keys as they appear after the first reset)
    "r1/Final reward", "r2/Final reward", "r3/Final reward", "Dense/Final reward",
    "Sparse/Final reward", "Guiding/Final reward", "Observed/Final reward",
)


@dataclass(frozen=True)
class EnvVariant:
    """What differs between the reference's env files."""

    name: str
    act_dim: int            # 1 / 2 / 4
    blend: str              # "linear" (KA_DDPG_*.py:633) | "switch" (KAMMA_medium.py:632-636)
    gamma1: str             # "decay" | "one" | "zero"  (action assist)
    gamma2: str             # "decay" | "one" | "zero"  (reward assist)
    guiding: str            # "pre" clip (KA_DDPG_1D.py:642) | "post" clip (KAMMA_medium.py:656)
    command: str            # "rand" | "gv" (KAMMA_curr1.py:577-599) | "bd" (KAMMA_curr2.py:601-641) | "eval" (KAMMA_eval.py:545-550)
    max_sim_steps: float
    eval_sums: bool = False
    action_low: float = 0.0


VARIANTS = {
    "KA_DDPG_1D": EnvVariant("KA_DDPG_1D", 1, "linear", "one", "one", "pre", "rand", 700000.0),
    "KA_DDPG_2D": EnvVariant("KA_DDPG_2D", 2, "linear", "one", "one", "pre", "rand", 700000.0),
    "KA_DDPG_4D": EnvVariant("KA_DDPG_4D", 4, "linear", "decay", "decay", "pre", "rand", 700000.0, action_low=-1.0),
    "KAMMA_medium": EnvVariant("KAMMA_medium", 4, "switch", "decay", "decay", "post", "rand", 500000.0),
    "KAMMA": EnvVariant("KAMMA", 4, "switch", "decay", "decay", "post", "rand", 400000.0),
    "KAMMA_curr1": EnvVariant("KAMMA_curr1", 4, "switch", "decay", "decay", "post", "gv", 500000.0),
    "KAMMA_curr2": EnvVariant("KAMMA_curr2", 4, "switch", "decay", "decay", "post", "bd", 500000.0),
    "KAMMA_g1d_g21": EnvVariant("KAMMA_g1d_g21", 4, "switch", "decay", "one", "post", "rand", 500000.0),
    "KAMMA_g11_g21": EnvVariant("KAMMA_g11_g21", 4, "switch", "one", "one", "post", "rand", 100000.0),   # KAMMA_g11_g21.py:179
    "KAMMA_eval": EnvVariant("KAMMA_eval", 4, "switch", "zero", "zero", "post", "eval", 50000.0, eval_sums=True),
    # eval twins of the 1d / 2d envs (KA_DDPG_1D_eval.py:57,180,193,224-226,543-548; KA_DDPG_2D_eval.py:57,180,193,544-549):
    # gamma frozen at 0, the eval ramp wins over RandSampling, the three eval sums
    "KA_DDPG_1D_eval": EnvVariant("KA_DDPG_1D_eval", 1, "linear", "zero", "zero", "pre", "eval", 50000.0, eval_sums=True),
    "KA_DDPG_2D_eval": EnvVariant("KA_DDPG_2D_eval", 2, "linear", "zero", "zero", "pre", "eval", 50000.0, eval_sums=True),
}


@dataclass
class EnvState:
    """Per-env state, body-frame twist; all float32 / int64."""

    px: np.ndarray
    py: np.ndarray
    yaw: np.ndarray
    vx: np.ndarray          # == v_forward_projected at step start (KAMMA_medium.py:710)
    vy: np.ndarray
    wz: np.ndarray          # == base_velocities[:, 5]
    w: np.ndarray           # [N,4] wheel speeds, DOF order [fl, rl, fr, rr]
    prev_lin: np.ndarray    # previous_linear_velocity[:, 0]
    prev_ang: np.ndarray    # previous_angular_velocity[:, 2]
    commands: np.ndarray    # [N,4] x, y, yaw-rate, heading
    sim_steps: np.ndarray   # int64
    episode_buf: np.ndarray  # int64
    episode_count: np.ndarray  # int64
    sums: np.ndarray        # [N, 8 (+3 eval)]
    old_torques: np.ndarray  # [N,4] (eval only)

    @staticmethod
    def zeros(n: int, n_sums: int = 11) -> "EnvState":
        z = lambda *s: np.zeros(s, dtype=F)
        return EnvState(z(n), z(n), z(n), z(n), z(n), z(n), z(n, 4), z(n), z(n), z(n, 4),
                        np.zeros(n, np.int64), np.zeros(n, np.int64), np.zeros(n, np.int64),
                        z(n, n_sums), z(n, 4))

    def copy(self) -> "EnvState":
        return EnvState(**{k: v.copy() for k, v in self.__dict__.items()})


@dataclass
class EnvParams:
    kp: float = 0.035               # control.stiffness, train/train_kamma_4d_nofifo_random_RLIL.py:64,82
    kp_omega: float = 0.665         # :613
    vehicle_mass: float = 25.0      # :174
    vehicle_inertia: float = 1.05   # :175
    dt: float = 0.1                 # decimation * sim dt, :260
    action_scale: float = 3.0       # :78,238
    torque_clip: float = 4.0        # :649
    max_episode_length: int = 100   # int(10.0 / 0.1 + 0.5), :262
    max_episode_length_s: float = 10.0
    bound: float = 50.0             # :187
    x_range: tuple = (1.0, 2.0)     # --linear_x, train/...:69,91
    y_range: tuple = (-0.5, 0.5)    # :72
    heading_range: tuple = (1.0, 1.1)  # "yaw", :73
    physics: bool = True
    skid: SkidSteerParams = field(default_factory=SkidSteerParams)


def _gamma(kind: str, sim_steps: np.ndarray, max_sim_steps: float) -> np.ndarray:
    # :627-628  torch.clamp(1.0 - (sim_steps.float() / max_sim_steps), min=0.0)
    if kind == "decay":
        return np.maximum(F(1.0) - sim_steps.astype(F) / F(max_sim_steps), F(0.0)).astype(F)
    return np.full(sim_steps.shape, 1.0 if kind == "one" else 0.0, dtype=F)


def _norm4(x: np.ndarray) -> np.ndarray:
    # torch.norm(x, dim=1) on [N,4] float32
    return np.sqrt((x * x).sum(axis=1, dtype=F)).astype(F)


class EnvOracle:
    def __init__(self, variant: EnvVariant, state: EnvState, params: EnvParams | None = None):
        self.v = variant
        self.s = state
        self.p = params or EnvParams()
        self.n = state.px.shape[0]
        self.any_reset_prev = False     # KAMMA_eval.py:470 zeroes old_torques of ALL envs on any reset
        self.force_gamma = None         # KA-DDPG 1d/2d: gamma_assist is a frozen attribute (KA_DDPG_1D.py:192,629)
        self.extras: dict = {}
        self.last: dict = {}

    # ------------------------------------------------------------------ A1: action expansion
    def expand_actions(self, actions: np.ndarray) -> np.ndarray:
        a = np.asarray(actions, dtype=F).reshape(self.n, self.v.act_dim)
        if self.v.act_dim == 1:      # KA_DDPG_1D.py:600-602  actions.repeat(1, 4)
            return np.repeat(a, 4, axis=1)
        if self.v.act_dim == 2:      # KA_DDPG_2D.py:624  [a0, a0, a1, a1]
            return np.stack([a[:, 0], a[:, 0], a[:, 1], a[:, 1]], axis=1)
        return a.copy()              # KAMMA_medium.py:600

    # ------------------------------------------------------------------ one policy step
    def step(self, actions, rand_mask=None, reset_uniforms=None):
        """``rand_mask`` [N]: the uniforms of KAMMA_medium.py:633; ``reset_uniforms`` [N,8]: the draws
        of reset_idx in reference order (pos x, pos y, angle, cmd x, cmd y, heading, loop x, loop omega)."""
        v, p, s, n = self.v, self.p, self.s, self.n
        act4 = self.expand_actions(actions)

        # ---- pre_physics_step :604-610
        desired_v = s.commands[:, 0].copy()
        desired_omega = s.commands[:, 2].copy()
        current_v = s.vx.copy()
        current_omega = s.wz.copy()
        v_delta = desired_v - current_v
        omega_delta = desired_omega - current_omega

        # ---- A2 criteria controller :613-624
        dt = F(p.dt)
        lin = F(p.kp) * (F(p.vehicle_mass) * (v_delta / dt))
        ang = F(p.kp_omega) * (F(p.vehicle_inertia) * (omega_delta / dt))
        ac_left = lin - ang
        ac_right = lin + ang
        criteria = np.stack([ac_left, ac_left, ac_right, ac_right], axis=1)

        # ---- A3 gamma, blend, clip :627-636,649
        g1 = _gamma(v.gamma1, s.sim_steps, v.max_sim_steps)
        g2 = _gamma(v.gamma2, s.sim_steps, v.max_sim_steps)
        if self.force_gamma is not None:
            g1 = np.full(n, self.force_gamma, dtype=F)
            g2 = g1.copy()
        scale = F(p.action_scale)
        if v.blend == "switch":
            mask = (np.asarray(rand_mask, dtype=F).reshape(n) < g1).astype(F)[:, None]
            torques = mask * criteria + (F(1.0) - mask) * (act4 * scale)
        else:   # KA_DDPG_1D.py:632-633
            g = g1[:, None]
            torques = (F(1.0) - g) * act4 * scale + g * criteria
        if v.guiding == "pre":   # KA_DDPG_1D.py:642
            guiding = -_norm4(act4 * scale - criteria)
        wheel_torqs = np.clip(torques, F(-p.torque_clip), F(p.torque_clip))

        # ---- A4 physics: decimation sub-steps with constant clipped torques :646-653
        if p.physics:
            st = {"px": s.px, "py": s.py, "yaw": s.yaw, "vx": s.vx, "vy": s.vy, "wz": s.wz, "w": s.w}
            st = {k: a.astype(np.float64) for k, a in st.items()}
            tau64 = wheel_torqs.astype(np.float64)
            for _ in range(p.skid.decimation):
                st = skid_substep(st, tau64, p.skid, np.float64)
            s.px, s.py, s.yaw = st["px"].astype(F), st["py"].astype(F), st["yaw"].astype(F)
            vx_new, s.vy, wz_new, s.w = st["vx"].astype(F), st["vy"].astype(F), st["wz"].astype(F), st["w"].astype(F)
        else:
            vx_new, wz_new = s.vx.copy(), s.wz.copy()

        if v.guiding == "post":  # :656
            guiding = -_norm4(wheel_torqs - criteria)

        if v.eval_sums:          # KAMMA_eval.py:677-685
            if self.any_reset_prev:
                s.old_torques[:] = 0
            diff = s.old_torques - wheel_torqs
            s.sums[:, 8] += (diff * diff).sum(axis=1, dtype=F) / F(10.0)
            s.sums[:, 10] += desired_v / F(10.0)
            s.old_torques = wheel_torqs.copy()

        # ---- post_physics_step :675-676
        s.episode_buf += 1
        s.sim_steps += 1

        # ---- A5 is_done :736-760 (has_fallen is identically False for a planar model)
        timeout = s.episode_buf >= (p.max_episode_length - 1)
        oob = (s.px < F(-p.bound)) | (s.px > F(p.bound)) | (s.py < F(-p.bound)) | (s.py > F(p.bound))
        reset = timeout | oob

        # ---- A6 calculate_metrics :767-810 (linear_acc lags one step: it uses the pre-physics projection)
        linear_acc = (current_v - s.prev_lin) / dt
        angular_acc = (wz_new - s.prev_ang) / dt
        s.prev_lin = current_v.copy()
        s.prev_ang = wz_new.copy()
        r1 = v_delta * v_delta + F(5.0) * (omega_delta * omega_delta)
        r2 = linear_acc * linear_acc + angular_acc * angular_acc
        r3 = (wheel_torqs * wheel_torqs).sum(axis=1, dtype=F)
        w1, w2, w3 = F(-25.0), F(-0.02), F(-0.09)
        rdense = w1 * r1 + w2 * r2 + w3 * r3
        sparse = np.where((np.abs(v_delta) < F(0.01)) & (np.abs(omega_delta) < F(0.01)), F(0.2), F(0.0)).astype(F)
        observed = rdense + sparse
        rew = (F(1.0) - g2) * observed + g2 * guiding
        s.sums[:, 0] += w1 * r1
        s.sums[:, 1] += w2 * r2
        s.sums[:, 2] += w3 * r3
        s.sums[:, 3] += rdense
        s.sums[:, 4] += sparse
        s.sums[:, 5] += guiding
        s.sums[:, 6] += observed
        s.sums[:, 7] += rew
        if v.eval_sums:          # KAMMA_eval.py:840
            s.sums[:, 9] += np.abs(v_delta) / F(10.0)

        # state after physics, before reset
        s.vx, s.wz = vx_new, wz_new

        # ---- A7r reset_idx :417-498
        ids = np.nonzero(reset)[0]
        self.any_reset_prev = False
        if ids.size:
            self._reset(ids, g1, g2, reset_uniforms)
            self.any_reset_prev = True

        # ---- A7 get_observations :833 (desired_* are views of commands -> post-reset values)
        obs = np.stack([s.commands[:, 0], s.commands[:, 2], v_delta, omega_delta, linear_acc, angular_acc], axis=1)

        self.last = dict(
            act4=act4, criteria=criteria, torques=torques, wheel_torqs=wheel_torqs, guiding=guiding,
            v_delta=v_delta, omega_delta=omega_delta, gamma1=g1, gamma2=g2, linear_acc=linear_acc,
            angular_acc=angular_acc, timeout=timeout.astype(np.int64), reset=reset, rew=rew, obs=obs,
            comp=(w1 * r1, w2 * r2, w3 * r3, rdense, sparse, observed), current_v=current_v,
            current_omega=current_omega, desired_v=desired_v, desired_omega=desired_omega,
        )
        return obs, rew, reset

    # ------------------------------------------------------------------ reset_idx
    def _reset(self, ids, g1, g2, U):
        v, p, s = self.v, self.p, self.s
        U = np.asarray(U, dtype=F)
        u = U[ids]
        # pose :422-434 ; base_init_state pos = (0, 0, 0.1), rot = identity
        s.px[ids] = F(0.0) + (F(1.0) * u[:, 0] + F(-0.5))
        s.py[ids] = F(0.0) + (F(1.0) * u[:, 1] + F(-0.5))
        s.yaw[ids] = F(2 * np.pi) * u[:, 2] + F(0.0)
        # velocities / wheel state :435-446
        s.vx[ids] = 0
        s.vy[ids] = 0
        s.wz[ids] = 0
        s.w[ids] = 0
        # commands :448-460
        cx = F(p.x_range[1] - p.x_range[0]) * u[:, 3] + F(p.x_range[0])
        cy = F(p.y_range[1] - p.y_range[0]) * u[:, 4] + F(p.y_range[0])
        ch = F(p.heading_range[1] - p.heading_range[0]) * u[:, 5] + F(p.heading_range[0])
        keep = (np.sqrt((cx * cx + cy * cy).astype(F)) > F(0.25)).astype(F)
        s.commands[ids, 0] = cx * keep
        s.commands[ids, 1] = cy * keep
        s.commands[ids, 2] = s.commands[ids, 2] * keep
        s.commands[ids, 3] = ch * keep
        s.episode_buf[ids] = 0      # :462
        s.episode_count[ids] += 1   # :463

        # episode statistics :470-488 (logging only; ratios may be nan/inf like the reference's)
        T = F(p.max_episode_length_s)
        n_live = 8 + (3 if v.eval_sums else 0)
        with np.errstate(divide="ignore", invalid="ignore"):
            fin = s.sums[:, 7]
            one_m = F(1.0) - g2
            ratios = [T * (F(100.0) * one_m * s.sums[:, k] / fin) for k in (0, 1, 2, 3, 4)]
            ratios.append(T * (F(100.0) * g2 * s.sums[:, 5] / fin))
            ratios.append(T * (F(100.0) * one_m * s.sums[:, 6] / fin))
            ep = {}
            keys = SUM_KEYS + (EVAL_SUM_KEYS if v.eval_sums else ())
            for k in range(n_live):
                ep["rew_" + keys[k]] = F(s.sums[ids, k].mean(dtype=F) / T)
            for k, name in enumerate(RATIO_KEYS):
                ep["rew_" + name] = F(ratios[k][ids].mean(dtype=F) / T)
        if v.blend == "linear":     # KA_DDPG_1D.py:487
            ep["gamma assist"] = F(g1.mean(dtype=F))
        else:                       # KAMMA_medium.py:486-487
            ep["gamma_1 action assist"] = F(g1.mean(dtype=F))
            ep["gamma_2 reward assist"] = F(g2.mean(dtype=F))
        self.extras["episode"] = ep
        s.sums[ids] = 0             # :485

        # per-env command :490-496 via sample_velocity_command
        if v.command == "rand":      # :569-575
            x = F(p.x_range[1] - p.x_range[0]) * u[:, 6] + F(p.x_range[0])
        elif v.command == "gv":      # KAMMA_curr1.py:577-599 (per-env generalisation of an N=1-only formula)
            progress = np.minimum(F(1.0), s.sim_steps[ids].astype(F) / F(v.max_sim_steps))
            factor = F(0.1) + F(0.9) * progress
            lo, hi = p.x_range
            c = (lo + hi) / 2.0
            x_min = F(c) - F(c - lo) * factor
            x_max = F(c) + F(hi - c) * factor
            x = (x_max - x_min) * u[:, 6] + x_min
        elif v.command == "bd":      # KAMMA_curr2.py:601-641 (per-env generalisation of an N=1-only formula)
            # u[:, 6] is the uniform of `torch.rand(1) < 0.5`, u[:, 7] the standard normal behind torch.normal(mean, 0.05)
            # (= randn * std + mean); the omega draws that follow are discarded by the reference (omega = 0.0)
            progress = np.minimum(F(1.0), s.sim_steps[ids].astype(F) / F(v.max_sim_steps))
            lo, hi = p.x_range
            c = (lo + hi) / 2.0
            left = F(c) - progress * F(c - lo)
            right = F(c) + progress * F(hi - c)
            x = u[:, 7] * F(0.05) + np.where(u[:, 6] < F(0.5), left, right).astype(F)
        elif v.command == "eval":    # KAMMA_eval.py:545-550
            progress = np.clip(s.sim_steps[ids].astype(F) / F(v.max_sim_steps), F(0.0), F(1.0))
            x = F(1.0) + progress * F(1.0)
        else:
            raise ValueError(v.command)
        s.commands[ids, 0] = np.maximum(x, F(0.0))
        s.commands[ids, 2] = 0
