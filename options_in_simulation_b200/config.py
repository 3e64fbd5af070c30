"""Frozen configuration of the env variants and the TASK_CFG dictionaries the
reference's scripts mutate (src/my_envs/KAMMA_medium.py:29-128, train/train_kamma_4d_nofifo_random_RLIL.py:77-91).
"""
from __future__ import annotations

import copy
from dataclasses import dataclass

from . import _lib as L


@dataclass(frozen=True)
class EnvVariant:
    name: str
    act_dim: int
    blend: int            # L.BLEND_*
    gamma1_mode: int      # L.GAMMA_*
    gamma1_const: float
    gamma2_mode: int
    gamma2_const: float
    guiding_post_clip: int
    command_mode: int     # L.CMD_*
    max_sim_steps: float
    eval_sums: bool = False
    action_low: float = 0.0
    relabel: int = L.RELABEL_EXECUTED   # trainer that the reference pairs with this env


_D, _K = L.GAMMA_DECAY, L.GAMMA_CONST

VARIANTS = {
    # KA-DDPG: linear blend, pre-clip guiding reward; 1d/2d never update gamma (KA_DDPG_1D.py:192,629)
    "KA_DDPG_1D": EnvVariant("KA_DDPG_1D", 1, L.BLEND_LINEAR, _K, 1.0, _K, 1.0, 0, L.CMD_RAND, 700000.0, relabel=L.RELABEL_NONE),
    "KA_DDPG_2D": EnvVariant("KA_DDPG_2D", 2, L.BLEND_LINEAR, _K, 1.0, _K, 1.0, 0, L.CMD_RAND, 700000.0, relabel=L.RELABEL_NONE),
    "KA_DDPG_4D": EnvVariant("KA_DDPG_4D", 4, L.BLEND_LINEAR, _D, 1.0, _D, 1.0, 0, L.CMD_RAND, 700000.0, action_low=-1.0),
    # KAMMA: Bernoulli switch, post-clip guiding reward
    "KAMMA_medium": EnvVariant("KAMMA_medium", 4, L.BLEND_SWITCH, _D, 1.0, _D, 1.0, 1, L.CMD_RAND, 500000.0),
    "KAMMA": EnvVariant("KAMMA", 4, L.BLEND_SWITCH, _D, 1.0, _D, 1.0, 1, L.CMD_RAND, 400000.0),
    "KAMMA_curr1": EnvVariant("KAMMA_curr1", 4, L.BLEND_SWITCH, _D, 1.0, _D, 1.0, 1, L.CMD_GV, 500000.0),
    "KAMMA_curr2": EnvVariant("KAMMA_curr2", 4, L.BLEND_SWITCH, _D, 1.0, _D, 1.0, 1, L.CMD_BD, 500000.0),
    "KAMMA_g1d_g21": EnvVariant("KAMMA_g1d_g21", 4, L.BLEND_SWITCH, _D, 1.0, _K, 1.0, 1, L.CMD_RAND, 500000.0),
    "KAMMA_g11_g21": EnvVariant("KAMMA_g11_g21", 4, L.BLEND_SWITCH, _K, 1.0, _K, 1.0, 1, L.CMD_RAND, 100000.0),   # KAMMA_g11_g21.py:179
    "KAMMA_eval": EnvVariant("KAMMA_eval", 4, L.BLEND_SWITCH, _K, 0.0, _K, 0.0, 1, L.CMD_EVAL, 50000.0, eval_sums=True),
    # eval twins of the 1d / 2d envs: gamma frozen at 0 (KA_DDPG_1D_eval.py:193), eval ramp before RandSampling
    # (:543-548; KA_DDPG_2D_eval.py:544-549), Smoothness / Tracking error / Desired velocity sums (:224-226,685-691,857)
    "KA_DDPG_1D_eval": EnvVariant("KA_DDPG_1D_eval", 1, L.BLEND_LINEAR, _K, 0.0, _K, 0.0, 0, L.CMD_EVAL, 50000.0,
                                  eval_sums=True, relabel=L.RELABEL_NONE),
    "KA_DDPG_2D_eval": EnvVariant("KA_DDPG_2D_eval", 2, L.BLEND_LINEAR, _K, 0.0, _K, 0.0, 0, L.CMD_EVAL, 50000.0,
                                  eval_sums=True, relabel=L.RELABEL_NONE),
}

# The parts of the reference's TASK_CFG the hot path reads (KAMMA_medium.py:29-128).  Scripts deep-copy
# and mutate it exactly like the reference's train scripts do.
TASK_CFG = {
    "test": False,
    "device_id": 0,
    "headless": True,
    "sim_device": "cuda:0",
    "seed": 42,
    "task": {
        "name": "TorqueDistributionTask",
        "env": {
            "numEnvs": 16,
            "clipObservations": 1000.0,
            "controlFrequencyInv": 10,
            "baseInitState": {"pos": [0.0, 0.0, 0.1], "rot": [1.0, 0.0, 0.0, 0.0],
                              "vLinear": [0.0, 0.0, 0.0], "vAngular": [0.0, 0.0, 0.0]},
            "dofInitTorques": [0.0, 0.0, 0.0, 0.0],
            "dofInitVelocities": [0.0, 0.0, 0.0, 0.0],
            "terrain": {"staticFriction": 0.85, "dynamicFriction": 0.85, "restitution": 0.0, "curriculum": False},
            "learn": {"terminalReward": 0.0, "episodeLength_s": 10.0},
            "randomCommandVelocityRanges": {"linear_x": [1.5, 1.5], "linear_y": [-0.5, 0.5], "yaw": [1.0, 1.1],
                                            "yaw_constant": 0.5},
            "control": {"decimation": 10, "stiffness": 1.0, "damping": 0.005, "actionScale": 3.0,
                        "wheel_radius": 0.1175},
        },
        "sim": {"dt": 0.01, "gravity": [0.0, 0.0, -9.81],
                "default_physics_material": {"static_friction": 1.0, "dynamic_friction": 1.0, "restitution": 0.0}},
        # robot constants for the skid-steer model that replaces PhysX
        # (src/my_assets/urdf/origin_sym.urdf:11-12,34-35,40,115-139; src/my_robots/origin_v18.py:70)
        "robot": {"body_mass": 21.0, "wheel_mass": 1.0, "body_izz": 1.01, "wheel_inertia": 0.01,
                  "half_wheelbase": 0.205, "half_track": 0.25, "wheel_speed_max": 64.0, "slip_eps": 0.05},
    },
}


def task_cfg() -> dict:
    return copy.deepcopy(TASK_CFG)


def env_params(variant: EnvVariant, cfg: dict, *, physics: bool = True, gamma_const: float | None = None) -> L.EnvParams:
    """Fold a TASK_CFG-style dict into the C parameter block (TorqueDistributionTask.update_config,
    KAMMA_medium.py:231-275, plus the hard-coded attributes at :174-179,613,649)."""
    t = cfg["task"] if "task" in cfg else cfg
    env, sim = t["env"], t["sim"]
    rob = t.get("robot", TASK_CFG["task"]["robot"])
    ctl, rng = env["control"], env["randomCommandVelocityRanges"]
    p = L.EnvParams()
    p.act_dim = variant.act_dim
    p.blend = variant.blend
    p.gamma1_mode, p.gamma2_mode = variant.gamma1_mode, variant.gamma2_mode
    p.gamma1_const = variant.gamma1_const if gamma_const is None else gamma_const
    p.gamma2_const = variant.gamma2_const if gamma_const is None else gamma_const
    p.guiding_post_clip = variant.guiding_post_clip
    p.command_mode = variant.command_mode
    p.physics = 1 if physics else 0
    p.eval_sums = 1 if variant.eval_sums else 0
    p.decimation = int(ctl["decimation"])
    p.sim_dt = float(sim["dt"])
    dt = p.decimation * float(sim["dt"])                       # KAMMA_medium.py:260
    p.dt = dt
    p.max_episode_length_s = float(env["learn"]["episodeLength_s"])
    p.max_episode_length = int(p.max_episode_length_s / dt + 0.5)   # :262
    p.max_sim_steps = variant.max_sim_steps
    p.kp = float(ctl["stiffness"])
    p.kp_omega = 0.665                                          # :613
    p.vehicle_mass, p.vehicle_inertia = 25.0, 1.05              # :174-175
    p.action_scale = float(ctl["actionScale"])
    p.torque_clip = 4.0                                         # :649
    p.bound = 50.0                                              # :187
    x, y, h = rng["linear_x"], rng["linear_y"], rng["yaw"]
    p.x_lo, p.x_hi, p.x_range = float(x[0]), float(x[1]), float(x[1]) - float(x[0])
    p.y_lo, p.y_range = float(y[0]), float(y[1]) - float(y[0])
    p.h_lo, p.h_range = float(h[0]), float(h[1]) - float(h[0])
    mass = float(rob["body_mass"]) + 4.0 * float(rob["wheel_mass"])
    a, b = float(rob["half_wheelbase"]), float(rob["half_track"])
    p.mass = mass
    p.yaw_inertia = float(rob["body_izz"]) + 4.0 * float(rob["wheel_inertia"]) + 4.0 * float(rob["wheel_mass"]) * (a * a + b * b)
    p.wheel_inertia = float(rob["wheel_inertia"])
    p.wheel_radius = float(ctl["wheel_radius"])
    p.half_wheelbase, p.half_track = a, b
    mu = float(sim["default_physics_material"]["dynamic_friction"])
    p.mu_n = mu * mass * abs(float(sim["gravity"][2])) / 4.0
    p.slip_eps = float(rob["slip_eps"])
    p.wheel_damping = float(ctl["damping"])
    p.wheel_speed_max = float(rob["wheel_speed_max"])
    return p


def script_cfg(num_envs: int, *, stiffness: float = 0.035, damping: float = 0.005, static_friction: float = 0.85,
               dynamic_friction: float = 0.85, yaw_constant: float = 0.5, linear_x=(1.0, 2.0), seed: int = 42,
               sim_device: str = "cuda:0") -> dict:
    """TASK_CFG after the mutations every train/eval script applies from its argparse defaults
    (train/train_kamma_4d_nofifo_random_RLIL.py:63-91), with ``numEnvs`` as given."""
    cfg = task_cfg()
    cfg["seed"] = seed
    cfg["sim_device"] = sim_device
    env = cfg["task"]["env"]
    env["numEnvs"] = int(num_envs)
    env["control"]["stiffness"] = stiffness
    env["control"]["damping"] = damping
    mat = cfg["task"]["sim"]["default_physics_material"]
    mat["static_friction"], mat["dynamic_friction"] = static_friction, dynamic_friction
    env["randomCommandVelocityRanges"]["yaw_constant"] = yaw_constant
    env["randomCommandVelocityRanges"]["linear_x"] = list(linear_x)
    return cfg
