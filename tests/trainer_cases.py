"""Shared loading of tests/golden/trainer_*.npz (oracle/gen_golden_trainer.py)."""
import os

import numpy as np

from conftest import GOLDEN

META = ("variant", "trainer", "seed", "n", "timesteps", "batch", "mem_steps", "random_timesteps", "learning_starts",
        "ou_scale", "T_e", "clip_low", "relabel", "stride", "n_updates", "clip_obs", "memory_index", "filled")
STATE_KEYS = ("px", "py", "yaw", "vx", "vy", "wz", "w", "prev_lin", "prev_ang", "commands", "sim_steps", "episode_buf",
              "episode_count", "sums", "old_torques")


def load_case(fname):
    d = np.load(os.path.join(GOLDEN, fname), allow_pickle=False)
    meta = {k: d[k].item() for k in META}
    draws = {k: d[f"draw_{k}"] for k in ("rand_mask", "reset_uniforms", "normals", "uniforms")}
    return d, meta, draws


def make_state(d):
    from oracle.env_oracle import EnvState

    return EnvState(**{k: d[f"s0_{k}"].copy() for k in STATE_KEYS})
