"""CPU: the numpy restatement (oracle/env_oracle.py) replayed against the golden vectors that
oracle/gen_golden.py recorded from the UNMODIFIED reference classes (src/my_envs/*.py)."""
import os

import numpy as np
import pytest

from conftest import GOLDEN, golden_files
from oracle.env_oracle import VARIANTS, EnvOracle, EnvParams, EnvState

F = np.float32


def load_case(fname):
    d = np.load(os.path.join(GOLDEN, fname), allow_pickle=False)
    variant = VARIANTS[str(d["variant"])]
    n, steps, physics = int(d["n"]), int(d["steps"]), bool(d["physics"])
    fg = float(d["force_gamma"])
    s0 = EnvState(**{k: d[f"s0_{k}"].copy() for k in EnvState.zeros(1).__dict__})
    return d, variant, n, steps, physics, (None if fg < 0 else fg), s0


def rel_err(ref, got):
    ref = np.asarray(ref, np.float64)
    return np.abs(ref - np.asarray(got, np.float64)) / np.maximum(1.0, np.abs(ref))


@pytest.mark.parametrize("fname", golden_files())
def test_env_oracle_matches_reference_vectors(fname):
    d, variant, n, steps, physics, fg, s0 = load_case(fname)
    o = EnvOracle(variant, s0, EnvParams(physics=physics))
    if fg is not None:
        o.force_gamma = F(fg)
    tol = 2e-5 if physics else 2e-6
    for t in range(steps):
        obs, rew, reset = o.step(d[f"t{t}_actions"], d[f"t{t}_rand_mask"], d[f"t{t}_reset_uniforms"])
        assert np.array_equal(reset, d[f"t{t}_reset"].astype(bool))            # integer/bool: bit-exact
        assert np.array_equal(o.last["timeout"], d[f"t{t}_timeout"])
        for k in ("sim_steps", "episode_buf", "episode_count"):
            assert np.array_equal(getattr(o.s, k), d[f"t{t}_state_{k}"]), k
        assert rel_err(d[f"t{t}_obs"], obs).max() <= tol
        assert rel_err(d[f"t{t}_rew"], rew).max() <= tol
        for k in ("wheel_torqs", "torques", "guiding"):
            assert rel_err(d[f"t{t}_{k}"], o.last[k]).max() <= tol, k
        for k in ("px", "py", "vx", "vy", "wz", "w", "prev_lin", "prev_ang", "commands"):
            assert rel_err(d[f"t{t}_state_{k}"], getattr(o.s, k)).max() <= tol, k
        ns = d[f"t{t}_state_sums"].shape[1]
        assert rel_err(d[f"t{t}_state_sums"], o.s.sums[:, :ns]).max() <= tol
        if f"t{t}_extras_keys" in d and "episode" in o.extras:
            for k, v in zip(d[f"t{t}_extras_keys"], d[f"t{t}_extras_vals"]):
                if str(k) not in o.extras["episode"]:
                    continue      # ratio keys are spelled differently per env file; logging only
                mine = float(o.extras["episode"][str(k)])
                if np.isfinite(v) and np.isfinite(mine):
                    assert abs(v - mine) <= 1e-4 * max(1.0, abs(v)), k


def test_golden_covers_all_blend_and_command_modes():
    seen = {str(np.load(os.path.join(GOLDEN, f))["variant"]) for f in golden_files()}
    assert {"KAMMA_medium", "KAMMA_eval", "KAMMA_curr1", "KAMMA_curr2", "KA_DDPG_1D", "KA_DDPG_2D", "KA_DDPG_4D"} <= seen


def test_dynamics_twin_float32_tracks_float64():
    """The increment form keeps a float32 evaluation of the SAME equations within 1e-6 of float64
    (this is the property the CUDA kernel relies on)."""
    from oracle.dynamics_twin import SkidSteerParams, random_state, skid_step

    rng = np.random.default_rng(5)
    st = random_state(512, rng)
    tau = rng.uniform(-4, 4, (512, 4))
    a = skid_step(st, tau, SkidSteerParams(), np.float64)
    b = skid_step({k: v.astype(np.float32) for k, v in st.items()}, tau.astype(np.float32), SkidSteerParams(), np.float32)
    for k in ("vx", "vy", "wz"):
        assert np.abs(a[k] - b[k]).max() < 5e-6, k


def test_dynamics_twin_physical_sanity():
    """Equal torques on a straight-moving robot keep it straight; zero torque decays wheel slip;
    steady state under the P-controller approaches the commanded speed from below (Kp is small)."""
    from oracle.dynamics_twin import SkidSteerParams, skid_step

    p = SkidSteerParams()
    n = 4
    st = {"px": np.zeros(n), "py": np.zeros(n), "yaw": np.zeros(n), "vx": np.full(n, 1.0), "vy": np.zeros(n),
          "wz": np.zeros(n), "w": np.full((n, 4), 1.0 / p.wheel_radius)}
    tau = np.full((n, 4), 1.0)
    for _ in range(20):
        st = skid_step(st, tau, p)
    assert np.abs(st["wz"]).max() < 1e-12 and np.abs(st["vy"]).max() < 1e-12 and np.abs(st["py"]).max() < 1e-12
    assert (st["vx"] > 1.0).all()
    # differential torque turns left (right wheels push harder -> positive yaw rate)
    st2 = {k: v.copy() for k, v in st.items()}
    tau2 = np.stack([np.full(n, 0.2), np.full(n, 0.2), np.full(n, 2.0), np.full(n, 2.0)], axis=1)
    for _ in range(10):
        st2 = skid_step(st2, tau2, p)
    assert (st2["wz"] > 0).all()
