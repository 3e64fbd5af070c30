"""TEST INFRASTRUCTURE ONLY — generates ``tests/golden/*.npz`` by running the UNMODIFIED reference
(/root/reference, through ``oracle/ref_harness``) on seeded synthetic inputs, and checks the numpy /
torch restatements in ``oracle/`` against it on the way.  Runs only in the build container:

    python -m oracle.gen_golden            # rewrites tests/golden/

The GPU box has no /root/reference; there the committed vectors are what the tests read.
Synthetic inputs follow SURVEY.md §8(d).
"""
from __future__ import annotations

import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

from oracle.env_oracle import EVAL_SUM_KEYS, SUM_KEYS, VARIANTS, EnvOracle, EnvParams, EnvState  # noqa: E402

GOLDEN = os.path.join(ROOT, "tests", "golden")
F = np.float32


# --------------------------------------------------------------------------------------------- inputs
def synth_state(variant, n: int, rng: np.random.Generator, *, physics: bool) -> EnvState:
    """SURVEY.md §8(d) synthetic state.  Without physics the heading is 0 so that the reference's
    world-frame projection (KAMMA_medium.py:701-710) reproduces the body-frame vx exactly."""
    s = EnvState.zeros(n)
    s.px[:] = rng.uniform(-0.5, 0.5, n)
    s.py[:] = rng.uniform(-0.5, 0.5, n)
    oob = rng.random(n) < 0.03                      # a few envs just outside the +-50 m bounds
    s.px[oob] = np.where(rng.random(oob.sum()) < 0.5, 50.5, -50.5)
    s.yaw[:] = rng.uniform(0, 2 * np.pi, n) if physics else 0.0
    s.vx[:] = rng.uniform(0.0, 2.0, n)
    s.vy[:] = rng.normal(0, 0.05, n)
    s.wz[:] = rng.normal(0, 0.05, n)
    s.w[:] = s.vx[:, None] / 0.1175 + rng.normal(0, 0.5, (n, 4))
    s.prev_lin[:] = s.vx + rng.normal(0, 0.02, n)
    s.prev_ang[:] = s.wz + rng.normal(0, 0.02, n)
    s.commands[:, 0] = rng.uniform(1.0, 2.0, n)
    s.commands[:, 1] = rng.uniform(-0.5, 0.5, n)
    yawcmd = rng.random(n) < 0.1
    s.commands[yawcmd, 2] = rng.uniform(-0.3, 0.3, yawcmd.sum())
    s.commands[:, 3] = rng.uniform(1.0, 1.1, n)
    near = rng.random(n) < 0.05                     # a few envs inside the sparse-reward window
    s.commands[near, 0] = s.vx[near] + F(0.004)
    s.commands[near, 2] = s.wz[near] - F(0.003)
    s.sim_steps[:] = rng.integers(0, int(variant.max_sim_steps * 1.1), n)
    s.episode_buf[:] = rng.integers(0, 99, n)
    s.episode_buf[rng.random(n) < 0.05] = 97        # time out on the first / second step
    s.episode_count[:] = rng.integers(1, 50, n)
    if n == 1:
        s.episode_buf[:] = 96
    s.sums[:, :8] = rng.normal(-20, 5, (n, 8))
    if variant.eval_sums:
        s.sums[:, 8:] = rng.uniform(0, 3, (n, 3))
        s.old_torques[:] = rng.uniform(-3, 3, (n, 4))
    else:
        s.sums[:, 8:] = 0
    return s


def load_state_into_reference(runner, variant, s: EnvState):
    t = runner.task
    f = lambda a: torch.as_tensor(np.asarray(a).copy())
    runner.set_state({"px": s.px, "py": s.py, "yaw": s.yaw, "vx": s.vx, "vy": s.vy, "wz": s.wz, "w": s.w})
    t.refresh_dof_state_tensors()
    t.refresh_body_state_tensors()
    t.v_forward_projected = f(s.vx)
    t.previous_linear_velocity[:, 0] = f(s.prev_lin)
    t.previous_angular_velocity[:, 2] = f(s.prev_ang)
    t.commands[:] = f(s.commands)
    t.sim_steps[:] = f(s.sim_steps)
    t.episode_buf[:] = f(s.episode_buf)
    t.episode_count[:] = f(s.episode_count)
    keys = SUM_KEYS + (EVAL_SUM_KEYS if variant.eval_sums else ())
    for k, name in enumerate(keys):
        t.episode_sums[name][:] = f(s.sums[:, k])
    if variant.eval_sums:
        t.old_torques = f(s.old_torques)


def read_state_from_reference(runner, variant) -> dict:
    t, v = runner.task, runner.view
    yaw = 2.0 * np.arctan2(v.quat[:, 3].double().numpy(), v.quat[:, 0].double().numpy())
    c, s_ = np.cos(yaw), np.sin(yaw)
    vwx, vwy = v.vel[:, 0].double().numpy(), v.vel[:, 1].double().numpy()
    keys = SUM_KEYS + (EVAL_SUM_KEYS if variant.eval_sums else ())
    sums = np.stack([t.episode_sums[k].numpy() for k in keys], axis=1)
    out = dict(
        px=v.pos[:, 0].numpy().copy(), py=v.pos[:, 1].numpy().copy(), yaw=yaw.astype(F),
        vx=(c * vwx + s_ * vwy).astype(F), vy=(-s_ * vwx + c * vwy).astype(F), wz=v.vel[:, 5].numpy().copy(),
        w=v.jvel.numpy().copy(), prev_lin=t.previous_linear_velocity[:, 0].numpy().copy(),
        prev_ang=t.previous_angular_velocity[:, 2].numpy().copy(), commands=t.commands.numpy().copy(),
        sim_steps=t.sim_steps.numpy().copy(), episode_buf=t.episode_buf.numpy().copy(),
        episode_count=t.episode_count.numpy().copy(), sums=sums.astype(F),
        v_fwd=t.v_forward_projected.numpy().copy(),
    )
    return out


def action_sample(variant, n, rng):
    return rng.uniform(variant.action_low, 1.0, (n, variant.act_dim)).astype(F)


# --------------------------------------------------------------------------------------------- case
def run_env_case(name: str, n: int, steps: int, *, physics: bool, seed: int, force_gamma=None) -> dict:
    from oracle.ref_harness.ref_env import RefRunner

    variant = VARIANTS[name]
    rng = np.random.default_rng(seed)
    init_u = rng.random((n, 8)).astype(F)
    runner = RefRunner(name, n, physics=physics, init_uniforms=init_u)
    s0 = synth_state(variant, n, rng, physics=physics)
    if name in ("KAMMA_curr1", "KAMMA_curr2"):
        assert n == 1, "the reference's gv / bd samplers are N=1 only (KAMMA_curr1.py:579, KAMMA_curr2.py:603)"
    load_state_into_reference(runner, variant, s0)
    if force_gamma is not None:     # KA-DDPG 1d/2d: gamma is a frozen attribute (KA_DDPG_1D.py:192,629)
        runner.task.gamma_assist = torch.full((n,), float(force_gamma))
    params = EnvParams(physics=physics)
    oracle = EnvOracle(variant, s0.copy(), params)
    if force_gamma is not None:
        oracle.force_gamma = F(force_gamma)

    rec = {"variant": name, "n": n, "steps": steps, "physics": physics,
           "force_gamma": -1.0 if force_gamma is None else float(force_gamma)}
    for k, a in s0.__dict__.items():
        rec[f"s0_{k}"] = a
    worst = {}
    for t in range(steps):
        act = action_sample(variant, n, rng)
        rm = rng.random(n).astype(F)
        # put some uniforms exactly around gamma to exercise the rand < gamma compare
        ru = rng.random((n, 8)).astype(F)
        if variant.command == "bd":     # column 7 is the standard normal behind torch.normal (KAMMA_curr2.py:617-619)
            ru[:, 7] = rng.standard_normal(n).astype(F)
        obs_r, rew_r, reset_r, extras_r = runner.step(act, rand_mask=rm if variant.blend == "switch" else None,
                                                      reset_uniforms=ru)
        obs_o, rew_o, reset_o = oracle.step(act, rm, ru)
        ref = read_state_from_reference(runner, variant)
        tk = runner.task
        step_out = dict(
            obs=obs_r.numpy().copy(), rew=rew_r.numpy().copy(), reset=reset_r.numpy().astype(np.uint8),
            wheel_torqs=tk.wheel_torqs.numpy().copy(), torques=tk.torques.numpy().copy(),
            guiding=tk.guiding_reward.numpy().copy(), timeout=tk.timeout_buf.numpy().copy(),
            linear_acc=tk.linear_acc.numpy().copy(), angular_acc=tk.angular_acc.numpy().copy(),
        )
        # ---- restatement vs reference
        L = oracle.last
        assert np.array_equal(step_out["reset"].astype(bool), reset_o), f"{name} step {t}: reset differs"
        assert np.array_equal(step_out["timeout"], L["timeout"]), f"{name} step {t}: timeout differs"
        for key, mine in (("obs", obs_o), ("rew", rew_o), ("wheel_torqs", L["wheel_torqs"]),
                          ("torques", L["torques"]), ("guiding", L["guiding"])):
            d = np.abs(step_out[key].astype(np.float64) - mine.astype(np.float64))
            tol = 1e-5 * np.maximum(1.0, np.abs(step_out[key])) if physics else 2e-6 * np.maximum(1.0, np.abs(step_out[key]))
            worst[key] = max(worst.get(key, 0.0), float((d / np.maximum(1.0, np.abs(step_out[key]))).max()))
            assert (d <= tol).all(), f"{name} step {t}: {key} differs by {d.max()}"
        so = oracle.s
        for key in ("sim_steps", "episode_buf", "episode_count"):
            assert np.array_equal(ref[key], getattr(so, key)), f"{name} step {t}: {key} differs"
        for key in ("px", "py", "yaw", "vx", "vy", "wz", "w", "prev_lin", "prev_ang", "commands", "sums"):
            a, b = ref[key].astype(np.float64), getattr(so, key).astype(np.float64)
            if key == "sums" and not variant.eval_sums:
                b = b[:, :8]
            if key == "yaw":    # compare as angles
                d = np.abs(np.angle(np.exp(1j * (a - b))))
            else:
                d = np.abs(a - b)
            rel = d / np.maximum(1.0, np.abs(a))
            worst[key] = max(worst.get(key, 0.0), float(rel.max()))
            assert rel.max() <= (2e-5 if physics else 2e-6), f"{name} step {t}: state {key} differs by {rel.max()}"
        assert np.abs(ref["v_fwd"] - so.vx).max() <= 2e-6 * max(1.0, np.abs(so.vx).max())
        if "episode" in extras_r and "episode" in oracle.extras:
            for k2, v2 in oracle.extras["episode"].items():
                rv = float(extras_r["episode"][k2])
                if np.isfinite(rv) and np.isfinite(v2):
                    assert abs(rv - float(v2)) <= 1e-4 * max(1.0, abs(rv)), (name, t, k2, rv, v2)
        rec[f"t{t}_actions"] = act
        rec[f"t{t}_rand_mask"] = rm
        rec[f"t{t}_reset_uniforms"] = ru
        for k2, v2 in step_out.items():
            rec[f"t{t}_{k2}"] = v2
        for k2, v2 in ref.items():
            rec[f"t{t}_state_{k2}"] = v2
        if "episode" in extras_r:
            keys = sorted(k for k in extras_r["episode"] if k.startswith("rew_") or k.startswith("gamma"))
            rec[f"t{t}_extras_keys"] = np.array(keys)
            rec[f"t{t}_extras_vals"] = np.array([float(extras_r["episode"][k]) for k in keys], dtype=np.float64)
    print(f"  {name:14s} N={n:<5d} steps={steps} physics={physics!s:5s} resets={int(sum(rec[f't{t}_reset'].sum() for t in range(steps)))}"
          f" worst rel: " + " ".join(f"{k}={v:.1e}" for k, v in worst.items()))
    return rec


ENV_CASES = [
    # (file tag, variant, N, steps, physics, force_gamma)
    ("kamma_medium_nophys", "KAMMA_medium", 64, 4, False, None),
    ("kamma_medium_phys", "KAMMA_medium", 64, 4, True, None),
    ("kamma_eval_nophys", "KAMMA_eval", 64, 4, False, None),
    ("kamma_eval_phys", "KAMMA_eval", 32, 3, True, None),
    ("kaddpg_1d_nophys", "KA_DDPG_1D", 64, 4, False, None),        # BASELINE.json configs[0]
    ("kaddpg_1d_g05_nophys", "KA_DDPG_1D", 64, 3, False, 0.5),
    ("kaddpg_1d_g0_phys", "KA_DDPG_1D", 64, 3, True, 0.0),
    ("kaddpg_2d_nophys", "KA_DDPG_2D", 64, 4, False, None),
    ("kaddpg_2d_g05_phys", "KA_DDPG_2D", 64, 3, True, 0.5),
    ("kaddpg_4d_nophys", "KA_DDPG_4D", 64, 4, False, None),
    ("kamma_curr1_n1_a", "KAMMA_curr1", 1, 6, False, None),
    ("kamma_curr1_n1_b", "KAMMA_curr1", 1, 6, True, None),
    ("kamma_g1d_g21_nophys", "KAMMA_g1d_g21", 32, 3, False, None),
    ("kamma_curr2_n1_a", "KAMMA_curr2", 1, 6, False, None),        # appended: the seed of a case is 1234 + its index
    ("kamma_curr2_n1_b", "KAMMA_curr2", 1, 6, True, None),
    ("kaddpg_1d_eval_nophys", "KA_DDPG_1D_eval", 64, 4, False, None),   # eval/eval_kaddpg_1d.py
    ("kaddpg_1d_eval_phys", "KA_DDPG_1D_eval", 32, 3, True, None),
    ("kaddpg_2d_eval_nophys", "KA_DDPG_2D_eval", 64, 4, False, None),   # eval/eval_kaddpg_2d.py
    ("kaddpg_2d_eval_phys", "KA_DDPG_2D_eval", 32, 3, True, None),
    ("kamma_nophys", "KAMMA", 64, 4, False, None),                      # train/train_kamma_4d_fifo_random_RLIL.py
    ("kamma_g11_g21_nophys", "KAMMA_g11_g21", 32, 3, False, None),      # train/train_kamma_controller.py
    ("kamma_g11_g21_phys", "KAMMA_g11_g21", 32, 3, True, None),
    ("kaddpg_4d_phys", "KA_DDPG_4D", 64, 3, True, None),
]


def main():
    os.makedirs(GOLDEN, exist_ok=True)
    only = set(sys.argv[1:])            # optional: file tags to (re)generate; default all
    print("env cases (reference vs oracle restatement):")
    for i, (tag, name, n, steps, physics, fg) in enumerate(ENV_CASES):
        if only and tag not in only:
            continue
        rec = run_env_case(name, n, steps, physics=physics, seed=1234 + i, force_gamma=fg)
        np.savez_compressed(os.path.join(GOLDEN, f"env_{tag}.npz"), **rec)
    if not only and os.path.exists(os.path.join(ROOT, 'oracle', 'gen_golden_agent.py')):
        from oracle import gen_golden_agent

        gen_golden_agent.main()


if __name__ == "__main__":
    main()
