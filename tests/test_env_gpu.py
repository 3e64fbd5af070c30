"""GPU parity tests of the fused env kernel (csrc/env_step.cu) through the C ABI / RLTask mirror:
  * every golden vector recorded from the UNMODIFIED reference (tests/golden/env_*.npz);
  * the numpy oracle at the bench size (N = 4096) on seeded synthetic state;
  * the skid-steer sub-steps against the float64 twin (<= 1e-6);
  * size-independent properties at full size with the in-kernel Philox streams.
Tolerances: integers / bools bit-exact; fp32 tensor math <= 1e-5 relative (observed: bit-identical
without physics); dynamics <= 1e-6 absolute vs float64."""
import os

import numpy as np
import pytest

from conftest import GOLDEN, golden_files

pytestmark = pytest.mark.gpu
F = np.float32


def _task(variant, n, *, physics=True, gamma_const=None, seed=42):
    import options_in_simulation_b200 as ois
    from options_in_simulation_b200.config import script_cfg

    if not os.path.exists(ois.LIB_PATH):
        raise RuntimeError("libois_b200.so missing on the GPU box — build() must run before the snapshot")
    return ois.TorqueDistributionTask(sim_config=script_cfg(n), variant=variant, physics=physics,
                                      gamma_const=gamma_const, seed=seed)


def rel_err(ref, got):
    ref = np.asarray(ref, np.float64)
    return np.abs(ref - np.asarray(got, np.float64)) / np.maximum(1.0, np.abs(ref))


def ang_err(a, b):
    return np.abs(np.angle(np.exp(1j * (np.asarray(a, np.float64) - np.asarray(b, np.float64)))))


@pytest.mark.parametrize("fname", golden_files())
def test_env_kernel_matches_reference_golden(fname):
    import torch

    d = np.load(os.path.join(GOLDEN, fname), allow_pickle=False)
    name, n, steps, physics = str(d["variant"]), int(d["n"]), int(d["steps"]), bool(d["physics"])
    fg = float(d["force_gamma"])
    task = _task(name, n, physics=physics, gamma_const=None if fg < 0 else fg)
    task.load_state(**{k: d[f"s0_{k}"] for k in ("px", "py", "yaw", "vx", "vy", "wz", "w", "prev_lin", "prev_ang",
                                                 "commands", "sim_steps", "episode_buf", "episode_count", "sums",
                                                 "old_torques")})
    tol = 1e-5 if physics else 1e-6
    for t in range(steps):
        task.supply_draws(rand_mask=d[f"t{t}_rand_mask"], reset_uniforms=d[f"t{t}_reset_uniforms"])
        task.pre_physics_step(torch.as_tensor(d[f"t{t}_actions"]).cuda())
        obs, rew, reset, _ = task.post_physics_step()
        torch.cuda.synchronize()
        # integer / bool: bit-exact
        assert np.array_equal(reset.cpu().numpy().astype(np.uint8), d[f"t{t}_reset"]), f"reset_buf step {t}"
        assert np.array_equal(task.timeout_buf.cpu().numpy(), d[f"t{t}_timeout"])
        st = task.dump_state()
        for k in ("sim_steps", "episode_buf", "episode_count"):
            assert np.array_equal(st[k], d[f"t{t}_state_{k}"]), f"{k} step {t}"
        # fp32
        assert rel_err(d[f"t{t}_obs"], obs.cpu().numpy()).max() <= tol, f"obs step {t}"
        assert rel_err(d[f"t{t}_rew"], rew.cpu().numpy()).max() <= tol, f"rew step {t}"
        assert rel_err(d[f"t{t}_wheel_torqs"], task.wheel_torqs.cpu().numpy()).max() <= tol
        assert rel_err(d[f"t{t}_torques"], task.torques.cpu().numpy()).max() <= tol
        assert rel_err(d[f"t{t}_guiding"], task.guiding_reward.cpu().numpy()).max() <= tol
        for k in ("px", "py", "vx", "vy", "wz", "w", "prev_lin", "prev_ang", "commands"):
            assert rel_err(d[f"t{t}_state_{k}"], st[k]).max() <= tol, f"{k} step {t}"
        assert ang_err(d[f"t{t}_state_yaw"], st["yaw"]).max() <= 1e-5
        ns = d[f"t{t}_state_sums"].shape[1]
        assert rel_err(d[f"t{t}_state_sums"], st["sums"][:, :ns]).max() <= 2e-5, f"sums step {t}"


@pytest.mark.parametrize("name,n", [("KAMMA_medium", 4096), ("KA_DDPG_2D", 4096), ("KA_DDPG_1D", 64),
                                     ("KAMMA_eval", 8192), ("KAMMA_curr1", 16384), ("KAMMA_curr2", 4096), ("KAMMA_medium", 4097),
                                     ("KAMMA_medium", 1), ("KA_DDPG_1D_eval", 8192), ("KA_DDPG_2D_eval", 4096),
                                     ("KAMMA_g11_g21", 4096), ("KAMMA", 4096), ("KA_DDPG_4D", 4096)])
def test_env_kernel_matches_oracle_at_bench_sizes(name, n):
    """Seeded synthetic state, 6 closed-loop steps, physics on: CUDA fp32 vs numpy oracle
    (float64 dynamics).  Ragged N (4097, 1) exercises the grid tail."""
    import torch

    from oracle.env_oracle import VARIANTS, EnvOracle, EnvParams
    from oracle.gen_golden import action_sample, synth_state

    variant = VARIANTS[name]
    rng = np.random.default_rng(1234 + n)
    s0 = synth_state(variant, n, rng, physics=True)
    fg = 0.5 if name in ("KA_DDPG_1D", "KA_DDPG_2D") else None     # the *_eval twins keep their own frozen gamma = 0
    oracle = EnvOracle(variant, s0.copy(), EnvParams(physics=True))
    if fg is not None:
        oracle.force_gamma = F(fg)
    task = _task(name, n, physics=True, gamma_const=fg)
    task.stats_interval = 1
    task.load_state(**{k: getattr(s0, k) for k in ("px", "py", "yaw", "vx", "vy", "wz", "w", "prev_lin", "prev_ang",
                                                   "commands", "sim_steps", "episode_buf", "episode_count", "sums",
                                                   "old_torques")})
    n_resets = 0
    for t in range(6):
        act = action_sample(variant, n, rng)
        rm, ru = rng.random(n).astype(F), rng.random((n, 8)).astype(F)
        obs_o, rew_o, reset_o = oracle.step(act, rm, ru)
        task.supply_draws(rand_mask=rm, reset_uniforms=ru)
        obs, rew, reset, extras = task.step(torch.as_tensor(act).cuda())
        torch.cuda.synchronize()
        assert np.array_equal(reset.cpu().numpy().astype(bool), reset_o), f"reset set differs at step {t}"
        n_resets += int(reset_o.sum())
        st = task.dump_state()
        for k in ("sim_steps", "episode_buf", "episode_count"):
            assert np.array_equal(st[k], getattr(oracle.s, k)), k
        assert rel_err(obs_o, obs.cpu().numpy()).max() <= 1e-5
        assert rel_err(rew_o, rew.cpu().numpy()).max() <= 1e-5
        assert rel_err(oracle.last["wheel_torqs"], task.wheel_torqs.cpu().numpy()).max() <= 1e-5
        for k in ("px", "py", "vx", "vy", "wz", "w", "prev_lin", "prev_ang", "commands"):
            assert rel_err(getattr(oracle.s, k), st[k]).max() <= 1e-5, f"{k} step {t}"
        ns = st["sums"].shape[1]
        assert rel_err(oracle.s.sums[:, :ns], st["sums"]).max() <= 2e-5
        if reset_o.any():   # episode statistics (device-side reduction vs the reference's means)
            ep_o, ep = oracle.extras["episode"], extras["episode"]
            for k, v in ep_o.items():
                if k in ep and np.isfinite(v) and not k.endswith("/Final reward"):
                    assert abs(float(ep[k]) - float(v)) <= 1e-3 * max(1.0, abs(float(v))), k
    assert n_resets > 0 or n < 16


def test_dynamics_kernel_vs_float64_twin():
    """Row A4: `decimation` sub-steps under constant torques, fp32 kernel vs float64 CPU twin of the SAME
    equations: <= 1e-6 absolute on twist, <= 1e-5 on wheel speeds (|w| ~ 17 rad/s -> 1e-6 relative)."""
    import torch

    from oracle.dynamics_twin import SkidSteerParams, random_state, skid_step

    n = 8192
    rng = np.random.default_rng(7)
    st = {k: v.astype(F) for k, v in random_state(n, rng).items()}
    tau = rng.uniform(-4, 4, (n, 4)).astype(F)
    ref = skid_step({k: v.astype(np.float64) for k, v in st.items()}, tau.astype(np.float64), SkidSteerParams())
    task = _task("KAMMA_medium", n)
    task.load_state(px=st["px"], py=st["py"], yaw=st["yaw"], vx=st["vx"], vy=st["vy"], wz=st["wz"], w=st["w"])
    task.dynamics_step(torch.as_tensor(tau).cuda())
    torch.cuda.synchronize()
    got = task.dump_state()
    for k in ("vx", "vy", "wz"):
        assert np.abs(ref[k] - got[k]).max() <= 1e-6, (k, np.abs(ref[k] - got[k]).max())
    for k in ("px", "py"):
        assert np.abs(ref[k] - got[k]).max() <= 1e-6, k
    # heading is stored unwrapped in [0, 2*pi + ...): 1e-6 relative to max(1, |yaw|) (fp32 ulp at 6.28 is 4.8e-7)
    assert (ang_err(ref["yaw"], got["yaw"]) / np.maximum(1.0, np.abs(ref["yaw"]))).max() <= 1e-6
    assert (np.abs(ref["w"] - got["w"]) / np.maximum(1.0, np.abs(ref["w"]))).max() <= 1e-6


def test_dynamics_closed_loop_100_steps_tracks_twin():
    """Closed loop under the knowledge controller only (gamma = 1): after 50 policy steps the fp32 kernel is
    still within 1e-4 of the float64 oracle and the robot tracks the command band of the reference's logs
    (steady v_delta ~ 0.14 m/s at Kp = 0.035, my_runs/.../reward_components_env0_timeseries.csv)."""
    import torch

    from oracle.env_oracle import VARIANTS, EnvOracle, EnvParams, EnvState

    n = 256
    rng = np.random.default_rng(3)
    s0 = EnvState.zeros(n)
    s0.commands[:, 0] = rng.uniform(1, 2, n)
    oracle = EnvOracle(VARIANTS["KAMMA_g11_g21"], s0.copy(), EnvParams(physics=True))
    task = _task("KAMMA_g11_g21", n)
    task.load_state(commands=s0.commands)
    for t in range(50):
        act = np.zeros((n, 4), F)
        rm, ru = rng.random(n).astype(F), rng.random((n, 8)).astype(F)
        oracle.step(act, rm, ru)
        task.supply_draws(rm, ru)
        task.step(torch.as_tensor(act).cuda())
    st = task.dump_state()
    assert np.abs(st["vx"] - oracle.s.vx).max() <= 1e-4
    v_delta = s0.commands[:, 0] - st["vx"]
    assert 0.0 < np.median(v_delta) < 0.5


def test_philox_mode_properties_full_size():
    """Throughput mode (in-kernel Philox): size-independent properties at N = 65536 (config 5 total)."""
    import torch

    n = 65536
    task = _task("KAMMA_medium", n, seed=9)
    task.post_reset()
    torch.cuda.synchronize()
    st = task.dump_state()
    assert (st["episode_count"] == 1).all() and (st["episode_buf"] == 0).all()
    assert (st["commands"][:, 0] >= 1.0).all() and (st["commands"][:, 0] <= 2.0).all()
    assert (st["commands"][:, 2] == 0).all()
    assert abs(st["commands"][:, 0].mean() - 1.5) < 0.01
    assert (np.abs(st["px"]) <= 0.5).all() and (st["yaw"] >= 0).all() and (st["yaw"] <= 2 * np.pi + 1e-6).all()
    # Bernoulli switch: with sim_steps = max/2 the criteria torque is applied with probability gamma = 0.5
    task.load_state(sim_steps=np.full(n, 250000, np.int64))
    act = torch.full((n, 4), 0.123, device="cuda")
    task.step(act)
    torch.cuda.synchronize()
    tq = task.torques.cpu().numpy()
    policy = np.isclose(tq[:, 0], 0.123 * 3.0)
    assert abs(policy.mean() - 0.5) < 0.01
    # counters advance for everyone; different steps draw different masks
    task.step(act)
    torch.cuda.synchronize()
    policy2 = np.isclose(task.torques.cpu().numpy()[:, 0], 0.123 * 3.0)
    assert 0.45 < (policy != policy2).mean() < 0.55
    st = task.dump_state()
    assert (st["sim_steps"] == 250002).all() and (st["episode_buf"] == 2).all()
    # timeouts: every env resets at exactly episode_buf == 99 and draws a fresh command
    task.load_state(episode_buf=np.full(n, 98, np.int64))
    task.step(act)
    torch.cuda.synchronize()
    assert task.reset_buf.all().item()
    st2 = task.dump_state()
    assert (st2["episode_buf"] == 0).all() and (st2["episode_count"] == 2).all()
    assert (st2["sums"] == 0).all() and (st2["vx"] == 0).all()
    assert not np.array_equal(st2["commands"][:, 0], st["commands"][:, 0])


def test_reset_idx_explicit_ids_match_oracle():
    import torch

    from oracle.env_oracle import VARIANTS, EnvOracle, EnvParams
    from oracle.gen_golden import synth_state

    n = 512
    rng = np.random.default_rng(11)
    variant = VARIANTS["KAMMA_medium"]
    s0 = synth_state(variant, n, rng, physics=True)
    task = _task("KAMMA_medium", n)
    task.load_state(**{k: getattr(s0, k) for k in ("px", "py", "yaw", "vx", "vy", "wz", "w", "prev_lin", "prev_ang",
                                                   "commands", "sim_steps", "episode_buf", "episode_count", "sums")})
    ids = np.sort(rng.choice(n, 100, replace=False))
    ru = rng.random((n, 8)).astype(F)
    o = EnvOracle(variant, s0.copy(), EnvParams())
    g = np.ones(n, F)
    o._reset(ids, g, g, ru)
    task.supply_draws(reset_uniforms=ru)
    task.reset_idx(torch.as_tensor(ids).cuda())
    torch.cuda.synchronize()
    st = task.dump_state()
    for k in ("episode_buf", "episode_count", "sim_steps"):
        assert np.array_equal(st[k], getattr(o.s, k)), k
    for k in ("px", "py", "yaw", "vx", "vy", "wz", "w", "prev_lin", "prev_ang", "commands"):
        assert rel_err(getattr(o.s, k), st[k]).max() <= 1e-6, k
    assert rel_err(o.s.sums[:, :8], st["sums"]).max() <= 1e-6


def test_rltask_surface():
    """Names and shapes the trainers / wrapper read."""
    import torch

    import options_in_simulation_b200 as ois

    task = _task("KAMMA_medium", 128)
    env = ois.wrap_env(task)
    obs, info = env.reset()
    assert obs.shape == (128, 6) and obs.dtype == torch.float32
    a = torch.rand(128, 4, device="cuda")
    obs, rew, term, trunc, extras = env.step(a)
    assert rew.shape == (128, 1) and term.shape == (128, 1) and term.dtype == torch.bool
    assert trunc.shape == (128, 1) and not trunc.any()
    assert task.wheel_torqs.shape == (128, 4) and task.action_scale == 3.0
    assert task.get_observations()["robot_view"]["obs_buf"] is task.obs_buf
    assert task.calculate_metrics() is task.rew_buf
    assert task.num_envs == 128 and task.num_observations == 6 and task.num_actions == 4
    assert task.observation_space.shape == (6,) and task.action_space.shape == (4,)
    task.stats_interval = 1
    env.step(a)
    oc = task.observed_components
    assert len(oc) == 26 and "env0_v_delta" in oc
    with pytest.raises(ois.OisError):
        ois.RandomMemory(memory_size=4, num_envs=2, device="cpu")   # no CPU path


def test_intermediate_outputs_off_changes_nothing_else():
    """task.intermediate_outputs = False (bench.py, throughput runs) skips the stores of the three reference attributes nothing
    on the path reads (pre-clip torques, guiding_reward, timeout_buf); everything else — observations, rewards, resets,
    wheel_torqs, the whole env state, the commands of non-reset envs — is bit-identical, with the in-kernel Philox streams."""
    import torch

    n = 20_000
    tasks = [_task("KAMMA_medium", n, seed=11) for _ in range(2)]
    tasks[1].intermediate_outputs = False
    for t in tasks:
        t.post_reset()
    g = torch.Generator(device="cuda").manual_seed(5)
    resets = 0
    for step in range(104):     # past the first episode end (99 steps): the reset path stores new commands
        act = torch.rand(n, 4, device="cuda", generator=g)
        outs = [t.step(act.clone()) for t in tasks]
        if step % 13 == 0 or step >= 97:
            for a, b in zip(outs[0][:3], outs[1][:3]):
                assert torch.equal(a, b)
            assert torch.equal(tasks[0].wheel_torqs, tasks[1].wheel_torqs)
            sa, sb = tasks[0].dump_state(), tasks[1].dump_state()
            for k in sa:
                assert np.array_equal(sa[k], sb[k]), (k, step)
            resets += int(outs[0][2].sum())
    assert resets > 0
    assert float(tasks[0].torques.abs().sum()) > 0 and float(tasks[1].torques.abs().sum()) == 0
    assert float(tasks[1].guiding_reward.abs().sum()) == 0 and int(tasks[1].timeout_buf.sum()) == 0
