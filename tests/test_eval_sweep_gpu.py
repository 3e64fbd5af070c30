"""GPU test of evaluate.tracking_error_sweep (BASELINE.json configs[4]: KAMMA-4d eval rollout, TE vs speed).

The CUDA path (fused env kernel + fp32 actor, SequentialTrainer.eval, device-side statistics log, binning) is compared
with the CPU oracle run in the same closed loop: numpy env oracle of src/my_envs/KAMMA_eval.py (float64 dynamics twin)
driven by the CPU-torch actor oracle with the same weights and the same supplied uniforms.  Bar: the steps at which
episodes end and the number of episodes are identical (integer work, bit-exact); the logged pairs
(rew_Desired velocity, rew_Tracking error) agree to 2e-3 relative after 230 closed-loop steps; the binned curve to 2e-3."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu
F = np.float32
N, STEPS = 96, 230


def test_tracking_error_sweep_matches_oracle_loop():
    import torch

    import options_in_simulation_b200 as ois
    from options_in_simulation_b200 import agent as A
    from options_in_simulation_b200.config import script_cfg
    from options_in_simulation_b200.evaluate import tracking_error_sweep
    from oracle.agent_oracle import AgentOracle, load_flat
    from oracle.env_oracle import VARIANTS, EnvOracle, EnvParams
    from oracle.gen_golden import synth_state

    variant = VARIANTS["KAMMA_eval"]
    rng = np.random.default_rng(99)
    s0 = synth_state(variant, N, rng, physics=True)
    s0.sim_steps[:] = rng.integers(0, 50000, N)          # spread over the 1 -> 2 m/s ramp (KAMMA_eval.py:545-550)
    s0.episode_buf[:] = rng.integers(0, 99, N)           # episodes end at different steps
    draws = [(rng.random(N).astype(F), rng.random((N, 8)).astype(F)) for _ in range(STEPS + 1)]

    # ---- CUDA path
    task = ois.TorqueDistributionTask(sim_config=script_cfg(N), variant="KAMMA_eval", physics=True, seed=5)
    task.init_done = True
    task.load_state(**{k: getattr(s0, k) for k in ("px", "py", "yaw", "vx", "vy", "wz", "w", "prev_lin", "prev_ang",
                                                   "commands", "sim_steps", "episode_buf", "episode_count", "sums",
                                                   "old_torques")})
    env = ois.wrap_env(task)
    osp, asp = task.observation_space, task.action_space
    torch.manual_seed(11)
    models = {"policy": A.DeterministicActor(osp, asp, "cuda:0"), "target_policy": A.DeterministicActor(osp, asp, "cuda:0"),
              "critic": A.Critic(osp, asp, "cuda:0"), "target_critic": A.Critic(osp, asp, "cuda:0")}
    memory = ois.RandomMemory(memory_size=4, num_envs=N, device="cuda:0")
    cfg = {"batch_size": 32, "random_timesteps": 0, "learning_starts": 0,
           "state_preprocessor": A.RunningStandardScaler, "state_preprocessor_kwargs": {"size": osp, "device": "cuda:0"},
           "exploration": {"noise": None}, "experiment": {"write_interval": 0, "checkpoint_interval": 0}}
    agent = A.DDPG(models=models, memory=memory, cfg=cfg, observation_space=osp, action_space=asp, device="cuda:0")

    def hook(trainer, t):       # env step k = t + 1 (k = 0 is the wrapper's reset step)
        task.supply_draws(rand_mask=draws[t + 1][0], reset_uniforms=draws[t + 1][1])

    task.supply_draws(rand_mask=draws[0][0], reset_uniforms=draws[0][1])
    out = tracking_error_sweep(env, agent, STEPS, lo=1.0, hi=2.0, bins=10, pre_step_hook=hook)

    # ---- oracle loop
    oracle = EnvOracle(variant, s0.copy(), EnvParams(physics=True))
    lo_a, hi_a = float(asp.low[0]), float(asp.high[0])
    ao = AgentOracle(6, 4, "sigmoid", ou=None, clip=(lo_a, hi_a))
    load_flat(ao.policy, agent.policy.flat.cpu())
    rows = []

    def env_step(k, a):
        o, _, reset = oracle.step(a, *draws[k])
        if reset.any():
            ep = oracle.extras["episode"]
            rows.append((k, float(ep["rew_Desired velocity"]), float(ep["rew_Tracking error"]), int(reset.sum())))
        return o

    obs = env_step(0, np.zeros((N, 4), F))                       # VecEnvRLGames.reset: one step with zero actions
    for t in range(STEPS):
        obs = env_step(t + 1, ao.act(np.clip(obs, -1000, 1000), t, STEPS).numpy())
    ref = np.asarray(rows, np.float64)

    got = out["series"].numpy()
    # the log's row k belongs to env step k (row 0 = the wrapper's reset step)
    assert got.shape == ref.shape, (got.shape, ref.shape)
    assert np.array_equal(got[:, 0], ref[:, 0]) and np.array_equal(got[:, 3], ref[:, 3])
    assert ref[:, 3].sum() >= 2 * N                      # every env ended at least two episodes
    assert np.abs(got[:, 1] - ref[:, 1]).max() <= 2e-3 * ref[:, 1].max()
    assert np.abs(got[:, 2] - ref[:, 2]).max() <= 2e-3 * max(1.0, ref[:, 2].max())
    # binned curve
    idx = np.clip(np.floor((ref[:, 1] - 1.0) / 1.0 * 10), 0, 9).astype(int)
    te_sum, cnt = np.zeros(10), np.zeros(10)
    np.add.at(te_sum, idx, ref[:, 2] * ref[:, 3])
    np.add.at(cnt, idx, ref[:, 3])
    assert np.array_equal(out["episodes"].numpy(), cnt)
    assert np.abs(out["tracking_error"].numpy() - te_sum / np.maximum(cnt, 1)).max() <= 2e-3
    assert (cnt > 0).sum() >= 5                          # the sweep covers the speed range
