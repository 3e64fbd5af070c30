"""GPU test of SURVEY.md §8 row F2: a run of ``SequentialTrainer`` with ``write_interval > 0`` leaves a TensorBoard
event file holding the tags the reference's artefact tools search for —

  * ``Reward_comp_env0 / {v_delta, desired_v, current_v, omega_delta, desired_omega, current_omega}`` (+ the other 20
    env-0 scalars), src/options_sim/cli/artifacts.py:23-30 <- src/my_trainers/base_KA_DDPG.py:229-254;
  * ``Info / rew_Tracking error`` and ``Info / rew_Desired velocity``, src/options_sim/cli/eval_artifacts.py:20-31
    <- src/my_envs/KAMMA_eval.py:488-493 via base_KA_DDPG.py:337-340;

and that the two eval series pair up step by step into the TE-vs-speed table ``options-sim-eval-artifacts`` builds
(eval_artifacts.py:65-174: first tag candidate present, align by step).  The event file is read back with TensorBoard's
own ``EventAccumulator``, the reader those tools use."""
import glob
import os

import numpy as np
import pytest

pytestmark = pytest.mark.gpu

ARTIFACT_TAGS = ["Reward_comp_env0 / v_delta", "Reward_comp_env0 / desired_v", "Reward_comp_env0 / current_v",
                 "Reward_comp_env0 / omega_delta", "Reward_comp_env0 / desired_omega", "Reward_comp_env0 / current_omega"]
TRACKING_ERR_CANDS = ["Info / rew_Tracking error", "Info/rew_Tracking error", "Info/rev_Tracking error", "tracking_error"]
DESIRED_VEL_CANDS = ["Info / rew_Desired velocity", "Info/rew_Desired velocity", "desired_velocity", "command/vel_x"]


def _world(variant, n, tmp_path, write_interval, noise=True):
    import torch

    import options_in_simulation_b200 as ois
    from options_in_simulation_b200 import agent as A
    from options_in_simulation_b200.config import script_cfg

    task = ois.TorqueDistributionTask(sim_config=script_cfg(n), variant=variant, seed=5)
    env = ois.wrap_env(task)
    osp, asp = task.observation_space, task.action_space
    torch.manual_seed(5)
    models = {"policy": A.DeterministicActor(osp, asp, "cuda:0"), "target_policy": A.DeterministicActor(osp, asp, "cuda:0"),
              "critic": A.Critic(osp, asp, "cuda:0"), "target_critic": A.Critic(osp, asp, "cuda:0")}
    memory = ois.RandomMemory(memory_size=64, num_envs=n, device="cuda:0", replacement=False)
    cfg = {"batch_size": 128, "discount_factor": 0.999, "polyak": 0.005, "actor_learning_rate": 3e-4,
           "critic_learning_rate": 1e-3, "random_timesteps": 0, "learning_starts": 4,
           "state_preprocessor": A.RunningStandardScaler, "state_preprocessor_kwargs": {"size": osp, "device": "cuda:0"},
           "exploration": {"noise": A.OrnsteinUhlenbeckNoise(0.15, 0.1, 0.05, device="cuda:0") if noise else None,
                           "final_scale": 1e-4, "timesteps": 1000},
           "experiment": {"directory": str(tmp_path), "experiment_name": "run", "write_interval": write_interval,
                          "checkpoint_interval": 0}}
    agent = A.DDPG(models=models, memory=memory, cfg=cfg, observation_space=osp, action_space=asp, device="cuda:0")
    return task, env, agent


def _read_events(run_dir):
    from tensorboard.backend.event_processing.event_accumulator import EventAccumulator

    assert glob.glob(os.path.join(run_dir, "events.out.tfevents.*")), "no TensorBoard event file was written"
    ea = EventAccumulator(run_dir, size_guidance={"scalars": 0})
    ea.Reload()
    return ea


def _first_series(ea, candidates):
    """eval_artifacts.py:_get_series: the first candidate tag that exists."""
    tags = set(ea.Tags().get("scalars", []))
    for c in candidates:
        if c in tags:
            ev = ea.Scalars(c)
            return c, np.array([e.step for e in ev]), np.array([e.value for e in ev])
    return None, None, None


def test_train_run_writes_the_env0_component_tags(tmp_path):
    import torch

    from options_in_simulation_b200.trainer import SequentialTrainer

    task, env, agent = _world("KAMMA_medium", 64, tmp_path, write_interval=10)
    trainer = SequentialTrainer(env=env, agents=agent, cfg={"timesteps": 40, "headless": True, "disable_progressbar": True})
    env.reset()
    task.episode_buf.copy_(98 - torch.arange(64, device="cuda") % 45)  # an episode ends on every one of the 40 steps
    trainer.train()
    agent.writer.flush()
    ea = _read_events(os.path.join(str(tmp_path), "run"))
    tags = set(ea.Tags()["scalars"])
    for t in ARTIFACT_TAGS:                                            # what options-sim-artifacts plots
        assert t in tags, f"missing {t}"
        assert len(ea.Scalars(t)) == 4 and [e.step for e in ea.Scalars(t)] == [10, 20, 30, 40]
    assert len([t for t in tags if t.startswith("Reward_comp_env0 / ")]) == 26      # base_KA_DDPG.py:229-254
    for t in ("Loss / Critic loss", "Loss / Policy loss", "Q-network / Q1 (mean)", "Target / Target (mean)",
              "Exploration / Exploration noise (max)", "Reward / Instantaneous reward (mean)"):      # ddpg.py:258-266,452-461
        assert t in tags, f"missing {t}"
    assert any(t.startswith("Info / rew_") for t in tags) and "Info / gamma_1 action assist" in tags
    v = np.array([e.value for e in ea.Scalars("Reward_comp_env0 / desired_v")])
    assert np.all((v >= 0.0) & (v <= 2.0))                             # commands are drawn from linear_x = [1, 2]


def test_eval_run_feeds_the_te_vs_speed_artifact(tmp_path):
    from options_in_simulation_b200.evaluate import phase_lock
    from options_in_simulation_b200.trainer import SequentialTrainer

    task, env, agent = _world("KAMMA_eval", 198, tmp_path, write_interval=20, noise=False)
    trainer = SequentialTrainer(env=env, agents=agent, cfg={"timesteps": 240, "headless": True, "disable_progressbar": True})
    env.reset()
    phase_lock(task)                                                   # two envs finish an episode on every step
    before = agent.policy.flat.clone()
    trainer.eval()
    agent.writer.flush()
    import torch

    assert torch.equal(before, agent.policy.flat)                      # eval: base post_interaction only (no _update)
    ea = _read_events(os.path.join(str(tmp_path), "run"))
    te_tag, te_steps, te = _first_series(ea, TRACKING_ERR_CANDS)
    dv_tag, dv_steps, dv = _first_series(ea, DESIRED_VEL_CANDS)
    assert te_tag == "Info / rew_Tracking error" and dv_tag == "Info / rew_Desired velocity"
    # eval_artifacts.py:_align_by_step: inner join on the step -> rows of tracking_error_vs_speed.csv
    common = np.intersect1d(te_steps, dv_steps)
    assert len(common) >= 10 and np.array_equal(te_steps, dv_steps)
    rows = np.stack([common, dv[np.isin(dv_steps, common)], te[np.isin(te_steps, common)]], axis=1)
    full = rows[rows[:, 0] > 120]                                      # past the partial first episodes
    assert np.all(np.isfinite(rows)) and np.all(rows[:, 2] >= 0.0)
    assert len(full) >= 4 and np.all((full[:, 1] >= 0.98) & (full[:, 1] <= 2.0))         # the 1 -> 2 m/s ramp of KAMMA_eval.py:545-550
