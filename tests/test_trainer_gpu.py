"""GPU: options_in_simulation_b200.trainer.SequentialTrainer (act -> fused env kernel -> relabel -> record ->
two updates per step) replays the vectors recorded from the UNMODIFIED reference trainer + task + agent
(tests/golden/trainer_*.npz), with the reference's random draws supplied to the kernels.  Both the fused replay
append (env kernel epilogue) and the separate ``ois_replay_append`` path must reproduce the reference's memory.
Tolerances: indices / terminated / ring position bit-exact; stored transitions 2e-5 relative (fp32 dynamics vs
the float64 twin the reference run used); weights after 12-14 fp32 updates 5e-5 absolute."""
import numpy as np
import pytest

from conftest import golden_files
from trainer_cases import STATE_KEYS, load_case

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("fused", [True, False], ids=["fused-append", "append-kernel"])
@pytest.mark.parametrize("fname", golden_files("trainer_"))
def test_sequential_trainer_matches_reference(fname, fused):
    import torch

    import options_in_simulation_b200 as ois
    from options_in_simulation_b200 import agent as A
    from options_in_simulation_b200.config import script_cfg
    from options_in_simulation_b200.trainer import SequentialTrainer

    d, m, draws = load_case(fname)
    n, T = m["n"], m["timesteps"]
    task = ois.TorqueDistributionTask(sim_config=script_cfg(n), variant=m["variant"], seed=1)
    task.init_done = True                       # the golden run starts from a loaded state, not from post_reset
    task.load_state(**{k: d[f"s0_{k}"] for k in STATE_KEYS if k != "old_torques"})
    env = ois.wrap_env(task)
    osp, asp = task.observation_space, task.action_space
    assert float(asp.low[0]) == m["clip_low"]
    torch.manual_seed(m["seed"])
    models = {"policy": A.DeterministicActor(osp, asp, "cuda:0"), "target_policy": A.DeterministicActor(osp, asp, "cuda:0"),
              "critic": A.Critic(osp, asp, "cuda:0"), "target_critic": A.Critic(osp, asp, "cuda:0")}
    memory = ois.RandomMemory(memory_size=m["mem_steps"], num_envs=n, device="cuda:0", replacement=False)
    cfg = {"batch_size": m["batch"], "discount_factor": 0.999, "polyak": 0.005, "actor_learning_rate": 3e-4,
           "critic_learning_rate": 1e-3, "random_timesteps": m["random_timesteps"], "learning_starts": m["learning_starts"],
           "state_preprocessor": A.RunningStandardScaler, "state_preprocessor_kwargs": {"size": osp, "device": "cuda:0"},
           "exploration": {"noise": A.OrnsteinUhlenbeckNoise(0.15, 0.1, m["ou_scale"], device="cuda:0"),
                           "final_scale": 1e-4, "timesteps": m["T_e"]},
           "experiment": {"write_interval": 0, "checkpoint_interval": 0}}
    agent = A.DDPG(models=models, memory=memory, cfg=cfg, observation_space=osp, action_space=asp, device="cuda:0")
    if fused:
        agent.fuse_replay_with(task)
    memory.supply_indexes(d["indexes"])
    acts = []

    def hook(trainer, t):       # env step k = t + 1 (k = 0 is the wrapper's reset step)
        agent.supply_draws(normals=draws["normals"][t], uniforms=draws["uniforms"][t])
        task.supply_draws(rand_mask=draws["rand_mask"][t + 1], reset_uniforms=draws["reset_uniforms"][t + 1])
        if t > 0:
            acts.append(task.actions.cpu().numpy().copy())

    task.supply_draws(rand_mask=draws["rand_mask"][0], reset_uniforms=draws["reset_uniforms"][0])
    trainer = SequentialTrainer(cfg={"timesteps": T, "headless": True, "disable_progressbar": True,
                                     "pre_step_hook": hook}, env=env, agents=agent)
    assert trainer.kind == ("KA_DDPG" if m["relabel"] else "KA")
    trainer.train()
    torch.cuda.synchronize()
    acts.append(task.actions.cpu().numpy().copy())

    assert agent._updates == m["n_updates"]
    assert memory.memory_index == m["memory_index"] and memory.filled == bool(m["filled"])
    for t in range(T):
        assert np.abs(acts[t] - d["actions"][t]).max() <= 1e-5, t
    assert np.array_equal(memory.tensors_view["terminated"].cpu().numpy().astype(np.uint8).reshape(-1, 1), d["mem_terminated"])
    assert not memory.tensors_view["truncated"].any().item()
    for name in ("states", "actions", "rewards", "next_states"):
        got, ref = memory.tensors_view[name].cpu().numpy(), d[f"mem_{name}"]
        assert np.array_equal(np.isnan(got), np.isnan(ref)), name
        assert np.nanmax(np.abs(got - ref)) <= 2e-5 * max(1.0, np.nanmax(np.abs(ref))), name
    last = T - 1
    assert np.array_equal(task.reset_buf.cpu().numpy().astype(np.uint8), d["reset"][last])
    assert np.abs(task.wheel_torqs.cpu().numpy() - d["wheel_torqs"][last]).max() <= 1e-5
    for name, model in (("policy", agent.policy), ("critic", agent.critic), ("target_policy", agent.target_policy),
                        ("target_critic", agent.target_critic)):
        flat = model.flat.cpu()
        assert np.abs(flat[::m["stride"]].numpy() - d[f"{name}_sub"]).max() <= 5e-5, name
        assert flat.double().sum().item() == pytest.approx(float(d[f"{name}_sum"]), abs=5e-3), name
    sc = agent._state_preprocessor
    assert np.allclose(sc.running_mean.cpu().numpy(), d["scaler_mean"], rtol=1e-5, atol=1e-7)
    assert np.allclose(sc.running_variance.cpu().numpy(), d["scaler_var"], rtol=1e-5, atol=1e-7)
    assert float(sc.current_count.item()) == float(d["scaler_count"])


def test_eval_loop_does_not_update():
    import torch

    import options_in_simulation_b200 as ois
    from options_in_simulation_b200 import agent as A
    from options_in_simulation_b200.config import script_cfg
    from options_in_simulation_b200.trainer import SequentialTrainer

    n = 64
    task = ois.TorqueDistributionTask(sim_config=script_cfg(n), variant="KAMMA_eval", seed=3)
    env = ois.wrap_env(task)
    osp, asp = task.observation_space, task.action_space
    models = {"policy": A.DeterministicActor(osp, asp, "cuda:0", output="tanh"),
              "target_policy": A.DeterministicActor(osp, asp, "cuda:0", output="tanh"),
              "critic": A.Critic(osp, asp, "cuda:0"), "target_critic": A.Critic(osp, asp, "cuda:0")}
    memory = ois.RandomMemory(memory_size=16, num_envs=n, device="cuda:0")
    cfg = {"batch_size": 32, "random_timesteps": 0, "learning_starts": 0,
           "exploration": {"noise": A.OrnsteinUhlenbeckNoise(0.15, 0.1, 0.0, device="cuda:0"), "timesteps": 5e4},
           "experiment": {"write_interval": 0, "checkpoint_interval": 0}}
    agent = A.DDPG(models=models, memory=memory, cfg=cfg, observation_space=osp, action_space=asp, device="cuda:0")
    before = agent.policy.flat.clone()
    SequentialTrainer(cfg={"timesteps": 20, "headless": True, "disable_progressbar": True}, env=env, agents=agent).eval()
    torch.cuda.synchronize()
    assert agent._updates == 0 and torch.equal(before, agent.policy.flat)
    assert memory.memory_index == 4 and memory.filled            # record_transition still writes (base_KA_DDPG.py:320-329)
    assert int(task.sim_steps.min().item()) == 21                # reset step + 20
