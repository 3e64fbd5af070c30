"""GPU tests of rollout.RolloutGraph: one policy step (DDPG.act -> env step with the fused replay append) replayed as a
CUDA graph whose per-step scalars live in a device-resident OisStepCtl (include/ois_b200.h).

The bar is bit-equality with the call-by-call path (which the other GPU tests hold to the oracle): observations,
rewards, reset flags, every replay tensor including the ring wrap, the OU noise state and the env state after the same
number of steps — with graph replays alone, interleaved with eager steps and DDPG updates, and through pinned host
buffers."""
import pytest

pytestmark = pytest.mark.gpu

T_TOTAL = 600000


def make_world(variant, n, precision, memory_steps=5, seed=7, t_e=5e5):
    import torch

    import options_in_simulation_b200 as ois
    from options_in_simulation_b200 import agent as A
    from options_in_simulation_b200.config import script_cfg

    torch.manual_seed(seed)
    dev = "cuda:0"
    task = ois.TorqueDistributionTask(sim_config=script_cfg(n, seed=seed, sim_device=dev), variant=variant, seed=seed)
    env = ois.wrap_env(task)
    osp, asp = task.observation_space, task.action_space
    models = {"policy": A.DeterministicActor(osp, asp, dev), "target_policy": A.DeterministicActor(osp, asp, dev),
              "critic": A.Critic(osp, asp, dev), "target_critic": A.Critic(osp, asp, dev)}
    memory = ois.RandomMemory(memory_size=memory_steps, num_envs=n, device=dev, replacement=False)
    cfg = {"batch_size": 512, "random_timesteps": 0, "learning_starts": 0, "mixed_precision": precision == "bf16",
           "sampler": "device", "graph_update": False,
           "state_preprocessor": A.RunningStandardScaler, "state_preprocessor_kwargs": {"size": osp, "device": dev},
           "exploration": {"noise": A.OrnsteinUhlenbeckNoise(0.15, 0.1, 0.3, device=dev),
                           "initial_scale": 1.0, "final_scale": 1e-4, "timesteps": t_e},
           "experiment": {"write_interval": 0, "checkpoint_interval": 0}}
    agent = A.DDPG(models=models, memory=memory, cfg=cfg, observation_space=osp, action_space=asp, device=dev)
    agent.init({"timesteps": T_TOTAL})
    agent.fuse_replay_with(task)
    obs, _ = env.reset()
    return dict(task=task, env=env, agent=agent, memory=memory, obs=obs)


def eager_step(w, t):
    actions = w["agent"].act(w["obs"], t, T_TOTAL)[0]
    w["obs"], rew, term, trunc, _ = w["env"].step(actions)
    w["agent"].record_transition(None, None, rew, w["obs"], term, trunc, None, t, T_TOTAL)
    return w["obs"], rew, term


def same(a, b):
    import torch

    return torch.equal(torch.nan_to_num(a.float(), nan=-77.0), torch.nan_to_num(b.float(), nan=-77.0))


def assert_worlds_equal(wa, wb):
    import torch

    for name in ("states", "actions", "rewards", "next_states", "terminated", "truncated"):
        assert same(wa["memory"].tensors[name], wb["memory"].tensors[name]), name
    assert wa["memory"].memory_index == wb["memory"].memory_index and wa["memory"].filled == wb["memory"].filled
    ta, tb = wa["task"], wb["task"]
    for name in ("_pose_vx", "_twist_prev", "dof_vel", "commands", "_sums_a", "_sums_b", "sim_steps", "episode_buf",
                 "episode_count", "obs_buf", "wheel_torqs"):
        assert torch.equal(getattr(ta, name), getattr(tb, name)), name
    assert torch.equal(wa["agent"]._exploration_noise.state, wb["agent"]._exploration_noise.state)
    assert ta._launches == tb._launches and wa["agent"]._act_calls == wb["agent"]._act_calls


@pytest.mark.parametrize("variant,n,precision", [("KAMMA_medium", 1000, "bf16"), ("KAMMA_medium", 4096, "fp32"),
                                                 ("KA_DDPG_2D", 333, "bf16"), ("KAMMA_eval", 640, "bf16"),
                                                 ("KA_DDPG_1D", 64, "fp32")])
def test_graph_replay_is_bit_identical_to_eager_steps(variant, n, precision):
    """12 steps over a 5-slot ring (wraps twice); the exploration horizon is crossed on the way (t_e = 1006), so the
    in-kernel noise-scale derivation is compared on both sides of ddpg.py:247."""
    import torch

    from options_in_simulation_b200.rollout import RolloutGraph

    wa = make_world(variant, n, precision, t_e=1006)
    wb = make_world(variant, n, precision, t_e=1006)
    g = RolloutGraph(wb["env"], wb["agent"], timestep=1000, timesteps=T_TOTAL)
    for k in range(12):
        oa, ra, da = eager_step(wa, 1000 + k)
        ob, rb, db = g.step()
        torch.cuda.synchronize()
        assert torch.equal(oa, ob) and torch.equal(ra, rb) and torch.equal(da, db), k
    assert g._graphs
    assert_worlds_equal(wa, wb)


def test_graph_replay_interleaves_with_eager_steps_and_updates():
    """graph x3, eager x2, two DDPG updates (the bf16 weight image must follow), graph x4, update, graph x2."""
    import torch

    from options_in_simulation_b200.rollout import RolloutGraph

    wa = make_world("KAMMA_medium", 1024, "bf16", memory_steps=4)
    wb = make_world("KAMMA_medium", 1024, "bf16", memory_steps=4)
    g = RolloutGraph(wb["env"], wb["agent"], timestep=1000, timesteps=T_TOTAL)
    plan = ["g", "g", "g", "e", "e", "u", "u", "g", "g", "g", "g", "u", "g", "g"]
    t = 1000
    for op in plan:
        if op == "u":
            wa["agent"]._update(t, T_TOTAL)
            wb["agent"]._update(t, T_TOTAL)
            continue
        oa, ra, da = eager_step(wa, t)
        if op == "g":
            g.timestep = t
            ob, rb, db = g.step()
        else:
            ob, rb, db = eager_step(wb, t)
        wb["obs"] = ob
        torch.cuda.synchronize()
        assert torch.equal(oa, ob) and torch.equal(ra, rb) and torch.equal(da, db), (op, t)
        t += 1
    assert torch.equal(wa["agent"].policy.flat, wb["agent"].policy.flat)
    assert_worlds_equal(wa, wb)


@pytest.mark.parametrize("mode", [True, "copy"], ids=["zero-copy", "copy-nodes"])
def test_graph_replay_with_host_buffers(mode):
    """host_io: observations in through pinned host memory, (obs | rew | reset) out to pinned host memory — read and
    written in place by the kernels (zero-copy) or through an H2D and ONE packed D2H copy node."""
    import torch

    from options_in_simulation_b200.rollout import RolloutGraph

    wa = make_world("KAMMA_medium", 4096, "bf16")
    wb = make_world("KAMMA_medium", 4096, "bf16")
    g = RolloutGraph(wb["env"], wb["agent"], timestep=1000, timesteps=T_TOTAL, host_io=mode)
    g.h_states.copy_(wb["obs"])
    for k in range(8):
        oa, ra, da = eager_step(wa, 1000 + k)
        ob, rb, db = g.step()
        assert torch.equal(oa.cpu(), ob) and torch.equal(ra.cpu(), rb) and torch.equal(da.cpu(), db), k
        g.h_states.copy_(ob)
    assert_worlds_equal(wa, wb)


def test_graph_refuses_what_it_cannot_replay():
    from options_in_simulation_b200 import _lib as L
    from options_in_simulation_b200.rollout import RolloutGraph

    w = make_world("KAMMA_medium", 256, "bf16")
    w["agent"]._random_timesteps = 2000
    with pytest.raises(L.OisError):
        RolloutGraph(w["env"], w["agent"], timestep=1000, timesteps=T_TOTAL)


@pytest.mark.parametrize("precision", ["bf16", "fp32"])
def test_trainer_with_graph_rollout_equals_call_by_call_trainer(precision):
    """SequentialTrainer(cfg graph_rollout) against the same trainer issuing the step call by call: 10 timesteps with the
    reference cadence (two updates per timestep, base_KA_DDPG.py:265,283) leave bit-identical weights and replay rows."""
    import torch

    from options_in_simulation_b200.trainer import SequentialTrainer

    worlds = []
    for graph in (False, True):
        w = make_world("KAMMA_medium", 512, precision, memory_steps=6)
        w["env"]._reset_once = False
        w["env"]._obs = w["obs"]
        tr = SequentialTrainer(env=w["env"], agents=w["agent"],
                               cfg={"timesteps": 10, "headless": True, "disable_progressbar": True, "graph_rollout": graph})
        tr.train()
        torch.cuda.synchronize()
        worlds.append(w)
    wa, wb = worlds
    for name in ("policy", "critic", "target_policy", "target_critic"):
        assert torch.equal(getattr(wa["agent"], name).flat, getattr(wb["agent"], name).flat), name
    assert wa["agent"]._updates == wb["agent"]._updates == 20
    assert_worlds_equal(wa, wb)


@pytest.mark.parametrize("variant,n", [("KAMMA_medium", 4096), ("KAMMA_medium", 1000), ("KA_DDPG_2D", 333), ("KA_DDPG_1D", 64),
                                       ("KAMMA_medium", 32 * 148 + 5)])
def test_fused_step_is_bit_identical_to_act_then_env_step(variant, n):
    """rollout.fused_step: actor + env step (+ replay append) as ONE kernel (csrc/mlp_tct.cu, kEnv) against the three
    separate calls, 11 steps over a 5-slot ring with an update in between; the launch counter proves which path ran
    (the last size is one env too many for one 32-env tile per CTA: it must fall back to the two launches)."""
    import ctypes

    import torch

    import options_in_simulation_b200 as ois
    from options_in_simulation_b200.rollout import fused_step

    lib = ois.lib()
    lib.ois_debug_tct_fused_launches.restype = ctypes.c_longlong
    wa = make_world(variant, n, "bf16", t_e=1006)
    wb = make_world(variant, n, "bf16", t_e=1006)
    before = lib.ois_debug_tct_fused_launches()
    for k in range(11):
        oa, ra, da = eager_step(wa, 1000 + k)
        ob, rb, db, _, _ = fused_step(wb["env"], wb["agent"], wb["obs"], 1000 + k, T_TOTAL)
        wb["obs"] = ob
        torch.cuda.synchronize()
        assert torch.equal(oa, ob) and torch.equal(ra, rb) and torch.equal(da, db), k
        if k == 5 and n * 5 >= 512:
            for w in (wa, wb):
                w["agent"]._update(1000 + k, T_TOTAL)
    fused = lib.ois_debug_tct_fused_launches() - before
    assert fused == (11 if n <= 32 * 148 else 0), fused
    assert_worlds_equal(wa, wb)
