"""GPU parity tests of the DDPG agent kernels (csrc/agent.cu, mlp_fp32.cu, mlp_tc.cu) through the
skrl-Agent mirror (options_in_simulation_b200/agent.py):
  * golden vectors recorded from the UNMODIFIED reference my_agents/ddpg.py (tests/golden/agent_*.npz);
  * oracle/agent_oracle.py at the bench sizes;
  * bf16 tcgen05 actor inference vs the fp32 oracle under the stated bf16 tolerance.
Tolerances (fp32 path): actions / Q statistics / updated weights <= 1e-5 relative to max(1, |ref|) —
GEMM summation order differs from CPU torch, so results are not bit-identical; scaler state (fp64) <= 1e-9.
Scaler statistics: the reference feeds float32 torch.mean / torch.var results into its fp64 running state, so
its state carries fp32 reduction-order noise (~1e-7 relative); the kernel reduces in fp64 and rounds once ->
compared at 1e-6 relative.
bf16 path: |action - fp32 oracle| <= 2e-2 absolute on the sigmoid/tanh output (bf16 has 8 mantissa bits:
three 512/128-wide bf16 layers give ~3e-3 typical, 2e-2 worst case; fp32 accumulation)."""
import numpy as np
import pytest

from agent_cases import BATCH_NAMES, NETS, STAT_NAMES, load, make_oracle
from conftest import golden_files

pytestmark = pytest.mark.gpu
BF16_ATOL = 2e-2


class _Box:
    def __init__(self, low, high, n):
        self.shape = (n,)
        self.low = np.full(n, low, np.float32)
        self.high = np.full(n, high, np.float32)


def build_agent(meta, *, mixed_precision=False, n_envs=None, sampler="host", rows=None, **extra_cfg):
    import torch

    import options_in_simulation_b200 as ois
    from options_in_simulation_b200 import agent as A

    obs_space, act_space = _Box(-1, 1, meta["obs_dim"]), _Box(meta["clip_low"], 1.0, meta["act_dim"])
    torch.manual_seed(meta["seed"])
    models = {"policy": A.DeterministicActor(obs_space, act_space, "cuda:0", output=meta["out_act"]),
              "target_policy": A.DeterministicActor(obs_space, act_space, "cuda:0", output=meta["out_act"]),
              "critic": A.Critic(obs_space, act_space, "cuda:0"), "target_critic": A.Critic(obs_space, act_space, "cuda:0")}
    n_envs = n_envs or meta["n_envs"]
    rows = rows or meta["rows"]
    memory = ois.RandomMemory(memory_size=rows // n_envs, num_envs=n_envs, device="cuda:0", replacement=False)
    cfg = {"batch_size": meta["batch"], "discount_factor": meta["discount"], "polyak": meta["polyak"],
           "actor_learning_rate": meta["lr_a"], "critic_learning_rate": meta["lr_c"], "random_timesteps": 10,
           "learning_starts": 10, "mixed_precision": mixed_precision, "sampler": sampler,
           "state_preprocessor": A.RunningStandardScaler,
           "state_preprocessor_kwargs": {"size": obs_space, "device": "cuda:0"},
           "exploration": {"noise": A.OrnsteinUhlenbeckNoise(0.15, 0.1, meta["base_scale"], device="cuda:0"),
                           "final_scale": 1e-4, "timesteps": meta["T_e"]},
           "experiment": {"write_interval": 0, "checkpoint_interval": 0}, "track_stats_interval": 1}
    cfg.update(extra_cfg)
    agent = A.DDPG(models=models, memory=memory, cfg=cfg, observation_space=obs_space, action_space=act_space,
                   device="cuda:0")
    agent.init({"timesteps": 1000})
    return agent, memory


def rel(ref, got):
    ref = np.asarray(ref, np.float64)
    return (np.abs(ref - np.asarray(got, np.float64)) / np.maximum(1.0, np.abs(ref))).max()


@pytest.mark.parametrize("fname", golden_files("agent_"))
def test_ddpg_fp32_matches_reference_golden(fname):
    import torch

    d, meta = load(fname)
    agent, memory = build_agent(meta)
    assert agent.policy.flat.double().sum().item() == pytest.approx(meta["init_policy_sum"], abs=1e-6)
    assert agent.critic.flat.double().sum().item() == pytest.approx(meta["init_critic_sum"], abs=1e-6)
    n = meta["n_envs"]
    for t in range(meta["rows"] // n):
        sl = slice(t * n, (t + 1) * n)
        memory.add_samples(**{k: torch.as_tensor(d[f"mem_{k}"][sl]).cuda().to(torch.bool if k in ("terminated", "truncated") else torch.float32)
                              for k in BATCH_NAMES})
    assert memory.filled and len(memory) == meta["rows"]
    # random phase: U[low, high] from the supplied uniforms — same fp32 expression, bit-exact
    agent.supply_draws(uniforms=d["rand_uniforms"])
    ra = agent.act(torch.as_tensor(d["rand_states"]).cuda(), 3, 1000)[0]
    assert np.array_equal(ra.cpu().numpy(), d["rand_actions"])
    orig = torch.randperm
    for k in range(meta["rounds"]):
        t = int(d[f"r{k}_timestep"])
        agent.supply_draws(normals=d[f"r{k}_normals"])
        a = agent.act(torch.as_tensor(d[f"r{k}_states"]).cuda(), t, 1000)[0]
        assert rel(d[f"r{k}_actions"], a.cpu().numpy()) <= 1e-5, f"actions round {k}"
        assert np.abs(agent._exploration_noise.state.cpu().numpy() - d[f"r{k}_ou_state"]).max() <= 1e-7
        idx = torch.as_tensor(d[f"r{k}_indices"])
        torch.randperm = lambda n_, **kw: torch.cat([idx, torch.zeros(n_ - idx.numel(), dtype=torch.long)])
        try:
            agent._update(t, 1000)
        finally:
            torch.randperm = orig
        assert torch.equal(memory.sampling_indexes.cpu(), idx)          # index path: bit-exact
        st = agent._stats.cpu().numpy()
        for j, name in enumerate(STAT_NAMES):
            ref = d[f"r{k}_stats"][j]
            assert abs(st[j] - ref) <= 2e-5 * max(1.0, abs(ref)), (name, st[j], ref)
        for net in NETS:
            flat = getattr(agent, net).flat
            sub = flat[:: meta["stride"]].cpu().numpy()
            assert rel(d[f"r{k}_{net}_sub"], sub) <= 1e-5, f"{net} round {k}"
            assert flat.double().sum().item() == pytest.approx(float(d[f"r{k}_{net}_sum"]), abs=2e-3)
        sc = agent._state_preprocessor
        assert rel(d[f"r{k}_scaler_mean"], sc.running_mean.cpu().numpy()) <= 1e-6
        assert rel(d[f"r{k}_scaler_var"], sc.running_variance.cpu().numpy()) <= 1e-6
        assert float(sc.current_count) == float(d[f"r{k}_scaler_count"])
        assert rel(d[f"r{k}_policy_m_sub"], agent.policy_optimizer.m[:: meta["stride"]].cpu().numpy()) <= 1e-5
        assert rel(d[f"r{k}_policy_v_sub"], agent.policy_optimizer.v[:: meta["stride"]].cpu().numpy()) <= 1e-5


@pytest.mark.parametrize("n", [1, 63, 4096, 16384])
def test_act_fp32_vs_oracle_sizes(n):
    import torch

    _, meta = load("agent_kamma_4d.npz")
    agent, _ = build_agent(meta)
    o = make_oracle(meta)
    g = torch.Generator().manual_seed(n)
    for step in range(3):
        states = torch.randn(n, 6, generator=g) * 2
        z = torch.randn(n, 4, generator=g)
        agent.supply_draws(normals=z.numpy())
        a = agent.act(states.cuda(), 200 + step, 1000)[0]
        ao = o.act(states, 200 + step, 1000, normals=z)
        assert rel(ao.numpy(), a.cpu().numpy()) <= 1e-5


def test_scaler_forward_train_matches_skrl_restatement():
    import torch

    from options_in_simulation_b200 import agent as A
    from oracle.ref_harness.skrl_shim import RunningStandardScaler as Ref

    mine, ref = A.RunningStandardScaler(6, device="cuda:0"), Ref(6)
    g = torch.Generator().manual_seed(0)
    for k in range(4):
        x = torch.randn(512, 6, generator=g) * (k + 1) + k
        y, yr = mine(x.cuda(), train=True), ref(x, train=True)
        assert rel(yr.numpy(), y.cpu().numpy()) <= 1e-6
        assert rel(ref.running_mean.numpy(), mine.running_mean.cpu().numpy()) <= 1e-6
        assert rel(ref.running_variance.numpy(), mine.running_variance.cpu().numpy()) <= 1e-6
        assert float(mine.current_count) == float(ref.current_count)
    x = torch.randn(1000, 6, generator=g) * 50
    mine.load_state_dict(ref.state_dict())                 # identical fp64 state ->
    assert torch.equal(mine(x.cuda()).cpu(), ref(x))      # inference form: same fp32 expression, bit-exact


def test_checkpoint_roundtrip_and_reference_format(tmp_path):
    """agent_{t}.pt is a dict of state_dicts with the reference's keys (ddpg.py:116-119,184-190): it loads
    into torch modules shaped like the reference's models, and back."""
    import torch

    from oracle.agent_oracle import flat_params, make_net

    _, meta = load("agent_kamma_4d.npz")
    agent, _ = build_agent(meta)
    path = str(tmp_path / "agent_100.pt")
    agent.save(path)
    ck = torch.load(path, map_location="cpu", weights_only=False)
    assert set(ck) == {"policy", "target_policy", "critic", "target_critic", "policy_optimizer", "critic_optimizer",
                       "state_preprocessor"}

    class RefModel(torch.nn.Module):    # attribute name `net` as in the train scripts
        def __init__(self, i, o, act):
            super().__init__()
            self.net = make_net(i, o, act)

    ref = RefModel(6, 4, "sigmoid")
    ref.load_state_dict(ck["policy"])
    assert torch.equal(flat_params(ref), agent.policy.flat.cpu())
    opt = torch.optim.Adam(ref.parameters(), lr=3e-4)
    opt.load_state_dict(ck["policy_optimizer"])
    assert ck["state_preprocessor"]["running_mean"].dtype == torch.float64
    # and the other way: a checkpoint written from torch modules loads into the B200 agent
    torch.manual_seed(99)
    ref2 = RefModel(6, 4, "sigmoid")
    ck["policy"] = ref2.state_dict()
    torch.save(ck, path)
    agent.load(path)
    assert torch.equal(agent.policy.flat.cpu(), flat_params(ref2))


def test_bf16_actor_inference_within_stated_tolerance():
    """Both tcgen05 actor kernels against the fp32 oracle: the transposed 64-env-tile kernel (csrc/mlp_tct.cu, the
    default while its 64-env tiles fit one wave of the SMs) and the 128-env-row-tile kernel (csrc/mlp_tc.cu), selected through the row threshold."""
    import torch

    import ctypes

    import options_in_simulation_b200 as ois

    lib = ois.lib()
    lib.ois_debug_set_tct_max_rows.argtypes = [ctypes.c_longlong]
    lib.ois_debug_tct_launches.restype = ctypes.c_longlong
    _, meta = load("agent_kamma_4d.npz")
    outs = {}
    try:
        for max_rows in (1 << 30, 0):
            lib.ois_debug_set_tct_max_rows(max_rows)
            # batch_size 512: the transposed kernel reads the update's weight image, which exists for batches that are a
            # multiple of 128 (csrc/update_tc.cu); the launch counter below proves which kernel served each call
            agent, _ = build_agent(meta, mixed_precision=True, batch_size=512)
            o = make_oracle(meta)
            g = torch.Generator().manual_seed(5)
            # 32-env tiles up to 32 * 148 rows (one wave), 64-env tiles beyond; several tiles per CTA from 64 * 148 rows on
            for n in (4096, 1, 31, 32, 33, 63, 64, 65, 100, 32 * 148, 32 * 148 + 1, 64 * 148 + 77, 128 * 148 + 77, 32768):
                states = torch.randn(n, 6, generator=g) * 2
                z = torch.randn(n, 4, generator=g)
                agent.supply_draws(normals=z.numpy())
                before = lib.ois_debug_tct_launches()
                a = agent.act(states.cuda(), 300, 1000)[0].cpu().numpy()
                assert lib.ois_debug_tct_launches() - before == (1 if max_rows else 0), (max_rows, n)
                ao = o.act(states, 300, 1000, normals=z).numpy()
                err = np.abs(a - ao)
                assert err.max() <= BF16_ATOL, (max_rows, n, err.max())
                assert err.mean() <= 3e-3
                outs[(max_rows, n)] = a
        for n in (4096, 100, 64 * 148 + 77):     # the two kernels agree with each other at least as closely
            assert np.abs(outs[(1 << 30, n)] - outs[(0, n)]).max() <= BF16_ATOL
    finally:
        lib.ois_debug_set_tct_max_rows(-1)
