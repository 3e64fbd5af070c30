"""GPU: replay memory kernels (csrc/replay.cu) — ring append, gather by index (bit-exact), device index
sampling, and the append fused into the env kernel vs the separate add_samples path.
Reference semantics: skrl Memory as restated by the inline FIFOMemory
(train/train_kaddpg_1d_nofifo_random_RLIL.py:105-206) and oracle/ref_harness/skrl_shim.py."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu
NAMES = ("states", "actions", "rewards", "next_states", "terminated", "truncated")


def _mem(cls, size, n, act_dim=4, **kw):
    import torch

    import options_in_simulation_b200 as ois

    m = getattr(ois, cls)(memory_size=size, num_envs=n, device="cuda:0", **kw)
    m.create_tensor("states", 6, torch.float32)
    m.create_tensor("next_states", 6, torch.float32)
    m.create_tensor("actions", act_dim, torch.float32)
    m.create_tensor("rewards", 1, torch.float32)
    m.create_tensor("terminated", 1, torch.bool)
    m.create_tensor("truncated", 1, torch.bool)
    return m


def _shim_mem(size, n, act_dim=4):
    import torch

    from oracle.ref_harness.skrl_shim import RandomMemory

    m = RandomMemory(memory_size=size, num_envs=n, device="cpu", replacement=False)
    m.create_tensor("states", 6, torch.float32)
    m.create_tensor("next_states", 6, torch.float32)
    m.create_tensor("actions", act_dim, torch.float32)
    m.create_tensor("rewards", 1, torch.float32)
    m.create_tensor("terminated", 1, torch.bool)
    m.create_tensor("truncated", 1, torch.bool)
    return m


@pytest.mark.parametrize("n,act_dim,size,steps", [(64, 1, 5, 12), (4096, 4, 3, 7), (1, 2, 4, 9), (333, 4, 2, 2)])
def test_append_wraps_like_skrl_and_gather_is_bit_exact(n, act_dim, size, steps):
    import torch

    g = torch.Generator().manual_seed(n + steps)
    mine, ref = _mem("FIFOMemory", size, n, act_dim), _shim_mem(size, n, act_dim)
    assert len(mine) == 0
    for t in range(steps):
        tr = dict(states=torch.randn(n, 6, generator=g), actions=torch.rand(n, act_dim, generator=g),
                  rewards=torch.randn(n, 1, generator=g), next_states=torch.randn(n, 6, generator=g),
                  terminated=torch.rand(n, 1, generator=g) < 0.1, truncated=torch.zeros(n, 1, dtype=torch.bool))
        ref.add_samples(**tr)
        mine.add_samples(**{k: v.cuda() for k, v in tr.items()})
        assert (mine.memory_index, mine.filled, len(mine)) == (ref.memory_index, ref.filled, len(ref))
    torch.cuda.synchronize()
    for k in NAMES:   # NaN-initialised slots compare equal bitwise
        a, b = mine.tensors[k].cpu(), ref.tensors[k]
        assert torch.equal(a.view(torch.uint8) if a.dtype == torch.bool else a.view(torch.int32),
                           b.view(torch.uint8) if b.dtype == torch.bool else b.view(torch.int32)), k
    batch = min(512, len(ref))
    idx = torch.randperm(len(ref), generator=g)[:batch]
    out = mine.sample_by_index(NAMES, idx)[0]
    exp = ref.sample_by_index(NAMES, idx)[0]
    for k, a, b in zip(NAMES, out, exp):
        assert a.shape == b.shape and a.dtype == b.dtype
        assert torch.equal(a.cpu(), b), k


def test_sample_host_randperm_reproduces_reference_indices():
    """sampler='host' consumes the torch CPU generator exactly like the reference's randperm(len)[:B]."""
    import torch

    n, size = 256, 8
    mine, ref = _mem("RandomMemory", size, n), _shim_mem(size, n)
    g = torch.Generator().manual_seed(0)
    for _ in range(5):
        tr = dict(states=torch.randn(n, 6, generator=g), actions=torch.rand(n, 4, generator=g),
                  rewards=torch.randn(n, 1, generator=g), next_states=torch.randn(n, 6, generator=g),
                  terminated=torch.zeros(n, 1, dtype=torch.bool), truncated=torch.zeros(n, 1, dtype=torch.bool))
        ref.add_samples(**tr)
        mine.add_samples(**{k: v.cuda() for k, v in tr.items()})
    torch.manual_seed(123)
    exp = ref.sample(NAMES, 512)[0]
    torch.manual_seed(123)
    out = mine.sample(NAMES, 512)[0]
    assert torch.equal(mine.sampling_indexes.cpu(), ref.sampling_indexes)
    for a, b in zip(out, exp):
        assert torch.equal(a.cpu(), b)


@pytest.mark.parametrize("length", [512, 513, 4096 * 245, 1_000_000, 10_000_019])
def test_device_sampler_distinct_in_range_uniform(length):
    import torch

    import options_in_simulation_b200 as ois
    from options_in_simulation_b200 import _lib as L

    batch = 512
    idx = torch.empty(batch, dtype=torch.int64, device="cuda")
    hist = np.zeros(8)
    for draw in range(64):
        L.check(ois.lib().ois_replay_sample_indices(idx.data_ptr(), batch, length, 77, draw, None))
        torch.cuda.synchronize()
        v = idx.cpu().numpy()
        assert v.min() >= 0 and v.max() < length
        assert len(np.unique(v)) == batch            # without replacement, like randperm(len)[:B]
        hist += np.histogram(v, bins=8, range=(0, length))[0]
    exp = hist.sum() / 8
    assert (np.abs(hist - exp) < 6 * np.sqrt(exp)).all()   # uniform over the memory
    with pytest.raises(ois.OisError):
        L.check(ois.lib().ois_replay_sample_indices(idx.data_ptr(), 512, 100, 0, 0, None))


def test_fused_append_equals_record_transition_path():
    """env kernel epilogue (row `memory_index` written in-kernel, KA relabel actions = wheel_torqs/scale,
    base_KA_DDPG.py:212) == wrapper.step + memory.add_samples done separately."""
    import torch

    import options_in_simulation_b200 as ois
    from options_in_simulation_b200.config import script_cfg

    n, size = 1024, 4
    tasks = [ois.TorqueDistributionTask(sim_config=script_cfg(n), variant="KAMMA_medium", seed=5) for _ in range(2)]
    fused, plain = _mem("RandomMemory", size, n), _mem("RandomMemory", size, n)
    tasks[0].bind_replay(fused)
    envs = [ois.wrap_env(t) for t in tasks]
    obs = [e.reset()[0].clone() for e in envs]
    assert torch.equal(obs[0], obs[1])
    g = torch.Generator(device="cuda").manual_seed(1)
    for t in range(6):
        a = torch.rand(n, 4, device="cuda", generator=g)
        o0, r0, d0, tr0, _ = envs[0].step(a)
        o1, r1, d1, tr1, _ = envs[1].step(a)
        # true fp32 division like CPU torch (CUDA torch multiplies by the reciprocal: 1 ulp apart)
        relabeled = (tasks[1].wheel_torqs.cpu() / tasks[1].action_scale).cuda()
        plain.add_samples(states=obs[1], actions=relabeled, rewards=r1, next_states=o1, terminated=d1, truncated=tr1)
        obs = [o0.clone(), o1.clone()]
        assert fused.memory_index == plain.memory_index and fused.filled == plain.filled
    torch.cuda.synchronize()
    for k in NAMES:
        a, b = fused.tensors[k], plain.tensors[k]
        if a.dtype == torch.bool:
            assert torch.equal(a, b), k
        else:
            assert torch.equal(a.view(torch.int32), b.view(torch.int32)), k


def test_sample_gather_matches_the_two_launch_path_and_advances_its_control_block():
    """ois_replay_sample_gather (what a graph-captured update replays) draws the indices ois_replay_sample_indices draws
    for the same (len, seed, offset), gathers the same rows, and moves the device-resident OisSampleCtl on."""
    import torch

    import options_in_simulation_b200 as ois
    from options_in_simulation_b200 import _lib as L

    n, size = 96, 40
    m = ois.RandomMemory(memory_size=size, num_envs=n, device="cuda:0", replacement=False, sampler="device", seed=5)
    names = ("states", "actions", "rewards", "next_states", "terminated", "truncated")
    for name, d, dt in (("states", 6, torch.float32), ("next_states", 6, torch.float32), ("actions", 4, torch.float32),
                        ("rewards", 1, torch.float32), ("terminated", 1, torch.bool), ("truncated", 1, torch.bool)):
        m.create_tensor(name, d, dt)
    g = torch.Generator(device="cuda").manual_seed(1)
    for name in ("states", "next_states", "actions", "rewards"):
        m.tensors[name].copy_(torch.randn(m.tensors[name].shape, device="cuda", generator=g))
    m.tensors["terminated"].copy_(torch.rand(size, n, 1, device="cuda", generator=g) < 0.1)
    m.memory_index, m.filled = 17, False          # 17 * 96 transitions to sample from
    length, batch = len(m), 512
    ctl = torch.zeros(4, dtype=torch.int64, device="cuda")
    for draw0, step0 in ((0, 0), (41, 1000)):
        L.check(ois.lib().ois_sample_ctl_set(ctl.data_ptr(), length, draw0, step0, None))
        m._draws = draw0
        for k in range(3):
            idx = torch.empty(batch, dtype=torch.int64, device="cuda")
            got = [t.clone() for t in m.sample_gather(names, idx, ctl)[0]]
            ref_idx = torch.empty(batch, dtype=torch.int64, device="cuda")
            L.check(ois.lib().ois_replay_sample_indices(ref_idx.data_ptr(), batch, length, m._seed, draw0 + k, None))
            assert torch.equal(idx, ref_idx) and idx.unique().numel() == batch and int(idx.max()) < length
            want = m.sample_by_index(names, ref_idx)[0]
            for a, b in zip(got, want):
                assert torch.equal(a, b)
            torch.cuda.synchronize()
            assert ctl.tolist()[:3] == [length, draw0 + k + 1, step0 + k + 1] and ctl.tolist()[3] == 0
        assert m._draws == draw0 + 3
