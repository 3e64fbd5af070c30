"""GPU tests of the bf16 tcgen05 update path (csrc/update_tc.cu):
  * the grouped-GEMM kernel in its four operand-orientation modes against torch on the same bf16-rounded operands
    (fp32 accumulation on both sides: <= 2e-3 relative to the largest |C|, i.e. accumulation-order noise only);
  * DDPG._update with ``mixed_precision=True`` against the exact fp32 path from the same state on the same batches.
Stated bf16 tolerance of the update (operands, saved activations and back-propagated gradients are rounded to bf16 —
8 mantissa bits — accumulation, master weights, Adam and polyak are fp32):
    losses / Q statistics            <= 2e-2 relative to max(1, |ref|)
    parameter gradients, per network ||g_bf16 - g_fp32||_2 <= 3e-2 ||g_fp32||_2   and cosine >= 0.999
    Adam first moments after the step  the same bound as the gradients (m = m0 + 0.1 (g - m0) carries the gradient)
Structure is exact: the same tensors change, the scaler statistics (fp64, shared kernels) are identical.
``test_bf16_update_vs_oracle`` holds the bf16 path DIRECTLY to oracle/agent_oracle.py (CPU torch autograd, pinned
to the reference's ddpg.py by oracle/gen_golden_agent.py) and to the reference's golden statistics, under the same
tolerances: this is the path bench.py times."""
import ctypes as C

import numpy as np
import pytest

from agent_cases import BATCH_NAMES, load

pytestmark = pytest.mark.gpu


def to_blocked(mat, rb):
    """[R, C] float -> tile-blocked bf16 bytes: [R/rb][C/8][rb][8] (update_tc.cu header)."""
    import torch

    R, Cc = mat.shape
    assert R % rb == 0 and Cc % 8 == 0
    return mat.to(torch.bfloat16).reshape(R // rb, rb, Cc // 8, 8).permute(0, 2, 1, 3).contiguous()


def debug_gemm(a_store, a_rb, b_store, b_rb, M, N, K, tile_n, mode, transposed=False):
    import torch

    from options_in_simulation_b200 import _lib as L

    fn = L.lib().ois_debug_tc_gemm
    fn.restype = C.c_int
    fn.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_void_p, C.c_int, C.c_int] + [C.c_int] * 5 + [C.c_void_p, C.c_int, C.c_int,
                                                                                                 C.c_void_p]
    a_blk, b_blk = to_blocked(a_store, a_rb).cuda(), to_blocked(b_store, b_rb).cuda()
    out = torch.full((N, M) if transposed else (M, N), float("nan"), device="cuda")
    L.check(fn(a_blk.data_ptr(), a_store.shape[1] // 8, a_rb, b_blk.data_ptr(), b_store.shape[1] // 8, b_rb, M, N, K, tile_n,
               mode, out.data_ptr(), out.shape[1], int(transposed), L.current_stream()))
    torch.cuda.synchronize()
    return out.cpu()


@pytest.mark.parametrize("M,N,K,tile_n", [(512, 512, 512, 128), (512, 128, 512, 128), (128, 512, 128, 128), (512, 512, 16, 128),
                                          (512, 16, 128, 16), (128, 16, 512, 16), (256, 256, 384, 128),
                                          (512, 512, 512, 64), (512, 128, 512, 64), (128, 512, 128, 64), (512, 512, 16, 64),
                                          (256, 256, 384, 64), (512, 512, 512, 32)])
@pytest.mark.parametrize("mode", [0, 1, 2, 3])
def test_tc_gemm_orientations(M, N, K, tile_n, mode):
    import torch

    k_stage = min(K, 128)
    g = torch.Generator().manual_seed(M * 7 + N * 3 + K + mode)
    A = torch.randn(M, K, generator=g).to(torch.bfloat16).float()
    B = torch.randn(N, K, generator=g).to(torch.bfloat16).float()
    ref = A @ B.T
    a_store, a_rb = (A.T.contiguous(), k_stage) if mode & 1 else (A, 128)        # stored [K, M]: contraction over rows
    # K-major B: stored in 128-row blocks (as the weight images are) when the tile is narrower than 128 -> the kernel's
    # one-copy-per-column-chunk path; otherwise row block == tile
    b_store, b_rb = (B.T.contiguous(), k_stage) if mode & 2 else (B, 128 if tile_n in (32, 64) and N % 128 == 0 else tile_n)
    if a_store.shape[1] % 8 or b_store.shape[1] % 8:
        pytest.skip("stored matrix needs a multiple of 8 columns")
    out = debug_gemm(a_store, a_rb, b_store, b_rb, M, N, K, tile_n, mode)
    assert torch.isfinite(out).all()
    assert (out - ref).abs().max() <= 2e-3 * ref.abs().max()
    if mode == 3 and tile_n == 16:
        out_t = debug_gemm(a_store, a_rb, b_store, b_rb, M, N, K, tile_n, mode, transposed=True)
        assert torch.equal(out_t.T.contiguous(), out)


def _fill(memory, d, meta):
    import torch

    n = meta["n_envs"]
    for t in range(meta["rows"] // n):
        sl = slice(t * n, (t + 1) * n)
        memory.add_samples(**{k: torch.as_tensor(d[f"mem_{k}"][sl]).cuda().to(torch.bool if k in ("terminated", "truncated") else torch.float32)
                              for k in BATCH_NAMES})


@pytest.mark.parametrize("fname", ["agent_kamma_4d.npz", "agent_kaddpg_2d.npz", "agent_kaddpg_1d.npz", "agent_kamma_eval_tanh.npz"])
def test_bf16_update_tracks_fp32_update(fname):
    import torch

    from test_agent_gpu import build_agent

    d, meta = load(fname)
    if meta["batch"] % 128:
        pytest.skip("tensor-core update needs batch % 128 == 0")
    exact, mem_e = build_agent(meta)
    fast, mem_f = build_agent(meta, mixed_precision=True)
    assert torch.equal(exact.policy.flat, fast.policy.flat) and torch.equal(exact.critic.flat, fast.critic.flat)
    _fill(mem_e, d, meta)
    _fill(mem_f, d, meta)
    orig = torch.randperm
    rounds = meta["rounds"]
    for k in range(rounds):
        idx = torch.as_tensor(d[f"r{k}_indices"])
        torch.randperm = lambda n_, **kw: torch.cat([idx, torch.zeros(n_ - idx.numel(), dtype=torch.long)])
        try:
            # the comparison is per update: start the fast agent from the exact agent's state
            for name in ("policy", "target_policy", "critic", "target_critic"):
                getattr(fast, name).flat.copy_(getattr(exact, name).flat)
            for name in ("policy_optimizer", "critic_optimizer"):
                getattr(fast, name).m.copy_(getattr(exact, name).m)
                getattr(fast, name).v.copy_(getattr(exact, name).v)
                getattr(fast, name).step = getattr(exact, name).step
            fast._state_preprocessor.load_state_dict(exact._state_preprocessor.state_dict())
            fast._weights_changed()
            t = int(d[f"r{k}_timestep"])
            exact._update(t, 1000)
            fast._update(t, 1000)
        finally:
            torch.randperm = orig
        se, sf = exact._stats.cpu().numpy(), fast._stats.cpu().numpy()
        assert np.all(np.abs(se - sf) <= 2e-2 * np.maximum(1.0, np.abs(se))), (k, se, sf)
        for ge, gf, name in ((exact._grad_critic, fast._grad_critic, "critic"), (exact._grad_actor, fast._grad_actor, "actor")):
            ge, gf = ge.double(), gf.double()
            assert torch.isfinite(gf).all()
            err = ((ge - gf).norm() / ge.norm()).item()
            cos = (torch.dot(ge, gf) / (ge.norm() * gf.norm())).item()
            assert err <= 3e-2 and cos >= 0.999, (name, k, err, cos)
        # the optimizer state carries the gradient linearly (exp_avg.lerp_(grad, 0.1)), unlike the weights, whose Adam
        # step is ~lr * sign(g) whatever the gradient: compare the first moments
        for name in ("policy_optimizer", "critic_optimizer"):
            me, mf = getattr(exact, name).m.double(), getattr(fast, name).m.double()
            assert ((me - mf).norm() / me.norm()).item() <= 3e-2, (name, k)
        sc_e, sc_f = exact._state_preprocessor, fast._state_preprocessor
        assert torch.equal(sc_e.running_mean, sc_f.running_mean) and torch.equal(sc_e.running_variance, sc_f.running_variance)


@pytest.mark.parametrize("fname", ["agent_kamma_4d.npz", "agent_kaddpg_2d.npz", "agent_kaddpg_1d.npz", "agent_kamma_eval_tanh.npz"])
def test_bf16_update_vs_oracle(fname):
    """mixed_precision=True DDPG._update against oracle/agent_oracle.py on the golden batches: losses, Q / target
    statistics, the two flat gradients (vs the oracle's autograd gradients) and the scaler state; the statistics are also
    held to the values the UNMODIFIED reference produced (r{k}_stats of the golden file)."""
    import torch

    from agent_cases import STAT_NAMES, make_oracle
    from oracle.agent_oracle import adam_moments, flat_params
    from test_agent_gpu import build_agent, rel

    d, meta = load(fname)
    if meta["batch"] % 128:
        pytest.skip("tensor-core update needs batch % 128 == 0")
    fast, mem = build_agent(meta, mixed_precision=True)
    o = make_oracle(meta)
    assert torch.equal(flat_params(o.policy), fast.policy.flat.cpu()) and torch.equal(flat_params(o.critic), fast.critic.flat.cpu())
    _fill(mem, d, meta)
    rows = {k: d[f"mem_{k}"] for k in BATCH_NAMES}
    orig = torch.randperm
    for k in range(meta["rounds"]):
        idx_np = d[f"r{k}_indices"]
        idx = torch.as_tensor(idx_np)
        # per-update comparison: the fast agent starts every round from the oracle's state
        for name in ("policy", "target_policy", "critic", "target_critic"):
            getattr(fast, name).flat.copy_(flat_params(getattr(o, name)).cuda())
        for opt_name, net in (("policy_optimizer", "policy"), ("critic_optimizer", "critic")):
            m, v, step = adam_moments(getattr(o, opt_name), getattr(o, net))
            opt = getattr(fast, opt_name)
            opt.m.copy_(m.cuda()); opt.v.copy_(v.cuda()); opt.step = step
        sc = fast._state_preprocessor
        sc.running_mean.copy_(o.scaler.running_mean); sc.running_variance.copy_(o.scaler.running_variance)
        sc.current_count.copy_(o.scaler.current_count)
        fast._weights_changed()
        torch.randperm = lambda n_, **kw: torch.cat([idx, torch.zeros(n_ - idx.numel(), dtype=torch.long)])
        try:
            fast._update(int(d[f"r{k}_timestep"]), 1000)
        finally:
            torch.randperm = orig
        st_o = o.update(*[rows[n][idx_np] for n in BATCH_NAMES])
        got = fast._stats.cpu().numpy()
        for j, name in enumerate(STAT_NAMES):
            assert abs(got[j] - st_o[name]) <= 2e-2 * max(1.0, abs(st_o[name])), (k, name, got[j], st_o[name])
            ref = float(d[f"r{k}_stats"][j])       # the unmodified reference's value for this round
            assert abs(got[j] - ref) <= 2e-2 * max(1.0, abs(ref)), (k, name, got[j], ref)
        for g_ref, g_got, name in ((o.last_critic_grad, fast._grad_critic, "critic"), (o.last_actor_grad, fast._grad_actor, "actor")):
            ge, gf = g_ref.double(), g_got.cpu().double()
            assert torch.isfinite(gf).all()
            err = ((ge - gf).norm() / ge.norm()).item()
            cos = (torch.dot(ge, gf) / (ge.norm() * gf.norm())).item()
            assert err <= 3e-2 and cos >= 0.999, (name, k, err, cos)
        for opt_name, net in (("policy_optimizer", "policy"), ("critic_optimizer", "critic")):
            m, _, step = adam_moments(getattr(o, opt_name), getattr(o, net))
            mf = getattr(fast, opt_name).m.cpu().double()
            assert ((m.double() - mf).norm() / m.double().norm()).item() <= 3e-2, (opt_name, k)
            assert getattr(fast, opt_name).step == step
        assert rel(o.scaler.running_mean.numpy(), sc.running_mean.cpu().numpy()) <= 1e-6
        assert rel(o.scaler.running_variance.numpy(), sc.running_variance.cpu().numpy()) <= 1e-6
        assert float(sc.current_count) == float(o.scaler.current_count)


def test_bf16_update_is_deterministic_and_learns():
    """Two identical bf16 runs give bit-identical weights (no atomics in the gradient path), and the critic loss on a
    fixed batch falls over 30 updates."""
    import torch

    from test_agent_gpu import build_agent

    d, meta = load("agent_kamma_4d.npz")
    finals, losses = [], []
    for rep in range(2):
        agent, mem = build_agent(meta, mixed_precision=True)
        _fill(mem, d, meta)
        idx = torch.as_tensor(d["r0_indices"])
        orig = torch.randperm
        torch.randperm = lambda n_, **kw: torch.cat([idx, torch.zeros(n_ - idx.numel(), dtype=torch.long)])
        try:
            ls = []
            for i in range(30):
                agent._update(100 + i, 1000)
                ls.append(float(agent._stats[0]))
        finally:
            torch.randperm = orig
        finals.append((agent.policy.flat.clone(), agent.critic.flat.clone()))
        losses.append(ls)
    assert torch.equal(finals[0][0], finals[1][0]) and torch.equal(finals[0][1], finals[1][1])
    assert losses[0] == losses[1]
    assert np.mean(losses[0][-5:]) < 0.7 * np.mean(losses[0][:5]), losses[0]


@pytest.mark.parametrize("mixed_precision", [True, False])
def test_graph_update_equals_eager_update(mixed_precision):
    """cfg graph_update=True replays the gather + three phases as one CUDA graph: same kernels, same order, same
    device-side sampler draws -> bit-identical weights, optimizer state and scaler statistics."""
    import torch

    from test_agent_gpu import build_agent

    d, meta = load("agent_kamma_4d.npz")
    agents = []
    for graph in (False, True):
        agent, mem = build_agent(meta, mixed_precision=mixed_precision, sampler="device", graph_update=graph)
        _fill(mem, d, meta)
        for i in range(6):
            agent._update(100 + i, 1000)
            if i == 2:      # inference between updates must see the updated weights (image rebuilt after a replay)
                states = torch.as_tensor(d["r0_states"]).cuda()
                agent.supply_draws(normals=np.zeros((states.shape[0], meta["act_dim"]), np.float32))
                agent._probe = agent.act(states, 500, 1000)[0].clone()
        torch.cuda.synchronize()
        agents.append(agent)
    e, g = agents
    assert g._graph is not None and e._graph is None
    for name in ("policy", "target_policy", "critic", "target_critic"):
        assert torch.equal(getattr(e, name).flat, getattr(g, name).flat), name
    assert torch.equal(e.policy_optimizer.m, g.policy_optimizer.m) and torch.equal(e.critic_optimizer.v, g.critic_optimizer.v)
    assert e.policy_optimizer.step == g.policy_optimizer.step == 6
    assert torch.equal(e._state_preprocessor.running_mean, g._state_preprocessor.running_mean)
    assert float(e._state_preprocessor.current_count) == float(g._state_preprocessor.current_count)
    assert torch.equal(e._probe, g._probe)
    assert torch.equal(e._stats, g._stats)


def test_incremental_weight_images_equal_full_repack():
    """Adam / polyak keep the bf16 tensor-core weight images current element by element; forcing a full repack from the
    fp32 masters before every update must give bit-identical results."""
    import torch

    from test_agent_gpu import build_agent

    d, meta = load("agent_kamma_4d.npz")
    out = []
    for repack in (False, True):
        agent, mem = build_agent(meta, mixed_precision=True, sampler="device")
        _fill(mem, d, meta)
        for i in range(5):
            if repack:
                agent._weights_changed()
            agent._update(100 + i, 1000)
        out.append(agent)
    for name in ("policy", "target_policy", "critic", "target_critic"):
        assert torch.equal(getattr(out[0], name).flat, getattr(out[1], name).flat), name


def test_bf16_agent_with_other_batch_sizes_still_acts_and_update_says_why():
    """The tensor-core update needs batch_size % 128 == 0; inference must not depend on it, and _update must fail
    loudly (no silent fp32 fallback) when it cannot run."""
    import torch

    from options_in_simulation_b200 import _lib as L
    from test_agent_gpu import build_agent

    d, meta = load("agent_kamma_4d.npz")
    meta = dict(meta, batch=64)
    agent, mem = build_agent(meta, mixed_precision=True)
    _fill(mem, d, meta)
    a = agent.act(torch.zeros(300, meta["obs_dim"], device="cuda"), 500, 1000)[0]
    assert a.shape == (300, meta["act_dim"]) and torch.isfinite(a).all()
    with pytest.raises(L.OisError, match="multiple of 128"):
        agent._update(100, 1000)
