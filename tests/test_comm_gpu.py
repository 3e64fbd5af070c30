"""GPU tests (ONE GPU is enough) of the peer-memory gradient exchange inside the Adam kernels (csrc/comm.cuh,
include/ois_b200.h ``ois_comm_*``; reference: the two reduce_parameters calls of src/my_agents/ddpg.py:409-410,431-432 =
all-reduce(SUM) then / world_size).

A single-GPU box cannot run ranks that wait for one another as separate launches, so the world is emulated inside one
process: ``world`` exchange blocks on the same device, connected with ``ois_comm_connect_local``, the peers' "published"
flags written by the host, and the peers' gradient buffers filled by the test between the update's phases.  What is
checked is the arithmetic and the plumbing of the kernel a real rank runs: the rank-ordered fp32 sum, the 1/world
scale, epochs advancing once per exchange, and bit-identity with a plain update whose gradient buffer holds the same sum.
The true multi-process run over CUDA IPC and NVLink is tests/scripts/dist_update_check.py (needs 2 GPUs)."""
import ctypes as C

import numpy as np
import pytest

from agent_cases import BATCH_NAMES, load

pytestmark = pytest.mark.gpu


class LocalWorld:
    """`world` OisComm blocks in this process; rank 0 is the one the agent under test uses."""

    def __init__(self, world, n_critic, n_actor):
        import torch

        from options_in_simulation_b200 import _lib as L
        from options_in_simulation_b200.parallel import _DeviceMemory

        self.L, self.world = L, world
        lib = L.lib()
        self.comms = []
        for r in range(world):
            h = C.c_void_p()
            L.check(lib.ois_comm_create(C.byref(h), r, world, n_critic, n_actor))
            self.comms.append(h)
        arr = (C.c_void_p * world)(*[h.value for h in self.comms])
        for h in self.comms:
            L.check(lib.ois_comm_connect_local(h, arr, world))
        self.grads = [[torch.as_tensor(_DeviceMemory(lib.ois_comm_grad(h, k), n), device="cuda:0")
                       for k, n in ((0, n_critic), (1, n_actor))] for h in self.comms]
        for k in (0, 1):      # every peer "has published" the next 1000 exchanges
            L.check(lib.ois_comm_debug_publish(self.comms[0], k, 1000))

    def status(self, r=0):
        w = (C.c_uint32 * 3)()
        self.L.check(self.L.lib().ois_comm_status(self.comms[r], w))
        return tuple(int(x) for x in w)

    def close(self):
        import torch

        torch.cuda.synchronize()
        self.grads = None
        for h in self.comms:
            self.L.lib().ois_comm_destroy(h)


def _fill(memory, d, meta):
    import torch

    n = meta["n_envs"]
    for t in range(meta["rows"] // n):
        sl = slice(t * n, (t + 1) * n)
        memory.add_samples(**{k: torch.as_tensor(d[f"mem_{k}"][sl]).cuda().to(torch.bool if k in ("terminated", "truncated") else torch.float32)
                              for k in BATCH_NAMES})


@pytest.mark.parametrize("world", [2, 3, 8])
@pytest.mark.parametrize("mixed", [False, True])
def test_exchanging_adam_equals_plain_adam_on_the_rank_ordered_sum(world, mixed):
    import torch

    from options_in_simulation_b200 import _lib as L
    from test_agent_gpu import build_agent

    d, meta = load("agent_kamma_4d.npz")
    plain, mem_p = build_agent(meta, mixed_precision=mixed)
    dp, mem_d = build_agent(meta, mixed_precision=mixed)
    _fill(mem_p, d, meta)
    _fill(mem_d, d, meta)
    lw = LocalWorld(world, dp.critic.num_params, dp.policy.num_params)
    dp._grad_critic, dp._grad_actor = lw.grads[0]
    lib = L.lib()
    gen = torch.Generator().manual_seed(world)
    scale = 1.0 / world
    try:
        for k in range(meta["rounds"]):
            idx = torch.as_tensor(d[f"r{k}_indices"])
            ios = []
            for ag, mem in ((plain, mem_p), (dp, mem_d)):
                ag._ensure_handle(ag._batch_size)
                batch = mem.sample_by_index(ag._tensors_names, idx)[0]
                io = ag._update_io(batch)
                ag.policy_optimizer.step += 1
                ag.critic_optimizer.step += 1
                io.adam_step = ag.critic_optimizer.step
                ios.append(io)
            io_p, io_d = ios
            io_d.comm = lw.comms[0].value
            st = L.current_stream()
            for ph, (gp, kind) in enumerate(((plain._grad_critic, 0), (plain._grad_actor, 1))):
                L.check(lib.ois_ddpg_update_phase(plain._handle, plain._scaler_ref(), C.byref(io_p), ph, scale, st))
                L.check(lib.ois_ddpg_update_phase(dp._handle, dp._scaler_ref(), C.byref(io_d), ph, scale, st))
                assert torch.equal(gp, lw.grads[0][kind])          # same state, same batch, same kernels
                # the peers' gradients of this exchange; the plain agent gets the rank-ordered fp32 sum
                total = gp.clone()
                for r in range(1, world):
                    other = gp * float(torch.rand((), generator=gen)) + 1e-3 * torch.randn(gp.shape, generator=gen).cuda()
                    lw.grads[r][kind].copy_(other)
                    total = total + other                            # ((g0 + g1) + g2) + ... in fp32, as the kernel adds
                gp.copy_(total)
            L.check(lib.ois_ddpg_update_phase(plain._handle, plain._scaler_ref(), C.byref(io_p), 2, scale, st))
            L.check(lib.ois_ddpg_update_phase(dp._handle, dp._scaler_ref(), C.byref(io_d), 2, scale, st))
            torch.cuda.synchronize()
            for name in ("policy", "critic", "target_policy", "target_critic"):
                assert torch.equal(getattr(plain, name).flat, getattr(dp, name).flat), (name, k)
            for name in ("policy_optimizer", "critic_optimizer"):
                assert torch.equal(getattr(plain, name).m, getattr(dp, name).m) and torch.equal(getattr(plain, name).v, getattr(dp, name).v)
            assert lw.status() == (k + 1, k + 1, 0)                # one epoch per exchange, no barrier timeout
            # the next act() must see the exchanged weights on the bf16 path too (images kept current by the kernel)
            x = torch.randn(256, meta["obs_dim"], generator=gen).cuda()
            plain.supply_draws(normals=np.zeros((256, meta["act_dim"]), np.float32))
            dp.supply_draws(normals=np.zeros((256, meta["act_dim"]), np.float32))
            assert torch.equal(plain.act(x, 500, 1000)[0], dp.act(x, 500, 1000)[0])
    finally:
        lw.close()


def test_comm_argument_checks():
    import torch

    from options_in_simulation_b200 import _lib as L

    lib = L.lib()
    h = C.c_void_p()
    assert lib.ois_comm_create(C.byref(h), 3, 2, 10, 10) == -1       # rank outside the world
    assert lib.ois_comm_create(C.byref(h), 0, 9, 10, 10) == -1       # more than OIS_MAX_RANKS
    L.check(lib.ois_comm_create(C.byref(h), 0, 2, 1000, 900))
    try:
        g0, g1 = lib.ois_comm_grad(h, 0), lib.ois_comm_grad(h, 1)
        assert g0 % 256 == 0 and g1 % 256 == 0 and g1 - g0 >= 4000   # 256-byte aligned, disjoint
        assert lib.ois_comm_grad(h, 2) is None
        buf = C.create_string_buffer(L.OIS_IPC_HANDLE_BYTES)
        L.check(lib.ois_comm_export(h, buf))
        assert any(buf.raw)                                          # a real IPC handle came back
        assert lib.ois_comm_connect(h, buf, 1) == -1                 # one handle per rank is required
    finally:
        torch.cuda.synchronize()
        lib.ois_comm_destroy(h)
