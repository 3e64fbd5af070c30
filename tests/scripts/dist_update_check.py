"""Run under torchrun on >= 2 GPUs (tests/test_multi_gpu.py spawns it; bench.py --gpus N runs the same assertions on
its own update leg): skrl data-parallel semantics of DDPG._update (src/my_agents/ddpg.py:122-127,409-410,431-432) —
parameters broadcast from rank 0, rank-local replay shards, flat critic / actor gradients summed over the ranks and
divided by the world size -> identical parameters on every rank after every update, different from what a single rank
would have computed alone.

Every combination the library offers is exercised, the one bench.py times first:
    precision bf16 (tcgen05 update) / fp32, batch 512 / 128, sampler device / host, one CUDA graph per update / eager,
    exchange "peer" (inside the Adam kernels over NVLink peer memory, csrc/comm.cuh) / "nccl" (dist.all_reduce).
"peer" and "nccl" must agree with each other (bit for bit at world 2, where a + b has one summation order)."""
import os
import sys

import numpy as np
import torch
import torch.distributed as dist

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
import options_in_simulation_b200 as ois
from options_in_simulation_b200 import agent as A


class Box:
    def __init__(self, lo, hi, n):
        self.shape, self.low, self.high = (n,), np.full(n, lo, np.float32), np.full(n, hi, np.float32)


CASES = [   # (name, mixed_precision, batch, sampler, graph_update, dp_exchange)
    ("bf16 B512 device graph peer", True, 512, "device", True, "peer"),
    ("bf16 B512 device graph nccl", True, 512, "device", True, "nccl"),
    ("bf16 B512 device eager peer", True, 512, "device", False, "peer"),
    ("fp32 B128 host eager peer", False, 128, "host", False, "peer"),
    ("fp32 B128 host eager nccl", False, 128, "host", False, "nccl"),
]


def run_case(name, mixed, batch, sampler, graph, exchange, rank, world, dev):
    osp, asp = Box(-1, 1, 6), Box(0, 1, 4)
    torch.manual_seed(100 + rank)          # different initial weights per rank: the broadcast must erase that
    models = {"policy": A.DeterministicActor(osp, asp, dev), "target_policy": A.DeterministicActor(osp, asp, dev),
              "critic": A.Critic(osp, asp, dev), "target_critic": A.Critic(osp, asp, dev)}
    n = 64
    memory = ois.RandomMemory(memory_size=16, num_envs=n, device=dev, replacement=False)
    cfg = {"batch_size": batch, "discount_factor": 0.999, "polyak": 0.005, "actor_learning_rate": 3e-4,
           "critic_learning_rate": 1e-3, "state_preprocessor": A.RunningStandardScaler,
           "state_preprocessor_kwargs": {"size": osp, "device": dev}, "mixed_precision": mixed, "sampler": sampler,
           "graph_update": graph, "dp_exchange": exchange,
           "exploration": {"noise": None}, "experiment": {"write_interval": 0, "checkpoint_interval": 0}}
    agent = A.DDPG(models=models, memory=memory, cfg=cfg, observation_space=osp, action_space=asp, device=dev)
    agent.init({"timesteps": 100})
    assert agent.is_distributed and agent.world_size == world
    assert agent._dp_exchange == exchange, f"{name}: exchange fell back to {agent._dp_exchange}"

    def gathered(flat):
        out = [torch.empty_like(flat) for _ in range(world)]
        dist.all_gather(out, flat.contiguous())
        return out

    for net in ("policy", "critic", "target_policy", "target_critic"):
        g = gathered(getattr(agent, net).flat)
        assert all(torch.equal(g[0], x) for x in g), f"{name}: {net} not identical after broadcast"
    solo = {k: getattr(agent, k).flat.clone() for k in ("policy", "critic")}

    gen = torch.Generator().manual_seed(7 + rank)      # rank-local replay shard
    for _ in range(16):
        memory.add_samples(states=torch.randn(n, 6, generator=gen).to(dev), actions=torch.rand(n, 4, generator=gen).to(dev),
                           rewards=(torch.randn(n, 1, generator=gen) * 5 - 5).to(dev),
                           next_states=torch.randn(n, 6, generator=gen).to(dev),
                           terminated=(torch.rand(n, 1, generator=gen) < 0.05).to(dev),
                           truncated=torch.zeros(n, 1, dtype=torch.bool).to(dev))
    torch.manual_seed(5 + rank)                        # rank-local sampling
    for step in range(4):                              # with graph_update: eager + capture, then replays
        agent._update(10 + step, 100)
        torch.cuda.synchronize()
        for net in ("policy", "critic", "target_policy", "target_critic"):
            g = gathered(getattr(agent, net).flat)
            assert all(torch.equal(g[0], x) for x in g), f"{name}: {net} diverged across ranks after update {step}"
    assert not torch.equal(solo["policy"], agent.policy.flat) and not torch.equal(solo["critic"], agent.critic.flat)
    # the ranks really trained on different data: their LOCAL gradients differ, the applied (summed) one does not
    gc = gathered(agent._grad_critic)
    assert not all(torch.equal(gc[0], x) for x in gc[1:]) or exchange == "nccl"   # nccl reduces in place
    # the scaler statistics stay rank-local (skrl does not reduce them)
    m = gathered(agent._state_preprocessor.running_mean.float())
    assert not all(torch.equal(m[0], x) for x in m[1:])
    if agent._exchange is not None:
        done_c, done_a, status = agent._exchange.status()
        assert status == 0 and done_c == 4 and done_a == 4, f"{name}: exchange status {(done_c, done_a, status)}"
    out = {k: getattr(agent, k).flat.clone() for k in ("policy", "critic")}
    dist.barrier()
    return out


def main():
    rank, local, world = int(os.environ["RANK"]), int(os.environ["LOCAL_RANK"]), int(os.environ["WORLD_SIZE"])
    torch.cuda.set_device(local)
    dev = f"cuda:{local}"
    dist.init_process_group("nccl", device_id=torch.device(dev))
    results = {}
    for case in CASES:
        results[case[0]] = run_case(*case, rank, world, dev)
        if rank == 0:
            print(f"  ok: {case[0]}", flush=True)
    # the two exchanges compute the same reduction (sum over ranks, then 1/world): same weights
    for a, b in (("bf16 B512 device graph peer", "bf16 B512 device graph nccl"), ("fp32 B128 host eager peer", "fp32 B128 host eager nccl"),
                 ("bf16 B512 device graph peer", "bf16 B512 device eager peer")):
        for net in ("policy", "critic"):
            wa, wb = results[a][net], results[b][net]
            if world == 2 or "eager peer" in b:
                assert torch.equal(wa, wb), f"{a} vs {b}: {net} differs"
            else:
                assert (wa - wb).abs().max().item() <= 1e-5, f"{a} vs {b}: {net} differs"
    dist.barrier()
    if rank == 0:
        print(f"dist_update_check ok: world {world}")
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
