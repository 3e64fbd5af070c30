"""World-size-2 ``gloo`` tests (CPU) of the data-parallel host logic in options_in_simulation_b200/parallel.py:
parameter broadcast, flat-gradient all-reduce (skrl Model.reduce_parameters: SUM then / world_size,
src/my_agents/ddpg.py:409-410,431-432), env sharding, per-rank seeds, eval-sweep bin reduction, max-over-ranks
timing.  The GPU box runs the same functions on ``nccl``."""
import os
import socket

import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from options_in_simulation_b200 import parallel as P


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _worker(rank, world, port, out):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), LOCAL_RANK=str(rank),
                      WORLD_SIZE=str(world))
    info = P.distributed_info()
    assert info.is_distributed and info.rank == rank and info.world_size == world
    P.ensure_process_group("gloo")
    assert dist.is_initialized()
    res = {}
    # 1. broadcast of the flat parameter buffer: every rank ends with rank 0's weights
    g = torch.Generator().manual_seed(100 + rank)
    flat = torch.randn(1000, generator=g)
    P.broadcast_flat(flat, 0)
    res["params"] = flat.clone()
    # 2. gradient all-reduce: SUM in place, consumer applies the returned 1/world factor
    grad = torch.arange(8, dtype=torch.float32) * (rank + 1)
    scale = P.all_reduce_gradients(grad)
    res["grad_mean"] = grad * scale
    # 3. a DP "update": identical params + averaged grads -> identical params on every rank
    res["stepped"] = flat - 0.1 * (torch.full_like(flat, float(rank)) * 1.0)   # local grads differ ...
    local = torch.full_like(flat, float(rank))
    s = P.all_reduce_gradients(local)
    res["stepped"] = flat - 0.1 * local * s                                       # ... the averaged ones do not
    # 4. shards, seeds
    res["shard"] = P.shard_range(65537, rank, world)
    res["seed"] = P.seed_for_rank(42)
    # 5. eval sweep bins
    bins = P.TrackingErrorBins(1.0, 2.0, 10)
    bins.add(torch.tensor([1.05, 1.95]) if rank == 0 else torch.tensor([1.05, 1.55]),
             torch.tensor([0.2, 0.4]) if rank == 0 else torch.tensor([0.4, 0.6]))
    bins.reduce()
    res["bins"] = (bins.te_sum.clone(), bins.count.clone(), bins.means().clone())
    # 6. timing: max over ranks
    res["tmax"] = P.max_over_ranks(1.0 + rank, "cpu")
    # 7. the one host-side exchange of the peer-memory gradient path: 64-byte IPC handles, concatenated in rank order
    res["handles"] = P.gather_handles(bytes([rank + 1]) * 64, world)
    out[rank] = res
    dist.barrier()
    dist.destroy_process_group()


@pytest.mark.timeout(120)
def test_world_size_2_gloo():
    world = 2
    mgr = mp.Manager()
    out = mgr.dict()
    mp.spawn(_worker, args=(world, _free_port(), out), nprocs=world, join=True)
    r0, r1 = out[0], out[1]
    g = torch.Generator().manual_seed(100)
    assert torch.equal(r0["params"], torch.randn(1000, generator=g)) and torch.equal(r0["params"], r1["params"])
    want = torch.arange(8, dtype=torch.float32) * 1.5          # (1x + 2x) / 2
    assert torch.equal(r0["grad_mean"], want) and torch.equal(r1["grad_mean"], want)
    assert torch.equal(r0["stepped"], r1["stepped"])
    assert r0["shard"] == (0, 32769) and r1["shard"] == (32769, 65537)
    assert (r0["seed"], r1["seed"]) == (42, 43)
    for r in (r0, r1):
        te, cnt, mean = r["bins"]
        assert cnt.tolist() == [2, 0, 0, 0, 0, 1, 0, 0, 0, 1]
        assert te[0].item() == pytest.approx(0.6) and mean[0].item() == pytest.approx(0.3)
        assert mean[5].item() == pytest.approx(0.6) and mean[9].item() == pytest.approx(0.4)
        assert r["tmax"] == 2.0
        assert r["handles"] == b"\x01" * 64 + b"\x02" * 64


def test_single_process_is_identity():
    for k in ("WORLD_SIZE", "RANK", "LOCAL_RANK"):
        os.environ.pop(k, None)
    assert not P.distributed_info().is_distributed
    t = torch.ones(4)
    assert P.all_reduce_gradients(t) == 1.0 and torch.equal(t, torch.ones(4))
    P.broadcast_flat(t)
    assert P.shard_range(10, 0, 1) == (0, 10)
    covered = [P.shard_range(4099, r, 8) for r in range(8)]
    assert covered[0][0] == 0 and covered[-1][1] == 4099
    assert all(covered[i][1] == covered[i + 1][0] for i in range(7))
