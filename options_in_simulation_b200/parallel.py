"""Data-parallel plumbing: one process per GPU, launched by torchrun.

skrl's DP semantics, as the reference uses them through ``Model.broadcast_parameters`` /
``Model.reduce_parameters`` (src/my_agents/ddpg.py:122-127,409-410,431-432) and its vendored ``Agent`` base
(src/my_agents/base.py:88-92): envs and replay shards are rank-local and never exchanged; parameters are
broadcast from rank 0 once; each update all-reduces (SUM) ONE flat gradient buffer per network and divides by
the world size (here the division is folded into the Adam kernel as ``grad_scale``).

Nothing here touches CUDA directly, so the functions run unchanged on the ``gloo`` backend with CPU tensors
(tests/test_distributed_cpu.py, world_size 2) and on ``nccl`` over NVLink 5 / NVSwitch on the GPU box.
"""
from __future__ import annotations

import os
from dataclasses import dataclass

import torch
import torch.distributed as dist


@dataclass(frozen=True)
class DistInfo:
    """skrl.config.torch.{is_distributed, rank, local_rank, world_size}, read from the torchrun environment."""
    is_distributed: bool
    rank: int
    local_rank: int
    world_size: int


def distributed_info() -> DistInfo:
    world = int(os.environ.get("WORLD_SIZE", "1"))
    return DistInfo(world > 1, int(os.environ.get("RANK", "0")), int(os.environ.get("LOCAL_RANK", "0")), world)


def ensure_process_group(backend: str = "nccl") -> None:
    """skrl Agent.__init__ (src/my_agents/base.py:88-92): init the default group if torchrun launched us."""
    if distributed_info().is_distributed and not dist.is_initialized():
        dist.init_process_group(backend)


def seed_for_rank(seed: int, rank: int | None = None) -> int:
    """skrl ``set_seed`` under torchrun: every rank uses ``seed + rank`` (env draws, OU noise, replay sampling)."""
    return int(seed) + (distributed_info().rank if rank is None else int(rank))


def shard_range(total_envs: int, rank: int, world_size: int) -> tuple[int, int]:
    """Contiguous env shard of a global population (eval sweep, BASELINE configs[4]): the first
    ``total % world`` ranks take one extra env."""
    base, extra = divmod(int(total_envs), int(world_size))
    lo = rank * base + min(rank, extra)
    return lo, lo + base + (1 if rank < extra else 0)


def broadcast_flat(flat: torch.Tensor, src: int = 0) -> None:
    """Model.broadcast_parameters: one broadcast of the flat parameter buffer."""
    if dist.is_initialized() and dist.get_world_size() > 1:
        dist.broadcast(flat, src)


def all_reduce_gradients(flat_grad: torch.Tensor) -> float:
    """Model.reduce_parameters: all-reduce(SUM) of ONE flat gradient buffer, in place.  Returns the factor the
    consumer must apply (1 / world_size) — the Adam kernel multiplies by it instead of a separate div pass."""
    if dist.is_initialized() and dist.get_world_size() > 1:
        dist.all_reduce(flat_grad, op=dist.ReduceOp.SUM)
        return 1.0 / dist.get_world_size()
    return 1.0


def reduce_sums(*tensors: torch.Tensor) -> None:
    """One-off SUM reductions of small statistics (per-speed-bin tracking-error sums of the eval sweep)."""
    if dist.is_initialized() and dist.get_world_size() > 1:
        for t in tensors:
            dist.all_reduce(t, op=dist.ReduceOp.SUM)


def max_over_ranks(value: float, device) -> float:
    """Timing rule of bench.py: a multi-GPU duration is the MAX over ranks."""
    if dist.is_initialized() and dist.get_world_size() > 1:
        t = torch.tensor([value], dtype=torch.float64, device=device)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())
    return float(value)


class TrackingErrorBins:
    """TE-vs-commanded-speed accumulator of the eval sweep (src/options_sim/cli/eval_artifacts.py:65-174 bins
    ``Info / rew_Tracking error`` by ``Info / rew_Desired velocity``): per-bin sums and counts on the device,
    reduced across ranks once at the end."""

    def __init__(self, lo: float = 1.0, hi: float = 2.0, bins: int = 10, device="cpu"):
        self.lo, self.hi, self.bins = float(lo), float(hi), int(bins)
        self.te_sum = torch.zeros(bins, dtype=torch.float64, device=device)
        self.count = torch.zeros(bins, dtype=torch.float64, device=device)

    def add(self, desired_v: torch.Tensor, te: torch.Tensor, mask: torch.Tensor | None = None) -> None:
        idx = ((desired_v.double() - self.lo) / (self.hi - self.lo) * self.bins).floor().clamp_(0, self.bins - 1).long()
        w = torch.ones_like(te, dtype=torch.float64) if mask is None else mask.double()
        self.te_sum.index_add_(0, idx, te.double() * w)
        self.count.index_add_(0, idx, w)

    def reduce(self) -> None:
        reduce_sums(self.te_sum, self.count)

    def means(self) -> torch.Tensor:
        return self.te_sum / self.count.clamp_min(1.0)
