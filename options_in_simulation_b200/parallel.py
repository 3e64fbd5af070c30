"""Data-parallel plumbing: one process per GPU, launched by torchrun.

skrl's DP semantics, as the reference uses them through ``Model.broadcast_parameters`` /
``Model.reduce_parameters`` (src/my_agents/ddpg.py:122-127,409-410,431-432) and its vendored ``Agent`` base
(src/my_agents/base.py:88-92): envs and replay shards are rank-local and never exchanged; parameters are
broadcast from rank 0 once; each update all-reduces (SUM) ONE flat gradient buffer per network and divides by
the world size (here the division is folded into the Adam kernel as ``grad_scale``).

The plain functions touch no CUDA API, so they run unchanged on the ``gloo`` backend with CPU tensors
(tests/test_distributed_cpu.py, world_size 2) and on ``nccl`` over NVLink 5 / NVSwitch on the GPU box.

``PeerExchange`` is the B200 data path of the same two reductions: every rank's flat gradient buffers live in an
exchange block that all peers map over NVLink (CUDA IPC), and the Adam kernels read all ranks' gradients in place
(csrc/comm.cuh) — no NCCL call, no host involvement per update, capturable in the update's CUDA graph.  The only thing
it asks of ``torch.distributed`` is to carry the 64-byte IPC handles between the ranks once, at construction.
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


class _DeviceMemory:
    """A raw device allocation owned by the C library, exposed to torch through ``__cuda_array_interface__``."""

    def __init__(self, ptr: int, count: int):
        self.__cuda_array_interface__ = {"shape": (int(count),), "typestr": "<f4", "data": (int(ptr), False), "version": 3}


def gather_handles(handle: bytes, world: int, group=None) -> bytes:
    """All ranks' IPC handles concatenated in rank order (the one host-side exchange, at construction)."""
    out = [None] * world
    dist.all_gather_object(out, bytes(handle), group=group)
    return b"".join(out)


class PeerExchange:
    """This rank's exchange block + the mappings of all peers' blocks (include/ois_b200.h: ``ois_comm_*``).

    ``critic_grad`` / ``actor_grad`` are torch views of the block's two gradient buffers: the update's backward kernels
    write them, the peers read them.  ``handle`` goes into ``OisUpdateIO.comm``."""

    def __init__(self, critic_params: int, actor_params: int, device, rank: int | None = None, world: int | None = None,
                 handles: bytes | None = None):
        import ctypes as C

        from . import _lib as L

        info = distributed_info()
        self.rank = info.rank if rank is None else int(rank)
        self.world = info.world_size if world is None else int(world)
        self.device = torch.device(device)
        self._L, self._C = L, C
        h = C.c_void_p()
        with torch.cuda.device(self.device):
            L.check(L.lib().ois_comm_create(C.byref(h), self.rank, self.world, int(critic_params), int(actor_params)))
            self.handle = h
            buf = C.create_string_buffer(L.OIS_IPC_HANDLE_BYTES)
            if self.world > 1 and handles != b"":       # b"": the caller connects later (create())
                if handles is None:
                    L.check(L.lib().ois_comm_export(h, buf))
                    handles = gather_handles(buf.raw, self.world)
                L.check(L.lib().ois_comm_connect(h, handles, self.world))
        lib = L.lib()
        self.critic_grad = torch.as_tensor(_DeviceMemory(lib.ois_comm_grad(h, 0), critic_params), device=self.device)
        self.actor_grad = torch.as_tensor(_DeviceMemory(lib.ois_comm_grad(h, 1), actor_params), device=self.device)

    @classmethod
    def create(cls, critic_params: int, actor_params: int, device):
        """Collective constructor for the torchrun case: every rank either gets a connected exchange or — if CUDA IPC
        is unavailable on ANY rank (container without a shared IPC namespace, no P2P path) — all of them get ``None``
        and fall back to NCCL together.  Two agreement points, so that no rank is left waiting in a collective."""
        import ctypes as C

        from . import _lib as L

        info = distributed_info()
        dev = torch.device(device)

        def all_ok(ok: bool) -> bool:
            t = torch.tensor([1 if ok else 0], dtype=torch.int32, device=dev)
            dist.all_reduce(t, op=dist.ReduceOp.MIN)
            return bool(t.item())

        self, handle, err = None, b"\0" * L.OIS_IPC_HANDLE_BYTES, None
        try:
            self = cls(critic_params, actor_params, dev, info.rank, info.world_size, handles=b"")
            buf = C.create_string_buffer(L.OIS_IPC_HANDLE_BYTES)
            with torch.cuda.device(dev):
                L.check(L.lib().ois_comm_export(self.handle, buf))
            handle = buf.raw
        except L.OisError as e:
            err = e
        ok = all_ok(err is None)
        handles = gather_handles(handle, info.world_size)
        if ok:
            try:
                with torch.cuda.device(dev):
                    L.check(L.lib().ois_comm_connect(self.handle, handles, info.world_size))
            except L.OisError as e:
                err = e
            ok = all_ok(err is None)
        if not ok:
            if info.rank == 0:
                print(f"[ois] peer-memory gradient exchange unavailable ({err or 'on another rank'}); using NCCL", flush=True)
            if self is not None and self.handle is not None:
                L.lib().ois_comm_destroy(self.handle)
                self.handle = None
            return None
        return self

    def status(self) -> tuple[int, int, int]:
        """(critic exchanges completed, actor exchanges completed, 0 or p + 1 if a barrier gave up waiting for rank p)."""
        w = (self._C.c_uint32 * 3)()
        with torch.cuda.device(self.device):
            self._L.check(self._L.lib().ois_comm_status(self.handle, w))
        return int(w[0]), int(w[1]), int(w[2])

    def close(self) -> None:
        if self.handle is not None:
            torch.cuda.synchronize(self.device)
            if dist.is_initialized() and self.world > 1:
                dist.barrier()              # no peer may still be reading this block
            self.critic_grad = self.actor_grad = None
            self._L.lib().ois_comm_destroy(self.handle)
            self.handle = None


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
