"""Replay memories with skrl's ``Memory`` surface (``create_tensor / add_samples / sample /
sample_by_index / __len__``, attributes ``memory_index, env_index, filled, tensors, tensors_view``).

Reference: skrl ``RandomMemory`` (train/train_kamma_4d_nofifo_random_RLIL.py:104) and the inline
``FIFOMemory`` (train/train_kaddpg_1d_nofifo_random_RLIL.py:105-206) — identical sampling
(``randperm(len)[:B]``) and identical multi-env append; they differ in capacity only.
Tensors are ``(memory_size, num_envs, dim)`` in HBM; append and gather are CUDA kernels (csrc/replay.cu).
"""
from __future__ import annotations

import collections

import torch

from . import _lib as L

_NAMES = ("states", "actions", "rewards", "next_states", "terminated", "truncated")


class Memory:
    def __init__(self, memory_size: int, num_envs: int = 1, device=None, replacement: bool = False,
                 sampler: str = "host", seed: int = 0, **_ignored):
        L.lib()
        self.memory_size = int(memory_size)
        self.num_envs = int(num_envs)
        self.device = torch.device("cuda:0" if device is None else device)
        if self.device.type != "cuda":
            raise L.OisError("replay memory lives in HBM; there is no CPU path")
        self.filled = False
        self.env_index = 0
        self.memory_index = 0
        self.tensors: dict[str, torch.Tensor] = {}
        self.tensors_view: dict[str, torch.Tensor] = {}
        self.sampling_indexes = None
        self._replacement = replacement
        self.sampler = sampler          # "host": torch.randperm on the CPU like the reference; "device": Philox/Feistel kernel
        self._seed = seed
        self._draws = 0
        self.kernel_launches = 0
        self._out = None
        self._supplied_indexes = collections.deque()
        self.generation = 0             # bumped whenever a tensor is (re)created: cached device addresses go stale

    def __len__(self) -> int:   # skrl Memory.__len__
        return self.memory_size * self.num_envs if self.filled else self.memory_index * self.num_envs + self.env_index

    def create_tensor(self, name: str, size, dtype=None, keep_dimensions: bool = True) -> bool:
        size = int(size if isinstance(size, int) else size.shape[0] if hasattr(size, "shape") else size[0])
        dtype = torch.float32 if dtype is None else dtype
        t = torch.zeros(self.memory_size, self.num_envs, size, device=self.device, dtype=dtype)
        if torch.is_floating_point(t):
            t.fill_(float("nan"))   # skrl fills float tensors with NaN
        self.tensors[name] = t
        self.tensors_view[name] = t.view(-1, size)
        self.generation += 1
        return True

    def get_tensor_by_name(self, name: str, keepdim: bool = True):
        return self.tensors[name] if keepdim else self.tensors_view[name]

    def advance(self) -> None:
        """Index bookkeeping of add_samples for one full multi-env step
        (train/train_kaddpg_1d_nofifo_random_RLIL.py:158-160)."""
        self.memory_index = (self.memory_index + 1) % self.memory_size
        if self.memory_index == 0:
            self.filled = True

    def add_samples(self, **tensors: torch.Tensor) -> None:
        if not tensors:
            raise ValueError("No samples to be recorded in memory. Pass samples as key-value arguments (where key is the tensor name)")
        tmp = tensors.get("states", tensors[next(iter(tensors))])
        if not (tmp.ndim == 2 and tmp.shape[0] == self.num_envs):
            raise ValueError(f"Expected shape (number of environments = {self.num_envs}, data size), got {tuple(tmp.shape)}")
        missing = [n for n in _NAMES if n not in tensors or n not in self.tensors]
        if missing:
            raise ValueError(f"add_samples needs the DDPG transition tensors, missing {missing}")
        f = lambda n, dt: tensors[n].to(device=self.device, dtype=dt).contiguous()
        s, a = f("states", torch.float32), f("actions", torch.float32)
        r, ns = f("rewards", torch.float32), f("next_states", torch.float32)
        te, tr = f("terminated", torch.bool), f("truncated", torch.bool)
        T = self.tensors
        L.check(L.lib().ois_replay_append(
            L.ptr(T["states"]), L.ptr(T["actions"]), L.ptr(T["rewards"]), L.ptr(T["next_states"]),
            L.ptr(T["terminated"]), L.ptr(T["truncated"]), self.memory_index, self.num_envs, s.shape[1], a.shape[1],
            L.ptr(s), L.ptr(a), L.ptr(r), L.ptr(ns), L.ptr(te), L.ptr(tr), L.current_stream()))
        self.kernel_launches += 1
        self.advance()

    def _alloc_out(self, batch: int):
        if self._out is None or self._out[0].shape[0] != batch:
            T = self.tensors
            self._out = tuple(torch.empty(batch, T[n].shape[2], device=self.device, dtype=T[n].dtype) for n in _NAMES)
        return self._out

    def sample_by_index(self, names, indexes, mini_batches: int = 1):
        """Memory.sample_by_index: gather rows of the flattened (memory_size*num_envs, dim) views."""
        idx = indexes.to(device=self.device, dtype=torch.int64).contiguous()
        out = self._alloc_out(idx.numel())
        T = self.tensors
        L.check(L.lib().ois_replay_gather(
            *[L.ptr(T[n]) for n in _NAMES], L.ptr(idx), idx.numel(), self.memory_size * self.num_envs,
            T["states"].shape[2], T["actions"].shape[2], *[L.ptr(o) for o in out], L.current_stream()))
        self.kernel_launches += 1
        by_name = dict(zip(_NAMES, out))
        return [[by_name[n] for n in names]]

    def supply_indexes(self, index_sets) -> None:
        """Parity mode: the next ``sample`` calls use these index vectors (what the reference's host
        ``torch.randperm(len)[:B]`` produced) instead of drawing their own."""
        self._supplied_indexes.extend(torch.as_tensor(i, dtype=torch.int64) for i in index_sets)

    def draw_indices(self, idx: torch.Tensor) -> None:
        """Device sampler into a caller-owned index buffer (the CUDA-graph update keeps one fixed buffer)."""
        L.check(L.lib().ois_replay_sample_indices(L.ptr(idx), idx.numel(), len(self), self._seed, self._draws,
                                                  L.current_stream()))
        self._draws += 1
        self.kernel_launches += 1
        self.sampling_indexes = idx

    def sample_gather(self, names, idx: torch.Tensor, ctl: torch.Tensor):
        """Sampler and gather as ONE launch driven by a device-resident OisSampleCtl (`ctl`: the struct as four int64
        words): what a CUDA-graph-captured DDPG._update replays.  The draw counter advances on the device; the host
        mirror follows arithmetically."""
        out = self._alloc_out(idx.numel())
        T = self.tensors
        L.check(L.lib().ois_replay_sample_gather(
            *[L.ptr(T[n]) for n in _NAMES], L.ptr(idx), idx.numel(), T["states"].shape[2], T["actions"].shape[2],
            *[L.ptr(o) for o in out], self._seed, L.ptr(ctl), L.current_stream()))
        self._draws += 1
        self.kernel_launches += 1
        self.sampling_indexes = idx
        by_name = dict(zip(_NAMES, out))
        return [[by_name[n] for n in names]]

    def sample(self, names, batch_size: int, mini_batches: int = 1, sequence_length: int = 1):
        size = len(self)
        if self._supplied_indexes:
            idx = self._supplied_indexes.popleft()
            if idx.numel() != batch_size or int(idx.max()) >= size:
                raise L.OisError("supplied sampling indexes do not fit this memory / batch size")
        elif self.sampler == "device":
            idx = torch.empty(batch_size, dtype=torch.int64, device=self.device)
            L.check(L.lib().ois_replay_sample_indices(L.ptr(idx), batch_size, size, self._seed, self._draws,
                                                      L.current_stream()))
            self._draws += 1
            self.kernel_launches += 1
        elif self._replacement:
            idx = torch.randint(0, size, (batch_size,))
        else:   # train/train_kaddpg_1d_nofifo_random_RLIL.py:139 — host randperm, like the reference
            idx = torch.randperm(size, dtype=torch.long)[:batch_size]
        self.sampling_indexes = idx
        return self.sample_by_index(names=names, indexes=idx, mini_batches=mini_batches)


class RandomMemory(Memory):
    """skrl RandomMemory(memory_size, num_envs, device, replacement=False)."""


class FIFOMemory(Memory):
    """The reference's inline ring buffer (same semantics; smaller capacity wraps within a run)."""


def memory_size_for(total_transitions: int, num_envs: int) -> int:
    """Capacity re-expressed in steps: the scripts' ``memory_size`` counts transitions at numEnvs = 1
    (1 000 000 / 35 000 / 10 000); at N envs keep the total."""
    return max(1, -(-int(total_transitions) // int(num_envs)))
