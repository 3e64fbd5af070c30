"""TEST INFRASTRUCTURE ONLY — minimal stand-in for the parts of ``skrl`` (Toni-SM/skrl 1.x, unpinned
in the reference: Apptainer/isaac-sim-custom.def:21-24 clones HEAD) that the reference's hot path
touches.  skrl is not vendored in /root/reference and not installable here, so every class below is
"restated, unverifiable offline" from the published skrl 1.x sources; the in-repo statements that DO
exist are followed where available:

* ``Agent``                      := the vendored copy src/my_agents/base.py:19 (bound in ``stubs.install``)
* ``Memory.add_samples/sample``  follow the inline FIFOMemory, train/train_kaddpg_1d_nofifo_random_RLIL.py:105-206
* everything else                SURVEY.md §8(c) row 3.

Call sites in the reference: src/my_agents/ddpg.py:11-14,125-136,201-206,230-239,313,377-389,410-444.
"""
from __future__ import annotations

import types

import numpy as np
import torch
import torch.nn as nn


# --------------------------------------------------------------------------- config / logger
class _TorchCfg:
    is_distributed = False
    rank = 0
    local_rank = 0
    world_size = 1


config = types.SimpleNamespace(torch=_TorchCfg())


class _Logger:
    def info(self, *a, **k):
        pass

    def warning(self, *a, **k):
        pass

    def error(self, *a, **k):
        pass


logger = _Logger()


def _space_size(space) -> int:
    if isinstance(space, int):
        return space
    if isinstance(space, (tuple, list)):
        return int(np.prod(space))
    return int(np.prod(space.shape))


# --------------------------------------------------------------------------- models
class Model(nn.Module):
    def __init__(self, observation_space, action_space, device=None):
        super().__init__()
        self.device = torch.device("cpu") if device is None else torch.device(device)
        self.observation_space = observation_space
        self.action_space = action_space
        self.num_observations = _space_size(observation_space)
        self.num_actions = _space_size(action_space)
        self._random_distribution = None

    def random_act(self, inputs, role=""):
        if self._random_distribution is None:
            self._random_distribution = torch.distributions.uniform.Uniform(
                low=torch.tensor(self.action_space.low[0], device=self.device, dtype=torch.float32),
                high=torch.tensor(self.action_space.high[0], device=self.device, dtype=torch.float32),
            )
        return (
            self._random_distribution.sample(sample_shape=(inputs["states"].shape[0], self.num_actions)),
            None,
            {},
        )

    def set_mode(self, mode: str) -> None:
        self.train(mode == "train")

    def freeze_parameters(self, freeze: bool = True) -> None:
        for p in self.parameters():
            p.requires_grad = not freeze

    def update_parameters(self, model, polyak: float = 1) -> None:
        with torch.no_grad():
            if polyak == 1:
                for p, q in zip(self.parameters(), model.parameters()):
                    p.data.copy_(q.data)
            else:
                for p, q in zip(self.parameters(), model.parameters()):
                    p.data.mul_(1 - polyak)
                    p.data.add_(polyak * q.data)

    def broadcast_parameters(self, rank: int = 0):
        object_list = [self.state_dict()]
        torch.distributed.broadcast_object_list(object_list, rank)
        self.load_state_dict(object_list[0])

    def reduce_parameters(self):
        gradients = [p.grad.view(-1) for p in self.parameters() if p.grad is not None]
        gradients = torch.cat(gradients)
        torch.distributed.all_reduce(gradients, op=torch.distributed.ReduceOp.SUM)
        offset = 0
        for p in self.parameters():
            if p.grad is not None:
                p.grad.data.copy_(
                    gradients[offset : offset + p.numel()].view_as(p.grad.data) / config.torch.world_size
                )
                offset += p.numel()


class DeterministicMixin:
    def __init__(self, clip_actions: bool = False, role: str = ""):
        self._clip_actions = clip_actions
        if clip_actions:
            self._clip_actions_min = torch.tensor(self.action_space.low, device=self.device, dtype=torch.float32)
            self._clip_actions_max = torch.tensor(self.action_space.high, device=self.device, dtype=torch.float32)

    def act(self, inputs, role=""):
        actions, outputs = self.compute(inputs, role)
        if self._clip_actions:
            actions = torch.clamp(actions, min=self._clip_actions_min, max=self._clip_actions_max)
        return actions, None, outputs


# --------------------------------------------------------------------------- memories
class Memory:
    def __init__(self, memory_size, num_envs=1, device=None, export=False, export_format="pt", export_directory=""):
        self.memory_size = memory_size
        self.num_envs = num_envs
        self.device = torch.device("cpu") if device is None else torch.device(device)
        self.filled = False
        self.env_index = 0
        self.memory_index = 0
        self.tensors = {}
        self.tensors_view = {}
        self.sampling_indexes = None

    def __len__(self) -> int:
        return self.memory_size * self.num_envs if self.filled else self.memory_index * self.num_envs + self.env_index

    def create_tensor(self, name, size, dtype=None, keep_dimensions=True) -> bool:
        size = _space_size(size)
        t = torch.zeros((self.memory_size, self.num_envs, size), device=self.device, dtype=dtype)
        self.tensors[name] = t
        self.tensors_view[name] = t.view(-1, size)
        for tensor in self.tensors.values():
            if torch.is_floating_point(tensor):
                tensor.fill_(float("nan"))
        return True

    def get_tensor_by_name(self, name, keepdim=True):
        return self.tensors[name] if keepdim else self.tensors_view[name]

    def add_samples(self, **tensors) -> None:
        tmp = tensors.get("states", tensors[next(iter(tensors))])
        dim, shape = tmp.ndim, tmp.shape
        if dim == 2 and shape[0] == self.num_envs:
            for name, tensor in tensors.items():
                if name in self.tensors:
                    self.tensors[name][self.memory_index].copy_(tensor)
            self.memory_index += 1
        elif dim == 2 and shape[0] < self.num_envs:
            for name, tensor in tensors.items():
                if name in self.tensors:
                    self.tensors[name][self.memory_index, self.env_index : self.env_index + shape[0]].copy_(tensor)
            self.env_index += shape[0]
        elif dim == 1:
            for name, tensor in tensors.items():
                if name in self.tensors:
                    self.tensors[name][self.memory_index, self.env_index].copy_(tensor)
            self.env_index += 1
        else:
            raise ValueError(f"Expected shape (number of environments = {self.num_envs}, data size), got {shape}")
        if self.env_index >= self.num_envs:
            self.env_index = 0
            self.memory_index += 1
        if self.memory_index >= self.memory_size:
            self.memory_index = 0
            self.filled = True

    def sample_by_index(self, names, indexes, mini_batches=1):
        return [[self.tensors_view[name][indexes] for name in names]]

    def sample_all(self, names, mini_batches=1, sequence_length=1):
        return [[self.tensors_view[name] for name in names]]


class RandomMemory(Memory):
    def __init__(self, memory_size, num_envs=1, device=None, export=False, export_format="pt",
                 export_directory="", replacement=True):
        super().__init__(memory_size, num_envs, device, export, export_format, export_directory)
        self._replacement = replacement

    def sample(self, names, batch_size, mini_batches=1, sequence_length=1):
        size = len(self)
        if self._replacement:
            indexes = torch.randint(0, size, (batch_size,))
        else:
            indexes = torch.randperm(size, dtype=torch.long)[:batch_size]
        self.sampling_indexes = indexes
        return self.sample_by_index(names=names, indexes=indexes, mini_batches=mini_batches)


# --------------------------------------------------------------------------- preprocessor
class RunningStandardScaler(nn.Module):
    def __init__(self, size, epsilon: float = 1e-8, clip_threshold: float = 5.0, device=None):
        super().__init__()
        self.epsilon = epsilon
        self.clip_threshold = clip_threshold
        self.device = torch.device("cpu") if device is None else torch.device(device)
        size = _space_size(size)
        self.register_buffer("running_mean", torch.zeros(size, dtype=torch.float64, device=self.device))
        self.register_buffer("running_variance", torch.ones(size, dtype=torch.float64, device=self.device))
        self.register_buffer("current_count", torch.ones((), dtype=torch.float64, device=self.device))

    def _parallel_variance(self, input_mean, input_var, input_count: int) -> None:
        delta = input_mean - self.running_mean
        total_count = self.current_count + input_count
        M2 = (
            (self.running_variance * self.current_count)
            + (input_var * input_count)
            + delta**2 * self.current_count * input_count / total_count
        )
        self.running_mean = self.running_mean + delta * input_count / total_count
        self.running_variance = M2 / total_count
        self.current_count = total_count

    def _compute(self, x, train=False, inverse=False):
        if train:
            self._parallel_variance(torch.mean(x, dim=0), torch.var(x, dim=0), x.shape[0])
        if inverse:
            return (
                torch.sqrt(self.running_variance.float())
                * torch.clamp(x, min=-self.clip_threshold, max=self.clip_threshold)
                + self.running_mean.float()
            )
        return torch.clamp(
            (x - self.running_mean.float()) / (torch.sqrt(self.running_variance.float()) + self.epsilon),
            min=-self.clip_threshold,
            max=self.clip_threshold,
        )

    def forward(self, x, train=False, inverse=False, no_grad=True):
        if no_grad:
            with torch.no_grad():
                return self._compute(x, train, inverse)
        return self._compute(x, train, inverse)


# --------------------------------------------------------------------------- noise
class OrnsteinUhlenbeckNoise:
    def __init__(self, theta, sigma, base_scale, mean=0, std=1, device=None):
        self.device = torch.device("cpu") if device is None else torch.device(device)
        self.state = 0
        self.theta = theta
        self.sigma = sigma
        self.base_scale = base_scale
        self.distribution = torch.distributions.Normal(
            loc=torch.tensor(mean, device=self.device, dtype=torch.float32),
            scale=torch.tensor(std, device=self.device, dtype=torch.float32),
        )

    def sample(self, size):
        if hasattr(self.state, "shape") and self.state.shape != torch.Size(size):
            self.state = 0
        self.state += -self.state * self.theta + self.sigma * self.distribution.sample(size)
        return self.base_scale * self.state


def set_seed(seed=None, deterministic=False):
    seed = 42 if seed is None else int(seed)
    import random

    random.seed(seed)
    np.random.seed(seed)
    torch.manual_seed(seed)
    return seed
