"""TEST INFRASTRUCTURE ONLY — see package docstring.  ``install()`` makes ``import my_envs.KAMMA_medium``,
``import my_agents.ddpg`` and ``import my_trainers.sequential`` work against the read-only
reference tree, by serving mocks for Isaac Sim / pxr / warp / OmniIsaacGymEnvs and the skrl shim.
"""
from __future__ import annotations

import importlib.abc
import importlib.machinery
import os
import sys
import types
from unittest.mock import MagicMock

import numpy as np
import torch

REFERENCE_ROOT = os.environ.get("OIS_REFERENCE_ROOT", "/root/reference")
_MOCK_ROOTS = ("omni", "pxr", "warp", "omniisaacgymenvs")


class _MockModule(types.ModuleType):
    """A package-like module whose every attribute is a MagicMock (and ``import *`` works)."""

    def __init__(self, name):
        super().__init__(name)
        self.__path__ = []
        self.__all__ = []

    def __getattr__(self, item):
        if item.startswith("__"):
            raise AttributeError(item)
        m = MagicMock(name=f"{self.__name__}.{item}")
        setattr(self, item, m)
        return m


class _MockFinder(importlib.abc.MetaPathFinder, importlib.abc.Loader):
    def find_spec(self, fullname, path=None, target=None):
        if fullname.split(".")[0] in _MOCK_ROOTS and fullname not in sys.modules:
            return importlib.machinery.ModuleSpec(fullname, self, is_package=True)
        return None

    def create_module(self, spec):
        return _MockModule(spec.name)

    def exec_module(self, module):
        pass


def _register(name: str, **attrs) -> types.ModuleType:
    mod = _MockModule(name)
    for k, v in attrs.items():
        setattr(mod, k, v)
    mod.__all__ = [k for k in attrs]
    sys.modules[name] = mod
    parent, _, child = name.rpartition(".")
    if parent:
        if parent not in sys.modules:
            _register(parent)
        setattr(sys.modules[parent], child, mod)
    return mod


# ---------------------------------------------------------------------- real stand-ins
class Box:
    """gym.spaces.Box / gymnasium.spaces.Box stand-in (only what the reference reads)."""

    def __init__(self, low, high, shape=None, dtype=np.float32):
        if shape is None:
            shape = np.asarray(low).shape
        self.shape = tuple(shape)
        self.dtype = dtype
        self.low = np.broadcast_to(np.asarray(low, dtype=dtype), self.shape).copy()
        self.high = np.broadcast_to(np.asarray(high, dtype=dtype), self.shape).copy()


class Space:
    pass


class BaseTask:
    def __init__(self, name=None, offset=None):
        self._name = name

    def set_up_scene(self, scene):
        pass


class RLTaskInterface(BaseTask):
    @property
    def num_envs(self):
        return self._num_envs

    @property
    def device(self):
        return self._device

    @property
    def num_observations(self):
        return self._num_observations

    @property
    def num_actions(self):
        return self._num_actions

    @property
    def num_states(self):
        return self._num_states

    @property
    def num_agents(self):
        return self._num_agents

    def get_states(self):
        return self.states_buf

    def get_extras(self):
        return self.extras


class ArticulationView:
    def __init__(self, *a, **k):
        pass


class Robot:
    def __init__(self, *a, **k):
        pass


class SimulationContext:
    """``SimulationContext.step(world, render=False)`` is the physics boundary
    (src/my_envs/KAMMA_medium.py:653); tests patch ``step`` with a callback."""

    step_callback = None

    @staticmethod
    def step(world, render=False):
        if SimulationContext.step_callback is not None:
            SimulationContext.step_callback()


def torch_rand_float(lower, upper, shape, device):
    return (upper - lower) * torch.rand(*shape, device=device) + lower


_INSTALLED = False


def install() -> None:
    global _INSTALLED
    if _INSTALLED:
        return
    if not os.path.isdir(REFERENCE_ROOT):
        raise RuntimeError(f"reference tree not found at {REFERENCE_ROOT}; the harness only runs in the build container")
    np.Inf = np.inf  # numpy >= 2 (src/my_envs/rl_task.py:99 uses np.Inf)

    from . import _isaac_rotations as rot
    from . import skrl_shim as sk

    sys.meta_path.insert(0, _MockFinder())

    _register("omni.isaac.core.utils.torch.rotations", **{n: getattr(rot, n) for n in rot.__all__})
    _register("omni.isaac.core.utils.torch.maths", torch_rand_float=torch_rand_float)
    _register("omni.isaac.core.articulations", ArticulationView=ArticulationView)
    _register("omni.isaac.core.robots.robot", Robot=Robot)
    _register("omni.isaac.core.tasks", BaseTask=BaseTask)
    _register("omni.isaac.gym.tasks.rl_task", RLTaskInterface=RLTaskInterface)
    _register("omni.isaac.core.simulation_context", SimulationContext=SimulationContext)

    spaces = _register("gym.spaces", Box=Box, Space=Space)
    gym = sys.modules["gym"]
    gym.spaces, gym.Space = spaces, Space
    gspaces = _register("gymnasium.spaces", Box=Box, Space=Space)
    gymn = sys.modules["gymnasium"]
    gymn.spaces, gymn.Space = gspaces, Space

    # skrl shim
    _register("skrl", config=sk.config, logger=sk.logger)
    _register("skrl.models.torch", Model=sk.Model, DeterministicMixin=sk.DeterministicMixin)
    _register("skrl.memories.torch", Memory=sk.Memory, RandomMemory=sk.RandomMemory)
    _register("skrl.resources.preprocessors.torch", RunningStandardScaler=sk.RunningStandardScaler)
    _register("skrl.resources.noises.torch", OrnsteinUhlenbeckNoise=sk.OrnsteinUhlenbeckNoise)
    _register("skrl.utils", set_seed=sk.set_seed)
    _register("skrl.agents.torch")
    _register("skrl.trainers.torch")

    src = os.path.join(REFERENCE_ROOT, "src")
    if src not in sys.path:
        sys.path.insert(0, src)

    # skrl.agents.torch.Agent := the vendored copy (src/my_agents/base.py:19)
    import my_agents.base as _base

    sys.modules["skrl.agents.torch"].Agent = _base.Agent
    sys.modules["skrl.agents.torch"].__all__ = ["Agent"]
    _INSTALLED = True
