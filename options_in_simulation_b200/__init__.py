"""ois-b200 — B200-native (sm_100a) rollout-and-update hot path of meesjansen/Options_In_Simulation.

Host code is Python + PyTorch (device memory, streams, torch.distributed); every hot operation is a
hand-written CUDA kernel in ``libois_b200.so`` reached through the C ABI of ``include/ois_b200.h``.
There is no CPU / eager fallback: without the library (or without a CUDA device) the product raises.
"""
from ._lib import LIB_PATH, OisError, build, lib  # noqa: F401

__all__ = ["build", "lib", "LIB_PATH", "OisError"]


def __getattr__(name):   # lazy: importing the package must not need torch.cuda
    if name in ("TorqueDistributionTask", "VecEnv", "wrap_env"):
        from . import envs
        return getattr(envs, name)
    if name in ("RandomMemory", "FIFOMemory", "Memory", "memory_size_for"):
        from . import memory
        return getattr(memory, name)
    if name in ("DDPG", "DDPG_DEFAULT_CONFIG", "DeterministicMLP"):
        from . import agent
        return getattr(agent, name)
    if name in ("SequentialTrainer",):
        from . import trainer
        return getattr(trainer, name)
    if name in ("RolloutGraph",):
        from . import rollout
        return getattr(rollout, name)
    raise AttributeError(name)
