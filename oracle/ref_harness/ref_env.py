"""TEST INFRASTRUCTURE ONLY — drives the UNMODIFIED reference task classes
(src/my_envs/*.py: TorqueDistributionTask) on CPU torch with a fake articulation view.

The physics boundary (``SimulationContext.step``, src/my_envs/KAMMA_medium.py:653) is replaced by a
callback: ``None`` (no-op; state tensors served unchanged) or a function ``(view, wheel_torqs) -> None``
that advances the fake view's state (used with the float64 skid-steer twin).
"""
from __future__ import annotations

import copy
import importlib
import types

import torch

from . import stubs


class FakeView:
    """Serves caller-provided state tensors (PhysX view stand-in) and records ``set_*`` calls.

    State layout follows OmniIsaacGymEnvs: positions [N,3], orientations wxyz [N,4], velocities
    [N,6] = (lin xyz, ang xyz) in the world frame, joint positions / velocities [N,4].
    """

    name = "robot_view"
    num_dof = 4

    def __init__(self, n: int):
        self.n = n
        self.pos = torch.zeros(n, 3)
        self.quat = torch.zeros(n, 4)
        self.quat[:, 0] = 1.0
        self.vel = torch.zeros(n, 6)
        self.jpos = torch.zeros(n, 4)
        self.jvel = torch.zeros(n, 4)
        self.efforts = torch.zeros(n, 4)
        self.effort_calls = 0

    # getters (clone=False returns the live buffer, as the GPU pipeline does)
    def get_joint_positions(self, clone=True):
        return self.jpos.clone() if clone else self.jpos

    def get_joint_velocities(self, clone=True):
        return self.jvel.clone() if clone else self.jvel

    def get_world_poses(self, clone=True):
        return (self.pos.clone(), self.quat.clone()) if clone else (self.pos, self.quat)

    def get_velocities(self, clone=True):
        return self.vel.clone() if clone else self.vel

    # setters
    def set_world_poses(self, positions=None, orientations=None, indices=None):
        idx = indices.long()
        self.pos[idx] = positions
        self.quat[idx] = orientations

    def set_velocities(self, velocities=None, indices=None):
        if indices is None:
            self.vel[:] = velocities
        else:
            self.vel[indices.long()] = velocities

    def set_joint_efforts(self, efforts, indices=None):
        self.effort_calls += 1
        if indices is None:
            self.efforts[:] = efforts
        else:
            self.efforts[indices.long()] = efforts

    def set_joint_velocities(self, velocities=None, indices=None):
        if indices is None:
            self.jvel[:] = velocities
        else:
            self.jvel[indices.long()] = velocities


def make_task(module: str, num_envs: int, *, stiffness=0.035, linear_x=(1.0, 2.0), physics=None, seed=0,
              _pre_post_reset=None):
    """Instantiate ``my_envs.<module>.TorqueDistributionTask`` the way the train scripts do
    (train/train_kamma_4d_nofifo_random_RLIL.py:77-95), on CPU, and run ``post_reset``.

    Returns ``(task, view, mod)``.
    """
    stubs.install()
    mod = importlib.import_module(f"my_envs.{module}")
    cfg = copy.deepcopy(mod.TASK_CFG)
    cfg["sim_device"] = "cpu"
    cfg["seed"] = seed
    cfg["task"]["env"]["numEnvs"] = num_envs
    cfg["task"]["env"]["control"]["stiffness"] = stiffness
    cfg["task"]["env"]["control"]["damping"] = 0.005
    cfg["task"]["env"]["randomCommandVelocityRanges"]["linear_x"] = list(linear_x)
    cfg["task"]["env"]["randomCommandVelocityRanges"]["yaw_constant"] = 0.5
    sim_config = types.SimpleNamespace(config=cfg, task_config=cfg["task"])
    world = types.SimpleNamespace(is_playing=lambda: True)
    env = types.SimpleNamespace(world=world, render_enabled=False)
    torch.manual_seed(seed)
    task = mod.TorqueDistributionTask(name="TorqueDistribution", sim_config=sim_config, env=env)
    view = FakeView(num_envs)
    task._robots = view

    def _step(world_, render=False):
        if physics is not None:
            physics(view, task.wheel_torqs)

    mod.SimulationContext.step = staticmethod(_step)
    if _pre_post_reset is not None:
        _pre_post_reset(task)
    import contextlib, io

    with contextlib.redirect_stdout(io.StringIO()):
        task.post_reset()
    return task, view, mod


# ---------------------------------------------------------------------------------------------
class RefRunner:
    """Runs the reference task with *supplied* random draws and (optionally) the float64 skid-steer
    twin as the physics callback, so the CUDA path / numpy oracle can be fed the identical numbers.

    Draw order inside ``reset_idx`` (src/my_envs/KAMMA_medium.py:424-456,571-572), per reset env:
    col 0,1 pos xy | 2 angle | 3 cmd x | 4 cmd y | 5 heading | 6 loop x_vel | 7 loop omega (unused).
    """

    def __init__(self, module: str, num_envs: int, *, physics: bool, init_uniforms, skid_params=None, **kw):
        import numpy as np
        from oracle.dynamics_twin import SkidSteerParams

        self.np = np
        self.n = num_envs
        self.skid = skid_params or SkidSteerParams()
        self._substep = 0
        self._uniforms = torch.as_tensor(np.asarray(init_uniforms, dtype=np.float32))
        self._ids = None
        self._call = 0
        stubs.install()
        self.mod = importlib.import_module(f"my_envs.{module}")
        self._orig_trf = self.mod.torch_rand_float
        self.mod.torch_rand_float = self._rand_float
        self.task, self.view, _ = make_task(module, num_envs, physics=self._physics if physics else None,
                                            _pre_post_reset=self._wrap_reset, **kw)

    # -- hooks
    def _wrap_reset(self, task):
        orig = task.reset_idx

        def reset_idx(env_ids):
            self._ids = env_ids.long()
            self._call = 0
            return orig(env_ids)

        task.reset_idx = reset_idx

    def _rand_float(self, lower, upper, shape, device):
        k = self._call
        self._call += 1
        U = self._uniforms
        if k == 0:
            u = U[self._ids][:, 0:2]
        elif k in (1, 2, 3, 4):
            u = U[self._ids][:, k + 1 : k + 2]
        else:
            env = self._ids[(k - 5) // 2]
            u = U[env, 6 + (k - 5) % 2].view(1, 1)
        assert tuple(u.shape) == tuple(shape), (k, u.shape, shape)
        return (upper - lower) * u + lower

    def _physics(self, view, wheel_torqs):
        # 10 identical calls per policy step (constant torques): integrate all sub-steps on the first
        np = self.np
        from oracle.dynamics_twin import skid_substep

        k = self._substep
        self._substep = (k + 1) % self.skid.decimation
        if k != 0:
            return
        yaw = 2.0 * np.arctan2(view.quat[:, 3].double().numpy(), view.quat[:, 0].double().numpy())
        c, s = np.cos(yaw), np.sin(yaw)
        vwx, vwy = view.vel[:, 0].double().numpy(), view.vel[:, 1].double().numpy()
        st = {"px": view.pos[:, 0].double().numpy(), "py": view.pos[:, 1].double().numpy(), "yaw": yaw,
              "vx": c * vwx + s * vwy, "vy": -s * vwx + c * vwy, "wz": view.vel[:, 5].double().numpy(),
              "w": view.jvel.double().numpy()}
        tau = wheel_torqs.double().numpy()
        for _ in range(self.skid.decimation):
            st = skid_substep(st, tau, self.skid, np.float64)
        self.set_state(st)

    def set_state(self, st):
        """Write a body-frame state dict (px, py, yaw, vx, vy, wz, w) into the fake PhysX view."""
        np, view = self.np, self.view
        f = lambda a: torch.as_tensor(np.asarray(a, dtype=np.float32))
        yaw = np.asarray(st["yaw"], dtype=np.float64)
        c, s = np.cos(yaw), np.sin(yaw)
        vx, vy = np.asarray(st["vx"], np.float64), np.asarray(st["vy"], np.float64)
        view.pos[:, 0], view.pos[:, 1] = f(st["px"]), f(st["py"])
        view.quat[:, 0], view.quat[:, 3] = f(np.cos(yaw / 2)), f(np.sin(yaw / 2))
        view.quat[:, 1:3] = 0
        view.vel[:, 0], view.vel[:, 1] = f(c * vx - s * vy), f(s * vx + c * vy)
        view.vel[:, 5] = f(st["wz"])
        view.jvel[:] = f(st["w"])

    def step(self, actions, rand_mask=None, reset_uniforms=None):
        import contextlib, io

        if reset_uniforms is not None:
            self._uniforms = torch.as_tensor(self.np.asarray(reset_uniforms, dtype=self.np.float32))
        orig_rand = torch.rand
        if rand_mask is not None:
            rm = torch.as_tensor(self.np.asarray(rand_mask, dtype=self.np.float32)).view(self.n, 1)

            def fake_rand(*size, **kw):
                assert tuple(size) == (self.n, 1), size
                return rm.clone()

            torch.rand = fake_rand
        orig_normal = torch.normal
        bimodal = getattr(self.task, "gridsampling", False)     # KAMMA_curr2.py:601-641 draws inside reset_idx
        calls = {"rand": 0, "normal": 0}

        def reset_rand(*size, **kw):
            # per reset env: `torch.rand(1) < 0.5` for x (supplied column 6), then once more for the discarded omega
            assert tuple(size) == (1,), size
            k = calls["rand"]
            calls["rand"] += 1
            env = self._ids[k // 2]
            return self._uniforms[env, 6].view(1).clone() if k % 2 == 0 else torch.full((1,), 0.75)

        def reset_normal(mean, std, **kw):
            # torch.normal(mean, std) == randn * std + mean; the standard normal is supplied column 7 (x draw only)
            k = calls["normal"]
            calls["normal"] += 1
            env = self._ids[k // 2]
            z = self._uniforms[env, 7] if k % 2 == 0 else torch.tensor(0.0)
            return (z * std + mean).to(torch.float32)

        try:
            with contextlib.redirect_stdout(io.StringIO()):     # KAMMA_eval prints every step
                self.task.pre_physics_step(torch.as_tensor(self.np.asarray(actions, dtype=self.np.float32)))
                torch.rand = orig_rand
                if bimodal:
                    torch.rand, torch.normal = reset_rand, reset_normal
                out = self.task.post_physics_step()
        finally:
            torch.rand, torch.normal = orig_rand, orig_normal
        return out
