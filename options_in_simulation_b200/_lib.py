"""ctypes binding of ``libois_b200.so`` (C ABI: ``include/ois_b200.h``).

There is no CPU fallback: if the shared library is missing or a call fails, an exception is raised.
The library is built in-tree by ``options_in_simulation_b200.build()`` (nvcc, sm_100a only).
"""
from __future__ import annotations

import ctypes as C
import os
import subprocess

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "libois_b200.so")
CSRC = os.path.join(_HERE, "csrc")

OIS_EP_STATS_LEN = 24
OIS_COMP0_LEN = 32
OIS_MAX_RANKS = 8
OIS_IPC_HANDLE_BYTES = 64
BLEND_LINEAR, BLEND_SWITCH = 0, 1
GAMMA_CONST, GAMMA_DECAY = 0, 1
CMD_RAND, CMD_GV, CMD_EVAL, CMD_BD = 0, 1, 2, 3
RELABEL_NONE, RELABEL_EXECUTED = 0, 1
ACT_NONE, ACT_SIGMOID, ACT_TANH = 0, 1, 2
PRECISION_FP32, PRECISION_BF16 = 0, 1

_f, _i32, _i64, _u64, _p = C.c_float, C.c_int32, C.c_int64, C.c_uint64, C.c_void_p


class OisError(RuntimeError):
    pass


class EnvParams(C.Structure):
    _fields_ = [(n, _i32) for n in (
        "act_dim", "blend", "gamma1_mode", "gamma2_mode", "guiding_post_clip", "command_mode", "physics",
        "eval_sums", "decimation", "max_episode_length")] + [(n, _f) for n in (
        "gamma1_const", "gamma2_const", "max_sim_steps", "kp", "kp_omega", "vehicle_mass", "vehicle_inertia", "dt",
        "action_scale", "torque_clip", "bound", "x_lo", "x_range", "y_lo", "y_range", "h_lo", "h_range", "x_hi",
        "max_episode_length_s", "mass", "yaw_inertia", "wheel_inertia", "wheel_radius", "half_wheelbase",
        "half_track", "mu_n", "slip_eps", "wheel_damping", "wheel_speed_max", "sim_dt")]


class EnvState(C.Structure):
    _fields_ = [(n, _p) for n in (
        "pose_vx", "twist_prev", "wheel_speed", "commands", "sums_a", "sums_b", "sums_eval", "old_torques",
        "sim_steps", "episode_buf", "episode_count")]


class EnvIO(C.Structure):
    _fields_ = [
        ("actions", _p), ("rand_mask", _p), ("reset_uniforms", _p), ("philox_seed", _u64), ("philox_offset", _u64),
        ("obs", _p), ("rew", _p), ("reset", _p), ("wheel_torqs", _p), ("timeout", _p), ("torques", _p),
        ("guiding", _p), ("ep_stats", _p), ("comp0", _p), ("any_reset_flags", _p), ("flag_slot", _i32),
        ("replay_states", _p), ("replay_actions", _p), ("replay_rewards", _p), ("replay_next_states", _p),
        ("replay_terminated", _p), ("replay_truncated", _p), ("relabel", _i32), ("clip_obs", _f),
        ("clip_actions", _f), ("obs_clipped", _p), ("ctl", _p), ("done_flag", _p),
    ]


class StepCtl(C.Structure):
    """OisStepCtl: per-step scalars kept on the device for CUDA-graph replay of the rollout step."""
    _fields_ = [("step", _i64), ("env_philox_offset", _u64), ("act_philox_offset", _u64), ("replay_row", _i64),
                ("replay_rows", _i64), ("flag_slot", _i32), ("noise_scale", _f), ("timestep", C.c_double),
                ("expl_timesteps", C.c_double), ("expl_initial_scale", C.c_double), ("expl_final_scale", C.c_double),
                ("self_advance", _i32), ("ticket", C.c_uint32)]


class SampleCtl(C.Structure):
    """OisSampleCtl: per-update scalars kept on the device for CUDA-graph replay of DDPG._update."""
    _fields_ = [("len", _i64), ("draws", _u64), ("adam_step", _i64), ("ticket", C.c_uint32), ("pad_", C.c_uint32)]


class MlpShape(C.Structure):
    _fields_ = [(n, _i32) for n in ("in_dim", "h0", "h1", "h2", "out_dim", "out_act")]


class Scaler(C.Structure):
    _fields_ = [("running_mean", _p), ("running_variance", _p), ("current_count", _p)]


class ActIO(C.Structure):
    _fields_ = [
        ("states", _p), ("actions", _p), ("ou_state", _p), ("normals", _p), ("uniforms", _p),
        ("philox_seed", _u64), ("philox_offset", _u64), ("random_phase", _i32),
        ("theta", _f), ("sigma", _f), ("base_scale", _f), ("noise_scale", _f), ("clip_low", _f), ("clip_high", _f),
        ("noise_stats", _p), ("ctl", _p),
    ]


class UpdateIO(C.Structure):
    _fields_ = [(n, _p) for n in ("states", "actions", "rewards", "next_states", "terminated", "truncated",
                                  "actor", "actor_target", "critic", "critic_target",
                                  "actor_m", "actor_v", "critic_m", "critic_v")] + [
        ("adam_step", _i64), ("actor_lr", _f), ("critic_lr", _f), ("beta1", _f), ("beta2", _f), ("adam_eps", _f),
        ("discount", _f), ("polyak", _f), ("actor_grad", _p), ("critic_grad", _p), ("stats", _p),
        ("adam_step_dev", _p), ("comm", _p), ("batch_rows", _i32)]


# every symbol include/ois_b200.h declares: (name, restype, argtypes)
_PP = C.POINTER
SYMBOLS = {
    "ois_env_step": (C.c_int, [_PP(EnvParams), _PP(EnvState), _PP(EnvIO), _i64, _p]),
    "ois_dynamics_step": (C.c_int, [_PP(EnvParams), _PP(EnvState), _p, _i64, _p]),
    "ois_env_reset_idx": (C.c_int, [_PP(EnvParams), _PP(EnvState), _p, _i64, _p, _u64, _u64, _i64, _p]),
    "ois_replay_append": (C.c_int, [_p] * 6 + [_i64, _i64, _i32, _i32] + [_p] * 6 + [_p]),
    "ois_replay_gather": (C.c_int, [_p] * 6 + [_p, _i64, _i64, _i32, _i32] + [_p] * 6 + [_p]),
    "ois_replay_sample_indices": (C.c_int, [_p, _i64, _i64, _u64, _u64, _p]),
    "ois_sample_ctl_set": (C.c_int, [_p, _i64, _u64, _i64, _p]),
    "ois_replay_sample_gather": (C.c_int, [_p] * 6 + [_p, _i64, _i32, _i32] + [_p] * 6 + [_u64, _p, _p]),
    "ois_agent_create": (C.c_int, [_PP(_p), _PP(MlpShape), _PP(MlpShape), _i64, _i32, _i32, _p]),
    "ois_agent_destroy": (C.c_int, [_p]),
    "ois_agent_invalidate": (C.c_int, [_p]),
    "ois_agent_act": (C.c_int, [_p, _p, _PP(Scaler), _PP(ActIO), _i64, _p]),
    "ois_agent_sync_weights": (C.c_int, [_p, _p, C.c_int64, _p]),
    "ois_rollout_fuse_begin": (C.c_int, []),
    "ois_rollout_fuse_end": (C.c_int, []),
    "ois_rollout_fused_launches": (_i64, []),
    "ois_scaler_forward": (C.c_int, [_PP(Scaler), _p, _p, _i64, _i32, _i32, _p]),
    "ois_mlp_forward": (C.c_int, [_p, _PP(MlpShape), _p, _p, _p, _i64, _p]),
    "ois_mlp_forward_static": (C.c_int, [_PP(MlpShape), _p, _p, _p, _i64, _p]),
    "ois_ddpg_update_phase": (C.c_int, [_p, _PP(Scaler), _PP(UpdateIO), _i32, _f, _p]),
    "ois_step_ctl_advance": (C.c_int, [_p, _p]),
    "ois_graph_launch": (C.c_int, [_p, _p]),
    "ois_host_wait": (C.c_int, [_p, C.c_uint32, _i64]),
    "ois_graph_launch_wait": (C.c_int, [_p, _p, _p, C.c_uint32, _i64]),
    "ois_comm_create": (C.c_int, [_PP(_p), _i32, _i32, _i64, _i64]),
    "ois_comm_export": (C.c_int, [_p, _p]),
    "ois_comm_connect": (C.c_int, [_p, _p, _i32]),
    "ois_comm_connect_local": (C.c_int, [_p, _PP(_p), _i32]),
    "ois_comm_grad": (_p, [_p, _i32]),
    "ois_comm_status": (C.c_int, [_p, _PP(C.c_uint32)]),
    "ois_comm_debug_publish": (C.c_int, [_p, _i32, C.c_uint32]),
    "ois_comm_destroy": (C.c_int, [_p]),
    "ois_last_error_string": (C.c_char_p, []),
    "ois_abi_version": (C.c_int, []),
    "ois_build_info": (C.c_char_p, []),
    "ois_struct_size": (C.c_int, [C.c_int]),
}

_lib = None


def build(verbose: bool = False) -> str:
    """Compile every CUDA source for sm_100a into ``libois_b200.so`` (in-tree, via csrc/Makefile)."""
    proc = subprocess.run(["make", "-C", CSRC, "-j8"], capture_output=True, text=True)
    if verbose or proc.returncode != 0:
        print(proc.stdout)
        print(proc.stderr)
    if proc.returncode != 0:
        raise OisError(f"building libois_b200.so failed (make exit {proc.returncode})")
    global _lib
    _lib = None
    return LIB_PATH


def lib() -> C.CDLL:
    """Load the shared library (once).  Raises if it has not been built: there is no other path."""
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise OisError(
                f"{LIB_PATH} is missing: build it with `python -c 'import options_in_simulation_b200 as o; o.build()'`"
                " — there is no CPU or eager fallback")
        handle = C.CDLL(LIB_PATH)
        for name, (res, args) in SYMBOLS.items():
            fn = getattr(handle, name)    # AttributeError if the symbol is not exported
            fn.restype = res
            fn.argtypes = args
        if handle.ois_abi_version() != 3:
            raise OisError("libois_b200.so ABI version mismatch")
        _lib = handle
    return _lib


def check(code: int) -> None:
    if code != 0:
        msg = lib().ois_last_error_string().decode("utf-8", "replace")
        raise OisError(f"libois_b200 error {code}: {msg}")


def ptr(t) -> int | None:
    """Device pointer of a torch tensor (None -> NULL).  The tensor must be contiguous and on CUDA."""
    if t is None:
        return None
    if not t.is_cuda:
        raise OisError("libois_b200 entry points take CUDA tensors only (no CPU fallback)")
    if not t.is_contiguous():
        raise OisError("tensor must be contiguous")
    return t.data_ptr()


def current_stream(device_index: int | None = None) -> int:
    """Raw cudaStream_t of torch's current stream (under torch.cuda.graph capture: the capture stream).  This sits on
    the host's critical path of every step, hence the direct binding: torch.cuda.current_stream() builds a Stream
    object and resolves the device through several Python layers (~5 us of the ~17 us a call used to cost)."""
    import torch

    if device_index is None:
        device_index = torch.cuda.current_device()
    return torch._C._cuda_getCurrentRawStream(device_index)
