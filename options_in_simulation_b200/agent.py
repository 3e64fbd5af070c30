"""``DDPG`` — B200-native mirror of the reference's agent (src/my_agents/ddpg.py:66, used for both
"KA-DDPG" and "KAMMA") with the skrl ``Agent`` surface the trainers call:

    act(states, timestep, timesteps) -> (actions, None, outputs)
    record_transition(states, actions, rewards, next_states, terminated, truncated, infos, timestep, timesteps)
    pre_interaction / post_interaction / _update / init / set_mode / set_running_mode / track_data / save / load

plus the model / preprocessor / noise classes the train scripts build it from
(train/train_kamma_4d_nofifo_random_RLIL.py:21-52,104-177): ``DeterministicActor``, ``Critic``,
``RunningStandardScaler``, ``OrnsteinUhlenbeckNoise``.

Parameters, Adam moments and gradients are flat fp32 buffers in HBM (``nn.Sequential`` order); every
arithmetic step is a kernel of ``libois_b200.so`` (csrc/agent.cu, csrc/mlp_fp32.cu, csrc/mlp_tc.cu).
``mixed_precision: False`` (the reference's setting, ddpg.py:163) selects the exact fp32 path;
``mixed_precision: True`` selects bf16 tcgen05 tensor-core inference AND update (fp32 accumulate, fp32 master weights) —
the counterpart of the reference's autocast hooks (ddpg.py:165-170,233).
"""
from __future__ import annotations

import collections
import copy
import ctypes as C
import datetime
import math
import os

import numpy as np
import torch

from . import _lib as L
from . import parallel as P

DDPG_DEFAULT_CONFIG = {   # src/my_agents/ddpg.py:19-61
    "gradient_steps": 1,
    "batch_size": 64,
    "discount_factor": 0.99,
    "polyak": 0.005,
    "actor_learning_rate": 1e-3,
    "critic_learning_rate": 1e-3,
    "learning_rate_scheduler": None,
    "learning_rate_scheduler_kwargs": {},
    "state_preprocessor": None,
    "state_preprocessor_kwargs": {},
    "random_timesteps": 0,
    "learning_starts": 0,
    "grad_norm_clip": 0,
    "exploration": {"noise": None, "initial_scale": 1.0, "final_scale": 1e-3, "timesteps": None},
    "rewards_shaper": None,
    "mixed_precision": False,
    "experiment": {"directory": "", "experiment_name": "", "write_interval": "auto", "checkpoint_interval": "auto",
                   "store_separately": False, "wandb": False, "wandb_kwargs": {}},
    # extensions of this implementation (absent from the reference)
    "sampler": "host",        # "host": torch.randperm like the reference; "device": Philox/Feistel kernel (no host sync)
    "track_stats_interval": 0,  # > 0: read losses / Q stats back every that many updates (the reference does every update)
    "graph_update": False,    # capture gather + the three update phases into ONE CUDA graph and replay it (needs
                              # sampler "device"): the launches of an update become one graph launch; under torchrun the
                              # gradient exchange is part of the graph
    "dp_exchange": "peer",    # world_size > 1, how the two gradient reductions of ddpg.py:409-410,431-432 run:
                              # "peer": inside the Adam kernels over NVLink peer memory (parallel.PeerExchange, csrc/comm.cuh);
                              # "nccl": dist.all_reduce between the phases (skrl's way), also capturable
}

_LAYER_KEYS = ("net.0", "net.2", "net.4", "net.6")   # nn.Sequential indices of the Linear layers


def _space_size(space) -> int:
    if isinstance(space, int):
        return space
    if isinstance(space, (tuple, list)):
        return int(np.prod(space))
    return int(np.prod(space.shape))


def distributed_info():
    """skrl.config.torch.{is_distributed, rank, local_rank, world_size} (read from the torchrun env)."""
    d = P.distributed_info()
    return d.is_distributed, d.rank, d.local_rank, d.world_size


# ------------------------------------------------------------------------------------------- models
class DeterministicMLP:
    """in -> 512 -> 512 -> 128 -> out, ELU, optional output activation: the model family of every train
    script.  Parameters are ONE flat fp32 device buffer (layout: include/ois_b200.h OisMlpShape); the
    state_dict uses the reference's ``net.{0,2,4,6}.{weight,bias}`` names so checkpoints interchange."""

    hidden = (512, 512, 128)

    def __init__(self, observation_space, action_space, device=None, clip_actions=False, *, in_dim, out_dim, out_act):
        self.device = torch.device("cuda:0" if device is None else device)
        self.observation_space, self.action_space = observation_space, action_space
        self.num_observations, self.num_actions = _space_size(observation_space), _space_size(action_space)
        self.shape = L.MlpShape(in_dim, *self.hidden, out_dim, out_act)
        dims = (in_dim, *self.hidden, out_dim)
        self.dims = dims
        self.offsets = []
        off = 0
        for l in range(4):
            w = (off, dims[l + 1], dims[l])
            off += dims[l + 1] * dims[l]
            b = (off, dims[l + 1])
            off += dims[l + 1]
            self.offsets.append((w, b))
        self.num_params = off
        # default nn.Linear initialisation, drawn in nn.Sequential order from the global CPU generator, so the
        # same torch seed gives the same weights as the reference's model classes
        flat = torch.empty(off, dtype=torch.float32)
        for l in range(4):
            lin = torch.nn.Linear(dims[l], dims[l + 1])
            (wo, r, c), (bo, _) = self.offsets[l]
            flat[wo:wo + r * c] = lin.weight.detach().reshape(-1)
            flat[bo:bo + r] = lin.bias.detach()
        self.flat = flat.to(self.device)
        self.training = False
        self._frozen = False
        self._agent = None   # set by DDPG: the C handle that owns workspaces

    # --- skrl Model surface
    def parameters(self):
        for (wo, r, c), (bo, _) in self.offsets:
            yield self.flat[wo:wo + r * c].view(r, c)
            yield self.flat[bo:bo + r]

    def state_dict(self):
        sd = collections.OrderedDict()
        for key, ((wo, r, c), (bo, _)) in zip(_LAYER_KEYS, self.offsets):
            sd[key + ".weight"] = self.flat[wo:wo + r * c].view(r, c).clone()
            sd[key + ".bias"] = self.flat[bo:bo + r].clone()
        return sd

    def load_state_dict(self, sd, strict=True):
        for key, ((wo, r, c), (bo, _)) in zip(_LAYER_KEYS, self.offsets):
            self.flat[wo:wo + r * c].copy_(sd[key + ".weight"].reshape(-1).to(self.device, torch.float32))
            self.flat[bo:bo + r].copy_(sd[key + ".bias"].to(self.device, torch.float32))
        self._master_changed()

    def set_mode(self, mode):
        self.training = mode == "train"

    def freeze_parameters(self, freeze=True):
        self._frozen = freeze

    def update_parameters(self, model, polyak=1):
        if polyak == 1:
            self.flat.copy_(model.flat)
        else:   # p.mul_(1 - polyak); p.add_(polyak * q)  — one fused pass is what the update kernel does; this
            # standalone call keeps the Model API and uses torch only for this cold path
            self.flat.mul_(1 - polyak).add_(model.flat, alpha=polyak)
        self._master_changed()

    def broadcast_parameters(self, rank=0):
        P.broadcast_flat(self.flat, rank)
        self._master_changed()

    def _master_changed(self):
        """The fp32 master was written behind the C handle's back: its cached bf16 images (tensor-core update, actor
        stream image) and any captured update graph are stale."""
        if self._agent is not None:
            self._agent._weights_changed()

    def reduce_parameters(self, flat_grad):
        """skrl Model.reduce_parameters on the flat gradient buffer; returns the 1/world_size factor to apply."""
        return P.all_reduce_gradients(flat_grad)

    def random_act(self, inputs, role=""):
        raise L.OisError("random_act is served by DDPG.act (fused explore kernel)")

    def _input(self, inputs):
        return inputs["states"]

    def act(self, inputs, role=""):
        """Model.act: fp32 forward of already-preprocessed inputs (exact path)."""
        if self._agent is None:
            raise L.OisError("model is not attached to a DDPG agent yet")
        x = self._input(inputs).to(self.device, torch.float32).contiguous()
        return self._agent._mlp_forward(self, x), None, {}

    def compute(self, inputs, role=""):
        return self.act(inputs, role)[0], {}


class DeterministicActor(DeterministicMLP):
    """train/train_kamma_4d_nofifo_random_RLIL.py:21-36 (Sigmoid head); eval scripts end in Tanh
    (eval/eval_kamma_4d.py:36) -> ``output="tanh"``."""

    def __init__(self, observation_space, action_space, device=None, clip_actions=False, output="sigmoid"):
        act = {"sigmoid": L.ACT_SIGMOID, "tanh": L.ACT_TANH, "none": L.ACT_NONE}[output]
        super().__init__(observation_space, action_space, device, clip_actions,
                         in_dim=_space_size(observation_space), out_dim=_space_size(action_space), out_act=act)


class Critic(DeterministicMLP):
    """train/train_kamma_4d_nofifo_random_RLIL.py:38-52: Q(cat(states, taken_actions))."""

    def __init__(self, observation_space, action_space, device=None, clip_actions=False):
        super().__init__(observation_space, action_space, device, clip_actions,
                         in_dim=_space_size(observation_space) + _space_size(action_space), out_dim=1,
                         out_act=L.ACT_NONE)

    def _input(self, inputs):
        return torch.cat([inputs["states"], inputs["taken_actions"]], dim=1)


# ------------------------------------------------------------------------------------------- preprocessor / noise
class RunningStandardScaler:
    """skrl RunningStandardScaler: fp64 running statistics in HBM, standardise + clamp(+-5) in a kernel."""

    def __init__(self, size, epsilon=1e-8, clip_threshold=5.0, device=None):
        if epsilon != 1e-8 or clip_threshold != 5.0:
            raise L.OisError("the scaler kernels implement skrl's defaults (epsilon 1e-8, clip_threshold 5)")
        self.device = torch.device("cuda:0" if device is None else device)
        self.size = _space_size(size)
        self.running_mean = torch.zeros(self.size, dtype=torch.float64, device=self.device)
        self.running_variance = torch.ones(self.size, dtype=torch.float64, device=self.device)
        self.current_count = torch.ones((), dtype=torch.float64, device=self.device)
        self._c = L.Scaler(self.running_mean.data_ptr(), self.running_variance.data_ptr(), self.current_count.data_ptr())

    def c_struct(self):
        return C.byref(self._c)

    def __call__(self, x, train=False, inverse=False, no_grad=True):
        if inverse:
            raise L.OisError("inverse scaling is not on the DDPG path")
        x = x.to(self.device, torch.float32).contiguous()
        y = torch.empty_like(x)
        L.check(L.lib().ois_scaler_forward(self.c_struct(), L.ptr(x), L.ptr(y), x.shape[0], x.shape[1], int(train),
                                           L.current_stream()))
        return y

    def state_dict(self):
        return collections.OrderedDict(running_mean=self.running_mean.clone(), running_variance=self.running_variance.clone(),
                                       current_count=self.current_count.clone())

    def load_state_dict(self, sd, strict=True):
        self.running_mean.copy_(sd["running_mean"])
        self.running_variance.copy_(sd["running_variance"])
        self.current_count.copy_(sd["current_count"])


class OrnsteinUhlenbeckNoise:
    """skrl OrnsteinUhlenbeckNoise(theta, sigma, base_scale): the state lives in HBM and is advanced inside
    the explore kernel (``state += -state*theta + sigma*N(0,1)``; noise = base_scale*state)."""

    def __init__(self, theta, sigma, base_scale, mean=0, std=1, device=None):
        if mean != 0 or std != 1:
            raise L.OisError("the explore kernel draws N(0,1) (the reference's setting)")
        self.theta, self.sigma, self.base_scale = float(theta), float(sigma), float(base_scale)
        self.device = torch.device("cuda:0" if device is None else device)
        self.state = None

    def _state_for(self, shape):
        if self.state is None or tuple(self.state.shape) != tuple(shape):   # skrl resets the state on a shape change
            self.state = torch.zeros(*shape, dtype=torch.float32, device=self.device)
        return self.state


class _AdamState:
    """torch.optim.Adam-compatible state for one flat parameter buffer (state_dict interchange)."""

    def __init__(self, model: DeterministicMLP, lr: float):
        self.model = model
        self.lr, self.betas, self.eps = float(lr), (0.9, 0.999), 1e-8
        self.m = torch.zeros_like(model.flat)
        self.v = torch.zeros_like(model.flat)
        self.step = 0

    def state_dict(self):
        state = {}
        i = 0
        for (wo, r, c), (bo, _) in self.model.offsets:
            for off, n, shp in ((wo, r * c, (r, c)), (bo, r, (r,))):
                state[i] = {"step": torch.tensor(float(self.step)), "exp_avg": self.m[off:off + n].view(shp).clone(),
                            "exp_avg_sq": self.v[off:off + n].view(shp).clone()}
                i += 1
        group = {"lr": self.lr, "betas": self.betas, "eps": self.eps, "weight_decay": 0, "amsgrad": False,
                 "maximize": False, "foreach": None, "capturable": False, "differentiable": False, "fused": None,
                 "params": list(range(i))}
        return {"state": state, "param_groups": [group]}

    def load_state_dict(self, sd):
        i = 0
        for (wo, r, c), (bo, _) in self.model.offsets:
            for off, n in ((wo, r * c), (bo, r)):
                if i in sd["state"]:
                    st = sd["state"][i]
                    self.m[off:off + n].copy_(st["exp_avg"].reshape(-1))
                    self.v[off:off + n].copy_(st["exp_avg_sq"].reshape(-1))
                    self.step = int(float(st["step"]))
                i += 1
        g = sd["param_groups"][0]
        self.lr, self.betas, self.eps = float(g["lr"]), tuple(g["betas"]), float(g["eps"])


# ------------------------------------------------------------------------------------------- agent
class DDPG:
    def __init__(self, models, memory=None, observation_space=None, action_space=None, device=None, cfg=None):
        L.lib()
        _cfg = copy.deepcopy(DDPG_DEFAULT_CONFIG)
        for k, v in (cfg or {}).items():
            if isinstance(v, dict) and isinstance(_cfg.get(k), dict):
                _cfg[k] = {**_cfg[k], **v}
            else:
                _cfg[k] = v
        self.cfg = _cfg
        self.models = models
        self.observation_space, self.action_space = observation_space, action_space
        self.device = torch.device("cuda:0" if device is None else device)
        if self.device.type != "cuda":
            raise L.OisError("DDPG runs on a CUDA device (B200, sm_100a) only; there is no CPU path")
        if isinstance(memory, (tuple, list)):
            self.memory, self.secondary_memories = memory[0], list(memory[1:])
        else:
            self.memory, self.secondary_memories = memory, []
        self.policy, self.target_policy = models.get("policy"), models.get("target_policy")
        self.critic, self.target_critic = models.get("critic"), models.get("target_critic")
        for m in (self.policy, self.target_policy, self.critic, self.target_critic):
            if not isinstance(m, DeterministicMLP):
                raise L.OisError("models must be options_in_simulation_b200 DeterministicActor / Critic instances")
        self.checkpoint_modules = {"policy": self.policy, "target_policy": self.target_policy, "critic": self.critic,
                                   "target_critic": self.target_critic}
        self.is_distributed, self.rank, self.local_rank, self.world_size = distributed_info()
        if self.is_distributed:   # ddpg.py:122-127 / base.py:88-92
            P.ensure_process_group("nccl")
            self.policy.broadcast_parameters()
            self.critic.broadcast_parameters()
        self.target_policy.freeze_parameters(True)       # ddpg.py:129-136
        self.target_critic.freeze_parameters(True)
        self.target_policy.update_parameters(self.policy, polyak=1)
        self.target_critic.update_parameters(self.critic, polyak=1)
        c = self.cfg
        self._gradient_steps, self._batch_size = c["gradient_steps"], int(c["batch_size"])
        self._discount_factor, self._polyak = c["discount_factor"], c["polyak"]
        self._actor_learning_rate, self._critic_learning_rate = c["actor_learning_rate"], c["critic_learning_rate"]
        if c["learning_rate_scheduler"] is not None:
            raise L.OisError("learning-rate schedulers are not used by any reference script and are not implemented")
        if c["grad_norm_clip"] > 0:
            raise L.OisError("grad_norm_clip > 0 is not used by any reference script and is not implemented")
        self._random_timesteps, self._learning_starts = c["random_timesteps"], c["learning_starts"]
        self._exploration_noise = c["exploration"]["noise"]
        self._exploration_initial_scale = c["exploration"]["initial_scale"]
        self._exploration_final_scale = c["exploration"]["final_scale"]
        self._exploration_timesteps = c["exploration"]["timesteps"]
        self._rewards_shaper = c["rewards_shaper"]
        self._mixed_precision = bool(c["mixed_precision"])
        self.policy_optimizer = _AdamState(self.policy, self._actor_learning_rate)      # ddpg.py:174-175
        self.critic_optimizer = _AdamState(self.critic, self._critic_learning_rate)
        self.checkpoint_modules["policy_optimizer"] = self.policy_optimizer
        self.checkpoint_modules["critic_optimizer"] = self.critic_optimizer
        sp = c["state_preprocessor"]
        if sp:
            self._state_preprocessor = sp(**c["state_preprocessor_kwargs"]) if isinstance(sp, type) else sp
            if not isinstance(self._state_preprocessor, RunningStandardScaler):
                raise L.OisError("state_preprocessor must be options_in_simulation_b200.agent.RunningStandardScaler")
            self.checkpoint_modules["state_preprocessor"] = self._state_preprocessor
        else:
            self._state_preprocessor = None
        self._grad_actor = torch.zeros_like(self.policy.flat)
        self._grad_critic = torch.zeros_like(self.critic.flat)
        # world_size > 1: the gradient buffers move into this rank's peer-mapped exchange block (csrc/comm.cuh)
        self._exchange = None
        self._dp_exchange = c["dp_exchange"] if self.is_distributed else "none"
        if self._dp_exchange not in ("peer", "nccl", "none"):
            raise L.OisError("dp_exchange must be 'peer' or 'nccl'")
        if self._dp_exchange == "peer":
            self._exchange = P.PeerExchange.create(self.critic.num_params, self.policy.num_params, self.device)
            if self._exchange is None:      # CUDA IPC unavailable on some rank: every rank falls back together
                self._dp_exchange = "nccl"
            else:
                self._grad_critic, self._grad_actor = self._exchange.critic_grad, self._exchange.actor_grad
        self._stats = torch.zeros(8, dtype=torch.float32, device=self.device)
        self._noise_stats = torch.zeros(3, dtype=torch.float32, device=self.device)
        self._handle = None
        self._handle_rows = 0
        self._handle_generation = 0    # bumped when the C handle is re-created: captured graphs hold pointers into it
        self._act_calls = 0
        self._updates = 0
        self._graph = None            # (torch.cuda.CUDAGraph, idx buffer) of one update, see _update_graphed
        self._step_dev = None
        self._seed = int(os.environ.get("OIS_SEED", "42")) + self.rank    # skrl set_seed convention: seed + rank
        self._fused_replay = False
        self._supplied = {}
        self.kernel_launches = 0
        # tracking (base.py:134-178)
        self.tracking_data = collections.defaultdict(list)
        exp = c["experiment"]
        self.write_interval = exp.get("write_interval", 0)
        self.checkpoint_interval = exp.get("checkpoint_interval", 0)
        self.checkpoint_store_separately = exp.get("store_separately", False)
        self.experiment_dir = None
        self.writer = None
        self._cumulative_rewards = None
        self._cumulative_timesteps = None
        self._track_rewards = collections.deque(maxlen=100)
        self._track_timesteps = collections.deque(maxlen=100)
        self.training = False
        for m in (self.policy, self.target_policy, self.critic, self.target_critic):
            m._agent = self

    # ------------------------------------------------------------------ handle / helpers
    def _ensure_handle(self, rows: int):
        if self._handle is not None and rows <= self._handle_rows:
            return
        if self._handle is not None:
            # a captured update graph / rollout.RolloutGraph holds raw pointers into the handle's workspaces and weight
            # images: drop ours, and let RolloutGraph see the new generation and re-capture
            torch.cuda.synchronize(self.device)
            self._graph = None
            L.lib().ois_agent_destroy(self._handle)
        self._handle_generation += 1
        rows = max(int(rows), self._batch_size)
        h = C.c_void_p()
        prec = L.PRECISION_BF16 if self._mixed_precision else L.PRECISION_FP32
        L.check(L.lib().ois_agent_create(C.byref(h), C.byref(self.policy.shape), C.byref(self.critic.shape), rows,
                                         self._batch_size, prec, L.current_stream()))
        self._handle, self._handle_rows = h, rows

    def _weights_changed(self):
        if self._handle is not None:
            L.check(L.lib().ois_agent_invalidate(self._handle))
        self._graph = None    # a captured update assumes the weight images its previous replay left behind

    def __del__(self):
        try:
            if self._handle is not None:
                L.lib().ois_agent_destroy(self._handle)
                self._handle = None
        except Exception:
            pass

    def _mlp_forward(self, model: DeterministicMLP, x):
        self._ensure_handle(x.shape[0])
        if self._mixed_precision:
            raise L.OisError("Model.act is the exact fp32 path; build the agent with mixed_precision=False to use it")
        y = torch.empty(x.shape[0], model.shape.out_dim, dtype=torch.float32, device=self.device)
        L.check(L.lib().ois_mlp_forward(self._handle, C.byref(model.shape), L.ptr(model.flat), L.ptr(x), L.ptr(y),
                                        x.shape[0], L.current_stream()))
        self.kernel_launches += 4
        return y

    def _scaler_ref(self):
        return self._state_preprocessor.c_struct() if self._state_preprocessor is not None else None

    def supply_draws(self, normals=None, uniforms=None):
        """Parity mode: the N(0,1) draws of the OU noise (ddpg.py:239) / the U(0,1) draws of random_act
        (ddpg.py:230) become inputs.  ``None`` returns that stream to the in-kernel Philox generator."""
        f = lambda a: None if a is None else torch.as_tensor(np.asarray(a, dtype=np.float32)).to(self.device).contiguous()
        self._supplied = {"normals": f(normals), "uniforms": f(uniforms)}

    # ------------------------------------------------------------------ skrl Agent surface
    def init(self, trainer_cfg=None):
        """ddpg.py:194-213 (+ base.py:134-178: experiment directory, TensorBoard writer, intervals)."""
        trainer_cfg = trainer_cfg or {}
        timesteps = int(trainer_cfg.get("timesteps", 0))
        if self.write_interval == "auto":
            self.write_interval = int(timesteps / 100)
        if self.checkpoint_interval == "auto":
            self.checkpoint_interval = int(timesteps / 10)
        exp = self.cfg["experiment"]
        if (self.write_interval > 0 or self.checkpoint_interval > 0) and self.rank == 0:
            directory = exp.get("directory") or os.path.join(os.getcwd(), "runs")
            name = exp.get("experiment_name") or "{}_{}".format(
                datetime.datetime.now().strftime("%y-%m-%d_%H-%M-%S-%f"), self.__class__.__name__)
            self.experiment_dir = os.path.join(directory, name)
            if self.checkpoint_interval > 0:
                os.makedirs(os.path.join(self.experiment_dir, "checkpoints"), exist_ok=True)
            if self.write_interval > 0:
                from torch.utils.tensorboard import SummaryWriter

                self.writer = SummaryWriter(log_dir=self.experiment_dir)
        elif self.rank != 0:   # only rank 0 logs / checkpoints (base.py:146-148)
            self.write_interval, self.checkpoint_interval = 0, 0
        self.set_mode("eval")
        if self.memory is not None:
            m = self.memory
            m.create_tensor(name="states", size=self.observation_space, dtype=torch.float32)
            m.create_tensor(name="next_states", size=self.observation_space, dtype=torch.float32)
            m.create_tensor(name="actions", size=self.action_space, dtype=torch.float32)
            m.create_tensor(name="rewards", size=1, dtype=torch.float32)
            m.create_tensor(name="terminated", size=1, dtype=torch.bool)
            m.create_tensor(name="truncated", size=1, dtype=torch.bool)
            self._tensors_names = ["states", "actions", "rewards", "next_states", "terminated", "truncated"]
            if self.cfg["sampler"] == "device":
                m.sampler, m._seed = "device", self._seed
        if self.action_space is not None:
            self.clip_actions_min = torch.tensor(self.action_space.low, device=self.device)
            self.clip_actions_max = torch.tensor(self.action_space.high, device=self.device)
            lo, hi = np.asarray(self.action_space.low), np.asarray(self.action_space.high)
            if not ((lo == lo.flat[0]).all() and (hi == hi.flat[0]).all()):
                raise L.OisError("per-dimension action bounds are not used by the reference and not implemented")
            self._clip = (float(lo.flat[0]), float(hi.flat[0]))
        else:
            self._clip = (-math.inf, math.inf)

    def fuse_replay_with(self, task):
        """Let the env kernel write this agent's memory rows (record_transition then only does bookkeeping)."""
        task.bind_replay(self.memory)
        self._fused_replay = True

    def set_mode(self, mode):
        self.training = mode == "train"
        for m in (self.policy, self.target_policy, self.critic, self.target_critic):
            m.set_mode(mode)

    def set_running_mode(self, mode):
        self.training = mode == "train"

    def track_data(self, tag, value):
        self.tracking_data[tag].append(value)

    def pre_interaction(self, timestep, timesteps):
        pass

    def act(self, states, timestep, timesteps, *, _ctl=None, _out=None, _states_ptr=None):
        """ddpg.py:215-268 as ONE C-ABI call: scaler -> actor MLP -> OU noise -> scale -> clamp
        (or U[low, high] while timestep < random_timesteps).
        ``_ctl`` / ``_out`` are for rollout.RolloutGraph: a device OisStepCtl the kernel takes its per-step scalars
        from, and a fixed output buffer (a captured launch cannot have either change between replays).
        ``_states_ptr``: device-visible address of ``states`` when it is a pinned HOST tensor the kernel reads in place
        (zero-copy over PCIe / C2C; the observations of one step are read exactly once, so a staging copy buys nothing)."""
        if _states_ptr is None:
            states = states.to(self.device, torch.float32).contiguous() if not (states.is_cuda and states.is_contiguous()) else states
        n = states.shape[0]
        self._ensure_handle(n)
        A = self.policy.shape.out_dim
        actions = torch.empty(n, A, dtype=torch.float32, device=self.device) if _out is None else _out
        io = L.ActIO()
        io.ctl = _ctl
        io.states, io.actions = (states.data_ptr() if _states_ptr is None else _states_ptr), actions.data_ptr()
        io.philox_seed, io.philox_offset = self._seed, self._act_calls
        io.clip_low, io.clip_high = self._clip
        io.normals = L.ptr(self._supplied.get("normals"))
        io.uniforms = L.ptr(self._supplied.get("uniforms"))
        random_phase = timestep < self._random_timesteps
        io.random_phase = int(random_phase)
        track = self.write_interval > 0
        noise = self._exploration_noise
        applied = False
        if not random_phase and noise is not None:
            io.ou_state = noise._state_for((n, A)).data_ptr()
            io.theta, io.sigma, io.base_scale = noise.theta, noise.sigma, noise.base_scale
            if self._exploration_timesteps is None:
                self._exploration_timesteps = timesteps
            if timestep <= self._exploration_timesteps:   # ddpg.py:247-251
                io.noise_scale = (1 - timestep / self._exploration_timesteps) * (
                    self._exploration_initial_scale - self._exploration_final_scale) + self._exploration_final_scale
                applied = True
            else:
                io.noise_scale = -1.0
            if track and applied:
                self._noise_stats.copy_(torch.tensor([-math.inf, math.inf, 0.0]))
                io.noise_stats = self._noise_stats.data_ptr()
        L.check(L.lib().ois_agent_act(self._handle, L.ptr(self.policy.flat), self._scaler_ref(), C.byref(io), n,
                                      L.current_stream(self.device.index)))
        self._act_calls += 1
        self.kernel_launches += self._act_launches(random_phase, noise is not None)
        if track and noise is not None and not random_phase:   # ddpg.py:258-266
            if applied:
                mx, mn, sm = self._noise_stats.tolist()
                self.track_data("Exploration / Exploration noise (max)", mx)
                self.track_data("Exploration / Exploration noise (min)", mn)
                self.track_data("Exploration / Exploration noise (mean)", sm / (n * A))
            else:
                for k in ("max", "min", "mean"):
                    self.track_data(f"Exploration / Exploration noise ({k})", 0)
        return actions, None, {}

    def _act_launches(self, random_phase, has_noise):
        if random_phase:
            return 1
        if self._mixed_precision:
            return 1          # scaler, MLP, OU noise and clamp are ONE tensor-core kernel (csrc/mlp_tc.cu)
        return (1 if self._state_preprocessor is not None else 0) + 4 + (1 if has_noise else 0)

    def record_transition(self, states, actions, rewards, next_states, terminated, truncated, infos, timestep, timesteps):
        """ddpg.py:270-329 (+ episode tracking of base.py:271-341 when write_interval > 0)."""
        if self.write_interval > 0:
            if self._cumulative_rewards is None:
                self._cumulative_rewards = torch.zeros_like(rewards, dtype=torch.float32)
                self._cumulative_timesteps = torch.zeros_like(rewards, dtype=torch.int32)
            self._cumulative_rewards.add_(rewards)
            self._cumulative_timesteps.add_(1)
            finished = (terminated + truncated).nonzero(as_tuple=False)
            if finished.numel():
                self._track_rewards.extend(self._cumulative_rewards[finished][:, 0].reshape(-1).tolist())
                self._track_timesteps.extend(self._cumulative_timesteps[finished][:, 0].reshape(-1).tolist())
                self._cumulative_rewards[finished] = 0
                self._cumulative_timesteps[finished] = 0
            self.tracking_data["Reward / Instantaneous reward (max)"].append(torch.max(rewards).item())
            self.tracking_data["Reward / Instantaneous reward (min)"].append(torch.min(rewards).item())
            self.tracking_data["Reward / Instantaneous reward (mean)"].append(torch.mean(rewards).item())
            if len(self._track_rewards):
                tr, tt = np.array(self._track_rewards), np.array(self._track_timesteps)
                self.tracking_data["Reward / Total reward (max)"].append(np.max(tr))
                self.tracking_data["Reward / Total reward (min)"].append(np.min(tr))
                self.tracking_data["Reward / Total reward (mean)"].append(np.mean(tr))
                self.tracking_data["Episode / Total timesteps (max)"].append(np.max(tt))
                self.tracking_data["Episode / Total timesteps (min)"].append(np.min(tt))
                self.tracking_data["Episode / Total timesteps (mean)"].append(np.mean(tt))
        if self.memory is not None:
            if self._rewards_shaper is not None:
                if self._fused_replay:
                    raise L.OisError("rewards_shaper cannot be combined with the fused replay append")
                rewards = self._rewards_shaper(rewards, timestep, timesteps)
            if not self._fused_replay:   # fused: the env kernel already wrote row memory_index and advanced it
                self.memory.add_samples(states=states, actions=actions, rewards=rewards, next_states=next_states,
                                        terminated=terminated, truncated=truncated)
                self.kernel_launches += 1
            for memory in self.secondary_memories:
                memory.add_samples(states=states, actions=actions, rewards=rewards, next_states=next_states,
                                   terminated=terminated, truncated=truncated)

    def post_interaction(self, timestep, timesteps):
        """ddpg.py:341-355 then base.py:645-669 (tracking data / checkpoints)."""
        if timestep >= self._learning_starts:
            self.set_mode("train")
            self._update(timestep, timesteps)
            self.set_mode("eval")
        self.base_post_interaction(timestep, timesteps)

    def base_post_interaction(self, timestep, timesteps):
        """skrl Agent.post_interaction (base.py:645-669), which single_agent_eval calls directly
        (base_KA_DDPG.py:334) to skip ``_update``."""
        timestep += 1
        if self.checkpoint_interval > 0 and not timestep % self.checkpoint_interval:
            self.write_checkpoint(timestep, timesteps)
        if self.write_interval > 0 and not timestep % self.write_interval:
            self.write_tracking_data(timestep, timesteps)

    def _update_io(self, batch):
        io = L.UpdateIO()
        s, a, r, ns, te, tr = batch
        io.states, io.actions, io.rewards, io.next_states = s.data_ptr(), a.data_ptr(), r.data_ptr(), ns.data_ptr()
        io.terminated, io.truncated = te.data_ptr(), tr.data_ptr()
        io.actor, io.actor_target = self.policy.flat.data_ptr(), self.target_policy.flat.data_ptr()
        io.critic, io.critic_target = self.critic.flat.data_ptr(), self.target_critic.flat.data_ptr()
        io.actor_m, io.actor_v = self.policy_optimizer.m.data_ptr(), self.policy_optimizer.v.data_ptr()
        io.critic_m, io.critic_v = self.critic_optimizer.m.data_ptr(), self.critic_optimizer.v.data_ptr()
        io.actor_lr, io.critic_lr = self.policy_optimizer.lr, self.critic_optimizer.lr
        io.beta1, io.beta2 = self.policy_optimizer.betas
        io.adam_eps = self.policy_optimizer.eps
        io.discount, io.polyak = self._discount_factor, self._polyak
        io.actor_grad, io.critic_grad = self._grad_actor.data_ptr(), self._grad_critic.data_ptr()
        io.stats = self._stats.data_ptr()
        io.comm = self._exchange.handle.value if (self._exchange is not None and self._dp_exchange == "peer") else None
        io.batch_rows = int(s.shape[0])
        return io

    def _reduce(self, model, grad):
        """The gradient reduction between two phases (ddpg.py:409-410,431-432).  "peer": nothing to do here — the next
        phase's Adam kernel exchanges the gradients itself; "nccl": skrl's all-reduce."""
        if self._dp_exchange == "nccl":
            model.reduce_parameters(grad)

    def _track_update_stats(self):
        every = self.cfg["track_stats_interval"] or (1 if self.write_interval > 0 else 0)
        if every and self._updates % every == 0:   # ddpg.py:452-461
            st = self._stats.tolist()
            self.track_data("Loss / Policy loss", st[1])
            self.track_data("Loss / Critic loss", st[0])
            self.track_data("Q-network / Q1 (max)", st[2])
            self.track_data("Q-network / Q1 (min)", st[3])
            self.track_data("Q-network / Q1 (mean)", st[4])
            self.track_data("Target / Target (max)", st[5])
            self.track_data("Target / Target (min)", st[6])
            self.track_data("Target / Target (mean)", st[7])

    def _update_graphed(self):
        """One update as ONE CUDA-graph launch and nothing else: the first node draws the batch indices and gathers the
        rows (ois_replay_sample_gather), the rest are the three phases.  What changes from update to update — memory
        length, draw counter, Adam step count — lives in a device-resident OisSampleCtl that the first node advances;
        the host re-seeds it (one single-thread launch) only when its mirrors moved for another reason."""
        lib, m = L.lib(), self.memory
        if m.sampler != "device" or m._supplied_indexes:
            raise L.OisError("graph_update needs sampler='device' and no supplied indexes")
        if len(m) < self._batch_size:
            raise L.OisError("graph_update needs len(memory) >= batch_size (raise learning_starts)")
        scale = 1.0 / self.world_size if self.is_distributed else 1.0
        grads = (self._grad_critic, self._grad_actor, None)
        nets = (self.critic, self.policy, None)
        if self._graph is None:
            idx = torch.empty(self._batch_size, dtype=torch.int64, device=self.device)
            # OisSampleCtl (len, draws, adam_step, ticket) on the device: the sampler / gather launch inside the graph
            # reads it and advances it, so a steady-state update is ONE graph launch and nothing else
            self._sample_ctl = torch.zeros(4, dtype=torch.int64, device=self.device)
            self._push_sample_ctl()
            batch = m.sample_gather(self._tensors_names, idx, self._sample_ctl)[0]   # allocates the batch buffers once
            io = self._update_io(batch)
            io.adam_step_dev = self._sample_ctl.data_ptr() + 16
            # this call's update runs eagerly (it also leaves the weight images in their steady state); the same
            # sequence is then captured — capture records, it does not execute — for every later call
            for ph in range(3):
                L.check(lib.ois_ddpg_update_phase(self._handle, self._scaler_ref(), C.byref(io), ph, scale, L.current_stream()))
                if ph < 2:
                    self._reduce(nets[ph], grads[ph])
            torch.cuda.synchronize()
            graph = torch.cuda.CUDAGraph()
            draws, launches = m._draws, m.kernel_launches
            # "thread_local": the NCCL watchdog thread may query events while this thread captures (dp_exchange "nccl")
            with torch.cuda.graph(graph, capture_error_mode="thread_local"):
                m.sample_gather(self._tensors_names, idx, self._sample_ctl)
                for ph in range(3):
                    L.check(lib.ois_ddpg_update_phase(self._handle, self._scaler_ref(), C.byref(io), ph, scale, L.current_stream()))
                    if ph < 2:
                        self._reduce(nets[ph], grads[ph])
            m._draws, m.kernel_launches = draws, launches      # the capture recorded a launch, it did not draw
            self._graph = (graph, idx, io, batch)
        else:
            graph, idx = self._graph[0], self._graph[1]
            if (len(m), m._draws, self.critic_optimizer.step) != self._sample_ctl_mirror:
                self._push_sample_ctl()      # the memory grew, or somebody else drew / stepped since the last replay
            graph.replay()
            m._draws += 1
            m.kernel_launches += 1
            m.sampling_indexes = idx
        self._sample_ctl_mirror = (len(m), m._draws, self.critic_optimizer.step + 1)
        n = 1
        self.policy_optimizer.step += n
        self.critic_optimizer.step += n
        self._updates += n
        self.kernel_launches += n * (UPDATE_LAUNCHES_BF16 if self._mixed_precision else UPDATE_LAUNCHES)   # + the memory's one
        # host-side "weights changed" flags of the library are not touched by a replay: the actor's inference image
        # must be rebuilt by the next act()
        L.check(lib.ois_agent_invalidate(self._handle))
        return n

    def _push_sample_ctl(self):
        """Device OisSampleCtl := the host mirrors (memory length, draw counter, optimizer steps completed)."""
        m = self.memory
        L.check(L.lib().ois_sample_ctl_set(self._sample_ctl.data_ptr(), len(m), m._draws, self.critic_optimizer.step,
                                           L.current_stream()))
        self.kernel_launches += 1

    def _update(self, timestep, timesteps):
        """ddpg.py:357-465: three C-ABI phases split where skrl all-reduces gradients."""
        self._ensure_handle(self._batch_size)
        lib = L.lib()
        if self.cfg["graph_update"]:
            done = 0
            while done < self._gradient_steps:
                done += self._update_graphed()
                self._track_update_stats()
            return
        for _ in range(self._gradient_steps):
            # the reference trains on whatever randperm(len)[:batch_size] returns, i.e. on a SHORTER batch while the
            # memory holds fewer than batch_size transitions (io.batch_rows tells the kernels; fp32 path only)
            batch = self.memory.sample(names=self._tensors_names, batch_size=self._batch_size)[0]
            if batch[0].shape[0] < 2:
                continue     # a 1-row batch has no variance (torch.var -> NaN in the reference); nothing sane to train on
            io = self._update_io(batch)
            self.policy_optimizer.step += 1
            self.critic_optimizer.step += 1
            io.adam_step = self.critic_optimizer.step
            scale = 1.0 / self.world_size if self.is_distributed else 1.0
            stream = L.current_stream()
            L.check(lib.ois_ddpg_update_phase(self._handle, self._scaler_ref(), C.byref(io), 0, scale, stream))
            self._reduce(self.critic, self._grad_critic)    # critic.reduce_parameters(), ddpg.py:409-410
            L.check(lib.ois_ddpg_update_phase(self._handle, self._scaler_ref(), C.byref(io), 1, scale, stream))
            self._reduce(self.policy, self._grad_actor)     # policy.reduce_parameters(), ddpg.py:431-432
            L.check(lib.ois_ddpg_update_phase(self._handle, self._scaler_ref(), C.byref(io), 2, scale, stream))
            self._updates += 1
            self.kernel_launches += UPDATE_LAUNCHES_BF16 if self._mixed_precision else UPDATE_LAUNCHES
            self._track_update_stats()

    # ------------------------------------------------------------------ tracking / checkpoints (base.py:193-249,364-395)
    def write_tracking_data(self, timestep, timesteps):
        if self.writer is not None:
            for k, v in self.tracking_data.items():
                if k.endswith("(min)"):
                    self.writer.add_scalar(k, np.min(v), timestep)
                elif k.endswith("(max)"):
                    self.writer.add_scalar(k, np.max(v), timestep)
                else:
                    self.writer.add_scalar(k, np.mean(v), timestep)
        self._track_rewards.clear()
        self._track_timesteps.clear()
        self.tracking_data.clear()

    def _modules_state(self):
        return {name: m.state_dict() for name, m in self.checkpoint_modules.items()}

    def write_checkpoint(self, timestep, timesteps):
        tag = str(timestep if timestep is not None else datetime.datetime.now().strftime("%y-%m-%d_%H-%M-%S-%f"))
        d = os.path.join(self.experiment_dir, "checkpoints")
        if self.checkpoint_store_separately:
            for name, sd in self._modules_state().items():
                torch.save(sd, os.path.join(d, f"{name}_{tag}.pt"))
        else:
            torch.save(self._modules_state(), os.path.join(d, f"agent_{tag}.pt"))

    def save(self, path):
        torch.save(self._modules_state(), path)

    def load(self, path):
        """Load an ``agent_{t}.pt`` written by this class or by the reference (dict of state_dicts)."""
        modules = torch.load(path, map_location="cpu", weights_only=False)
        if not isinstance(modules, dict):
            raise L.OisError("unsupported checkpoint format")
        for name, data in modules.items():
            module = self.checkpoint_modules.get(name)
            if module is not None and data is not None:
                module.load_state_dict(data)
        self._weights_changed()


# kernels per _update on the fp32 path: 2 scaler + (4 fwd + 2 copy) + (4 fwd) + td + 2 copy + 4 fwd + loss +
# critic bwd (4 wgrad + 4 colsum + 3 dgrad) | adam + 4 fwd + 2 copy + 4 fwd + loss + critic dgrad 4 + out-grad +
# actor bwd 11 | adam + 2 polyak
UPDATE_LAUNCHES = (2 + 4 + 2 + 4 + 1 + 2 + 4 + 1 + 11) + (1 + 4 + 2 + 4 + 1 + 4 + 1 + 11) + 3
# on the bf16 tcgen05 path (csrc/update_tc.cu; steady state, weight images kept current by Adam / polyak):
# batch prep + 4 grouped fwd (target actor || critic) + 4 grouped fwd (target critic || actor) + TD / loss / first data
# gradient + critic bwd (2 dgrad + ONE grouped launch of the weight and bias gradients) | adam + policy pass through the
# critic (3 fwd, layer 3 emits the first data gradient; [layer 4 || dgrad]; dgrad; dx) + sigma' / actor's first data
# gradient / policy-loss statistics + actor bwd 3 | adam(actor) + both polyaks in ONE launch
UPDATE_LAUNCHES_BF16 = (1 + 4 + 4 + 1 + 3) + (1 + 6 + 1 + 3) + 1
