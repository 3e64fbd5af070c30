"""``RolloutGraph`` — one policy step of the trainer's loop body replayed as ONE CUDA-graph launch.

The step is what src/my_trainers/base_KA_DDPG.py:175-219 does per timestep::

    actions = agent.act(states, t, T)                       src/my_agents/ddpg.py:215-268
    next_states, rewards, terminated, ... = env.step(actions)   src/my_envs/KAMMA_medium.py:596-727
    agent.record_transition(...)                             src/my_agents/ddpg.py:270-329 (fused into the env kernel)

Issued call by call, that is two C-ABI calls plus (with a host in the loop) four torch copies per step: at 4096 envs
the GPU work is ~30 us and the host issue cost is of the same order.  Captured once, the step is one
``cudaGraphLaunch``.  A captured launch has frozen arguments, so everything that changes from step to step (Philox
offsets, the decayed exploration scale, the replay ring slot, the eval flag slot) lives in a device-resident
``OisStepCtl`` block (include/ois_b200.h) that the last CTA of the env kernel to finish advances to the next step.
The host mirrors (``agent._act_calls``, ``task._launches``, ``memory.memory_index`` / ``filled``) are advanced
arithmetically per replay, so eager calls and graph replays can be interleaved and produce the same stream of results
as eager calls alone (tests/test_rollout_graph_gpu.py checks bit-equality).

Limits: the agent must be past ``random_timesteps`` (that branch launches a different kernel), noise statistics
tracking and the strided env statistics are not gathered on replayed steps (``write_interval`` / ``stats_interval``
reads need a host round trip per step anyway), supplied parity draws are not supported.
"""
from __future__ import annotations

import ctypes as C

import torch

from . import _lib as L


def _device_synchronize():
    """cudaDeviceSynchronize without the Python layers of torch.cuda.synchronize (this call is on the per-step path)."""
    fn = getattr(torch._C, "_cuda_synchronize", None)
    if fn is None:
        torch.cuda.synchronize()
    else:
        fn()


class RolloutGraph:
    """``host_io=False``: device-resident loop — ``step()`` returns the (fixed) device tensors
    ``(next_states, rewards[N,1], terminated[N,1])`` without synchronising.
    ``host_io=True``: the observations of each step are read from the pinned host tensor ``h_states`` and the step's
    results are returned in pinned host tensors once the step is complete — the host-in-the-loop shape ``bench.py``
    times as ``e2e``.  The two kernels access the pinned buffers in place (zero-copy: the actor kernel loads its rows of
    ``h_states`` over PCIe, the env kernel stores obs | rew | reset straight to host memory), so the graph is two kernel
    nodes; every byte still crosses the bus every step, but without the copy-engine round trips.  There are two pinned
    result packs and two captured graphs that alternate: step t reads its observations from the pack step t-1 wrote
    (``states = next_states``, base_KA_DDPG.py:268-275, without a host copy — the caller may still overwrite
    ``h_states`` before ``step()``) and writes the other one, so the results of step t stay readable while step t+1
    runs.  Completion is signalled by the env kernel itself (``OisEnvIO.done_flag``: a pinned word the last CTA
    stores the step count to) and ``step()`` is ONE C call, ``ois_graph_launch_wait``: no cudaDeviceSynchronize.
    ``step(wait=False)`` only launches; ``wait()`` then blocks until that step's results are in host memory (one step
    may be in flight while the host consumes the previous one).
    ``host_io="copy"`` keeps explicit copy nodes instead (H2D of the observations, ONE packed D2H) and a device
    synchronise per step."""

    def __init__(self, env, agent, timestep: int, timesteps: int, host_io: bool = False, fuse: bool = True):
        task = env.task
        if not agent._fused_replay or task._replay is None:
            raise L.OisError("RolloutGraph needs agent.fuse_replay_with(task) (the append is part of the captured kernel)")
        if timestep < agent._random_timesteps:
            raise L.OisError("RolloutGraph starts after random_timesteps (that phase launches a different kernel)")
        if agent._supplied or task._supplied_mask is not None or task._supplied_reset is not None:
            raise L.OisError("RolloutGraph does not take supplied parity draws")
        if agent.write_interval > 0:
            raise L.OisError("RolloutGraph does not gather the per-step noise statistics (write_interval must be 0)")
        self.env, self.task, self.agent, self.memory = env, task, agent, task._replay
        self.timestep, self.timesteps = int(timestep), int(timesteps)
        self.host_io = host_io
        self.fuse = bool(fuse)           # actor + env step as one kernel where the library can (csrc/mlp_tct.cu, kEnv)
        self.kernel_launches = 0
        dev, n, A = task.device, task.num_envs, task.num_actions
        self._dev_index = dev.index
        self._ctl = torch.zeros(C.sizeof(L.StepCtl), dtype=torch.uint8, device=dev)
        self._actions = torch.empty(n, A, dtype=torch.float32, device=dev)
        # the state the captured actor launch reads: the env's own clamped-observation buffer (device loop) or the
        # landing buffer of the host->device copy (host loop)
        self.states = task._obs_clipped[0]
        self._zero_copy = host_io is True
        self._steps_done = 0             # steps issued through the step-control block (the value done_flag receives)
        if host_io:
            # obs [N,6] f32 | rew [N] f32 | terminated [N] u8: the task keeps them in ONE allocation (envs.py: _out_pack)
            packs = 2 if self._zero_copy else 1
            self._h_packs = [torch.zeros(n * 29, dtype=torch.uint8).pin_memory() for _ in range(packs)]
            view = lambda p: (p[: n * 24].view(torch.float32).view(n, 6), p[n * 24: n * 28].view(torch.float32).view(n, 1),
                              p[n * 28:].view(torch.bool).view(n, 1))
            self._h_views = [view(p) for p in self._h_packs]
            self._cur = 0                # zero-copy: the pack whose observations the NEXT step reads
            if self._zero_copy:
                self._h_done = torch.zeros(16, dtype=torch.int32).pin_memory()     # [0]: steps completed (done_flag)
            else:
                self._h_in = torch.empty(n, 6, dtype=torch.float32).pin_memory()
                self._d_states = torch.empty(n, 6, dtype=torch.float32, device=dev)
            self.h_next_states, self.h_rewards, self.h_terminated = self._h_views[0]
        self._graphs = {}                # captured step per input pack (zero-copy: 0 and 1; otherwise only 0)
        self._graph_generation = -1      # agent._handle_generation the captured launches point into
        self._warm = False
        self._per_step = (1, 1)
        self._wait_timeout_ns = 30_000_000_000

    @property
    def h_states(self):
        """Pinned host buffer the next ``step()`` reads its observations from (fill or overwrite it before the call)."""
        return self._h_views[self._cur][0] if self._zero_copy else self._h_in

    # ------------------------------------------------------------------ the per-step scalars
    def _write_ctl(self):
        """Host -> device copy of the block for the NEXT step, from the host mirrors (so eager steps may have run)."""
        a, t, m = self.agent, self.task, self.memory
        c = L.StepCtl()
        c.step, c.self_advance, c.ticket = self._steps_done, 1, 0
        c.env_philox_offset, c.act_philox_offset = t._launches, a._act_calls
        c.replay_row, c.replay_rows = m.memory_index, m.memory_size
        c.flag_slot = t._launches % 3
        if a._exploration_timesteps is None:
            a._exploration_timesteps = self.timesteps      # ddpg.py:244-245
        T_e = a._exploration_timesteps
        c.timestep, c.expl_timesteps = float(self.timestep), float(T_e)
        c.expl_initial_scale, c.expl_final_scale = a._exploration_initial_scale, a._exploration_final_scale
        c.noise_scale = ((1 - self.timestep / T_e) * (a._exploration_initial_scale - a._exploration_final_scale)
                         + a._exploration_final_scale) if self.timestep <= T_e else -1.0
        host = torch.frombuffer(bytearray(bytes(c)), dtype=torch.uint8)
        self._ctl.copy_(host)
        self._mirror = (t._launches, a._act_calls, m.memory_index, self.timestep)

    def _mirrors_moved(self):
        t, a, m = self.task, self.agent, self.memory
        return self._mirror != (t._launches, a._act_calls, m.memory_index, self.timestep)

    def _issue(self, src: int = 0):
        """The launches of one step, on the current stream (eagerly once to warm up, then under capture).  ``src``: the
        pinned pack the observations are read from (zero-copy mode; the results go to the other one)."""
        ctl = self._ctl.data_ptr()
        n = self.task.num_envs
        if self.fuse:
            # actor + env step as ONE kernel where the library can serve it (ois_rollout_fuse_begin: the act call is
            # recorded, the env step that consumes its actions launches both); otherwise the same two launches as before
            L.check(L.lib().ois_rollout_fuse_begin())
        try:
            self._issue_calls(src, ctl, n)
        finally:
            if self.fuse:
                L.check(L.lib().ois_rollout_fuse_end())

    def _issue_calls(self, src, ctl, n):
        if self.host_io == "copy":
            self._d_states.copy_(self._h_in, non_blocking=True)
            self.agent.act(self._d_states, self.timestep, self.timesteps, _ctl=ctl, _out=self._actions)
            self.task.step(self._actions, True, _ctl=ctl)
            self._h_packs[0].copy_(self.task._out_pack, non_blocking=True)
        elif self.host_io:
            h_in = self._h_views[src][0]                 # pinned memory: the same address is valid on the device (UVA)
            hp = self._h_packs[1 - src].data_ptr()
            self.agent.act(h_in, self.timestep, self.timesteps, _ctl=ctl, _out=self._actions, _states_ptr=h_in.data_ptr())
            self.task.step(self._actions, True, _ctl=ctl, _host_out=(hp, hp + n * 24, hp + n * 28),
                           _done_flag=self._h_done.data_ptr())
        else:
            self.agent.act(self.states, self.timestep, self.timesteps, _ctl=ctl, _out=self._actions)
            self.task.step(self._actions, True, _ctl=ctl)

    def _host_state(self):
        t, a, m = self.task, self.agent, self.memory
        return (t._launches, t.kernel_launches, a._act_calls, a.kernel_launches, m.memory_index, m.filled)

    def _set_host_state(self, s):
        t, a, m = self.task, self.agent, self.memory
        t._launches, t.kernel_launches, a._act_calls, a.kernel_launches, m.memory_index, m.filled = s

    def _capture(self, src: int = 0):
        saved = self._host_state()            # capture runs the Python bookkeeping of one step without executing it
        graph = torch.cuda.CUDAGraph()
        # "relaxed": Python's collector may run a finaliser while the step is being captured (an old agent handle or
        # graph going away -> cudaFree), which the stricter modes treat as an error; nothing captured here depends on
        # memory that such a call could touch (every buffer is held by this object, the task, the agent or the memory)
        with torch.cuda.graph(graph, capture_error_mode="relaxed"):
            self._issue(src)
        self._set_host_state(saved)
        raw = None
        if self._zero_copy:
            try:                      # the instantiated cudaGraphExec_t, for the one-call launch + wait in C
                raw = int(graph.raw_cuda_graph_exec())
            except Exception:         # a torch without the accessor: torch launches, ois_host_wait still waits
                raw = None
        self._graphs[src] = (graph, raw)
        self._graph_generation = self.agent._handle_generation

    def _advance_host(self):
        """What one eager act + step would have done to the host mirrors."""
        t, a, m = self.task, self.agent, self.memory
        t._launches += 1
        t.kernel_launches += self._per_step[0]
        a._act_calls += 1
        a.kernel_launches += self._per_step[1]
        m.advance()
        t.obs_clipped = t._obs_clipped[0]
        self.env._obs = t.obs_clipped
        self.timestep += 1
        self._mirror = (t._launches, a._act_calls, m.memory_index, self.timestep)

    # ------------------------------------------------------------------ public
    def wait(self, lag: int = 0):
        """Zero-copy mode: block (spinning on the pinned completion word, no driver call) until all but the last ``lag``
        launched steps are complete; with lag 0 returns the last step's pinned result tensors."""
        L.check(L.lib().ois_host_wait(self._h_done.data_ptr(), (self._steps_done - lag) & 0xFFFFFFFF, self._wait_timeout_ns))
        return self.h_next_states, self.h_rewards, self.h_terminated

    def step(self, wait: bool = True):
        """One policy step.  The first call issues the launches eagerly (lazy initialisation inside the library must
        not happen under capture), the next ones capture, every later call is one graph launch."""
        a, t = self.agent, self.task
        if not self.host_io and t.obs_clipped.data_ptr() != self.states.data_ptr():
            self.states.copy_(t.obs_clipped)          # an eager step left the current observation in the other buffer
        if not self._warm or self._mirrors_moved():
            self._write_ctl()                          # (re)seed the device block from the host mirrors
        src = self._cur if self._zero_copy else 0
        launched_and_waited = False
        if not self._warm:
            before = self._host_state()
            fused0 = L.lib().ois_rollout_fused_launches()
            self._issue(src)
            after = self._host_state()
            fused = L.lib().ois_rollout_fused_launches() - fused0        # act + env step went out as one kernel
            self._per_step = (after[1] - before[1] - fused, after[3] - before[3])
            self._set_host_state(before)
            self._warm = True
        else:
            # the bf16 image of the actor weights follows DDPG._update; a no-op when nothing changed
            stream = L.current_stream(self._dev_index)
            L.check(L.lib().ois_agent_sync_weights(a._handle, L.ptr(a.policy.flat), self.task.num_envs, stream))
            if self._graph_generation != a._handle_generation:
                self._graphs = {}      # the agent re-created its C handle: the old graphs' pointers are dead
            if src not in self._graphs:
                self._capture(src)
            graph, raw = self._graphs[src]
            if raw is None:
                graph.replay()
            elif wait:                 # launch + spin on the completion word: ONE C call, no device synchronise
                L.check(L.lib().ois_graph_launch_wait(raw, stream, self._h_done.data_ptr(), (self._steps_done + 1) & 0xFFFFFFFF,
                                                      self._wait_timeout_ns))
                launched_and_waited = True
            else:
                L.check(L.lib().ois_graph_launch(raw, stream))
        self._advance_host()
        self._steps_done += 1
        if self._zero_copy:
            self._cur = 1 - src
            self.h_next_states, self.h_rewards, self.h_terminated = self._h_views[self._cur]
            if wait and not launched_and_waited:
                self.wait()
            return self.h_next_states, self.h_rewards, self.h_terminated
        if self.host_io:
            _device_synchronize()     # copy-node mode: one device synchronise per step
            return self.h_next_states, self.h_rewards, self.h_terminated
        return t.obs_clipped, t.rew_buf.view(-1, 1), t.reset_buf.view(-1, 1)


def fused_step(env, agent, states, timestep: int, timesteps: int):
    """One policy step of the trainer's loop body — ``agent.act`` -> ``env.step`` -> ``agent.record_transition``
    (src/my_trainers/base_KA_DDPG.py:175-219) — issued eagerly as ONE kernel launch where the library can serve it
    (bf16 agent, at most 32 x SMs envs, fused replay append, past ``random_timesteps``, ``write_interval`` 0), and as the
    usual two launches otherwise.  Returns what ``env.step`` returns.  Results are bit-identical to the three separate
    calls (tests/test_rollout_graph_gpu.py)."""
    fuse = agent.write_interval <= 0 and agent._fused_replay
    lib = L.lib()
    before = lib.ois_rollout_fused_launches()
    if fuse:
        L.check(lib.ois_rollout_fuse_begin())
    try:
        actions = agent.act(states, timestep, timesteps)[0]
        out = env.step(actions)
    finally:
        if fuse:
            L.check(lib.ois_rollout_fuse_end())
    env.task.kernel_launches -= lib.ois_rollout_fused_launches() - before     # act + env step were one launch
    next_states, rewards, terminated, truncated, _ = out
    agent.record_transition(None, None, rewards, next_states, terminated, truncated, None, timestep, timesteps)
    return out
