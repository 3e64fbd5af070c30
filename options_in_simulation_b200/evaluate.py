"""Tracking-error-vs-speed sweep: the eval rollout behind BASELINE.json configs[4] (KAMMA-4d eval, 65 536 envs over 8 GPUs).

What the reference does (single env, TensorBoard in the loop): ``single_agent_eval`` (src/my_trainers/base_KA_DDPG.py:295-350)
steps ``KAMMA_eval`` whose commanded speed ramps from 1 to 2 m/s with the env's own step count
(src/my_envs/KAMMA_eval.py:545-550); each time an episode ends, ``reset_idx`` logs the episode means
``rew_Tracking error`` (mean |v_delta|) and ``rew_Desired velocity`` (src/my_envs/KAMMA_eval.py:224-226,488-493,681-683,840);
``options-sim-eval-artifacts`` then pairs the two TensorBoard series step by step and plots TE against speed
(src/options_sim/cli/eval_artifacts.py:65-174).

Here every rank rolls its shard of envs through ``trainer.SequentialTrainer.eval``; the fused env kernel reduces the
statistics of the episodes that end at step k into row k of a device-resident log (``task.stats_log``), so the loop has
no host synchronisation; at the end the log is read once, turned into the same (Desired velocity, Tracking error)
pairs the reference logs, binned by speed (``parallel.TrackingErrorBins``), and the bins are summed over ranks with ONE
all-reduce.  Each pair is weighted by the number of episodes behind it (the reference has one env, so weight 1).
"""
from __future__ import annotations

import torch

from . import _lib as L
from . import parallel as P

# rows of OisEnvIO.ep_stats (include/ois_b200.h): episode sums [0..10] in envs.SUM_KEYS + EVAL_SUM_KEYS order, count [18]
_I_TE, _I_DV, _I_SMOOTH, _I_COUNT = 9, 10, 8, 18


def phase_lock(task, ramp_steps: float | None = None) -> int:
    """Arrange the envs so that one vectorised rollout IS the speed sweep.  The reference sweeps by running one env for
    50 000 steps; its logged pair is a mean over the envs whose episode ended at that step, so envs that end together
    must share a commanded speed.  Env i gets phase = i mod P (P = episode length in steps): its step counter is placed
    at phase / P of the ramp, its command is re-sampled there (reset_idx) and its episode clock is advanced by `phase`,
    so the group of phase p ends its episodes at steps P - p, 2P - p, ... and commands the speed of ramp position p / P.
    Returns P; the first P steps of the sweep hold partial episodes (pass ``skip_steps=P``)."""
    n = task.num_envs
    period = int(task.max_episode_length) - 1                      # timeout = episode_buf >= max_episode_length - 1
    ramp = float(task.variant.max_sim_steps if ramp_steps is None else ramp_steps)
    phase = torch.arange(n, device=task.device, dtype=torch.int64) % period
    task.sim_steps.copy_((phase.double() * (ramp / period)).long())
    task.reset_idx(torch.arange(n, device=task.device, dtype=torch.int64))
    task.episode_buf.copy_(phase)
    return period


def tracking_error_sweep(env, agent, timesteps: int, lo: float = 1.0, hi: float = 2.0, bins: int = 10,
                         pre_step_hook=None, skip_steps: int = 0):
    """Run ``timesteps`` eval steps and return a dict:

    ``series``     [k, 4] float64 (cpu): step, Desired velocity, Tracking error, episodes ended — one row per step at which
                   episodes ended, i.e. the two TensorBoard series of the reference aligned by step
    ``speed``      [bins] bin centres, ``tracking_error`` [bins] episode-weighted mean TE per bin, ``episodes`` [bins]
                   (summed over ranks when torch.distributed is initialised)
    """
    from .trainer import SequentialTrainer

    task = env.task
    if not task.variant.eval_sums:
        raise L.OisError("tracking_error_sweep needs an eval variant (KAMMA_eval): the sums it bins are eval-only")
    # row k = env step k; when the wrapper has not been reset yet, step 0 is its reset step (zero actions), hence + 1
    log = torch.zeros(int(timesteps) + 1, L.OIS_EP_STATS_LEN, dtype=torch.float32, device=task.device)
    task.stats_log, task._stats_row = log, 0
    try:
        trainer = SequentialTrainer(env=env, agents=agent, cfg={
            "timesteps": int(timesteps), "headless": True, "disable_progressbar": True, "log_interval": 0,
            "pre_step_hook": pre_step_hook})
        trainer.eval()
    finally:
        task.stats_log = None
    st = log.double()                                    # ONE device->host read for the whole sweep (below)
    count = st[:, _I_COUNT].clone()
    count[: int(skip_steps)] = 0                          # partial first episodes (see phase_lock)
    T = float(task.max_episode_length_s)
    safe = count.clamp_min(1.0)
    dv, te = st[:, _I_DV] / safe / T, st[:, _I_TE] / safe / T      # KAMMA_eval.py:488-493: mean over ended envs / T
    acc = P.TrackingErrorBins(lo, hi, bins, device=task.device)
    ended = count > 0
    acc.te_sum.index_add_(0, _bin_index(dv, lo, hi, bins), te * count)     # weight = episodes behind the pair
    acc.count.index_add_(0, _bin_index(dv, lo, hi, bins), count)
    acc.reduce()
    steps = torch.nonzero(ended).flatten()
    series = torch.stack([steps.double(), dv[steps], te[steps], count[steps]], dim=1).cpu()
    width = (hi - lo) / bins
    return {"series": series,
            "speed": torch.tensor([lo + (b + 0.5) * width for b in range(bins)], dtype=torch.float64),
            "tracking_error": acc.means().cpu(), "episodes": acc.count.cpu(),
            "smoothness": (st[:, _I_SMOOTH] / safe / T)[steps].cpu()}


def _bin_index(x, lo, hi, bins):
    return ((x - lo) / (hi - lo) * bins).floor().clamp_(0, bins - 1).long()
