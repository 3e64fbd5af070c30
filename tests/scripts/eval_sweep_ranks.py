"""Run under torchrun on >= 2 GPUs (tests/test_multi_gpu.py spawns it): BASELINE.json configs[4] in miniature — every rank
rolls its own shard of KAMMA_eval envs (8192 per rank, as 65 536 over 8 GPUs), no data-path collective; the per-speed-bin
tracking-error sums and episode counts are all-reduced ONCE at the end (evaluate.tracking_error_sweep), so every rank
holds the same curve and the episode total is the sum of the ranks' own."""
import os
import sys

import torch
import torch.distributed as dist

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
import options_in_simulation_b200 as ois
from options_in_simulation_b200 import agent as A
from options_in_simulation_b200.config import script_cfg
from options_in_simulation_b200.evaluate import phase_lock, tracking_error_sweep


def main():
    rank, local, world = int(os.environ["RANK"]), int(os.environ["LOCAL_RANK"]), int(os.environ["WORLD_SIZE"])
    torch.cuda.set_device(local)
    dev = f"cuda:{local}"
    dist.init_process_group("nccl", device_id=torch.device(dev))
    n, steps = 8192, 210
    task = ois.TorqueDistributionTask(sim_config=script_cfg(n, sim_device=dev), variant="KAMMA_eval", seed=42 + rank)
    env = ois.wrap_env(task)
    env.reset()
    period = phase_lock(task)      # spread the envs over the speed ramp (KAMMA_eval.py:545-550), one speed per phase group
    osp, asp = task.observation_space, task.action_space
    torch.manual_seed(3)                               # same policy on every rank (an eval of ONE checkpoint)
    models = {"policy": A.DeterministicActor(osp, asp, dev), "target_policy": A.DeterministicActor(osp, asp, dev),
              "critic": A.Critic(osp, asp, dev), "target_critic": A.Critic(osp, asp, dev)}
    memory = ois.RandomMemory(memory_size=2, num_envs=n, device=dev)
    cfg = {"batch_size": 32, "mixed_precision": True, "state_preprocessor": A.RunningStandardScaler,
           "state_preprocessor_kwargs": {"size": osp, "device": dev}, "exploration": {"noise": None},
           "experiment": {"write_interval": 0, "checkpoint_interval": 0}}
    agent = A.DDPG(models=models, memory=memory, cfg=cfg, observation_space=osp, action_space=asp, device=dev)
    before = agent.policy.flat.clone()
    out = tracking_error_sweep(env, agent, steps, skip_steps=period)
    torch.cuda.synchronize()
    assert torch.equal(before, agent.policy.flat), "eval must not update the policy"
    local_eps = out["series"][:, 3].sum().to(dev)
    total = local_eps.clone()
    dist.all_reduce(total)
    assert float(out["episodes"].sum()) == float(total), (float(out["episodes"].sum()), float(total))
    assert float(local_eps) >= n - 100, float(local_eps)             # one full episode per env after the skipped ones
    curve = torch.stack([out["tracking_error"], out["episodes"]]).to(dev)
    gathered = [torch.empty_like(curve) for _ in range(world)]
    dist.all_gather(gathered, curve)
    assert all(torch.equal(gathered[0], g) for g in gathered), "ranks disagree on the reduced curve"
    assert int((out["episodes"] > 0).sum()) >= 8 and bool(torch.isfinite(out["tracking_error"]).all())
    if rank == 0:
        print("eval_sweep_ranks ok", [round(float(x), 4) for x in out["tracking_error"]], flush=True)
    dist.barrier()
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
