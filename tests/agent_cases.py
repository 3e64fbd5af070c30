"""Shared replay logic for the agent golden vectors (tests/golden/agent_*.npz, written by
oracle/gen_golden_agent.py from the UNMODIFIED reference my_agents/ddpg.py)."""
import os

import numpy as np

from conftest import GOLDEN

STAT_NAMES = ("critic_loss", "policy_loss", "q_max", "q_min", "q_mean", "y_max", "y_min", "y_mean")
NETS = ("policy", "critic", "target_policy", "target_critic")
BATCH_NAMES = ("states", "actions", "rewards", "next_states", "terminated", "truncated")


def load(fname):
    d = np.load(os.path.join(GOLDEN, fname), allow_pickle=False)
    meta = {k: d[k].item() for k in ("seed", "obs_dim", "act_dim", "out_act", "batch", "base_scale", "clip_low", "T_e",
                                     "lr_a", "lr_c", "discount", "polyak", "rows", "n_envs", "stride", "rounds",
                                     "init_policy_sum", "init_critic_sum")}
    return d, meta


def make_oracle(meta):
    import torch

    from oracle.agent_oracle import AgentOracle

    torch.manual_seed(meta["seed"])
    return AgentOracle(meta["obs_dim"], meta["act_dim"], meta["out_act"], batch_size=meta["batch"],
                       discount=meta["discount"], polyak=meta["polyak"], actor_lr=meta["lr_a"], critic_lr=meta["lr_c"],
                       ou=(0.15, 0.1, meta["base_scale"]), final_scale=1e-4, exploration_timesteps=meta["T_e"],
                       random_timesteps=10, clip=(meta["clip_low"], 1.0))
