"""CPU: the restated training loop (oracle/trainer_oracle.py = EnvOracle + AgentOracle in the reference's call
order) replays the vectors recorded from the UNMODIFIED reference trainer + task + agent
(tests/golden/trainer_*.npz, written by oracle/gen_golden_trainer.py).  Pins SURVEY.md §8 row A12: relabel,
two updates per step, the wrapper's reset step, ring-buffer wrap."""
import numpy as np
import pytest

from conftest import golden_files
from trainer_cases import load_case, make_state


@pytest.mark.parametrize("fname", golden_files("trainer_"))
def test_trainer_oracle_matches_reference(fname):
    import torch

    from oracle.agent_oracle import flat_params
    from oracle.env_oracle import VARIANTS
    from oracle.trainer_oracle import TrainerOracle

    d, m, draws = load_case(fname)
    torch.manual_seed(m["seed"])
    orc = TrainerOracle(VARIANTS[m["variant"]], make_state(d), n=m["n"], batch=m["batch"], mem_steps=m["mem_steps"],
                        random_timesteps=m["random_timesteps"], learning_starts=m["learning_starts"],
                        ou_scale=m["ou_scale"], T_e=m["T_e"], clip_low=m["clip_low"], relabel=bool(m["relabel"]),
                        clip_obs=m["clip_obs"])
    orc.run(m["timesteps"], draws, d["indexes"])
    assert orc.updates == m["n_updates"] == 2 * (m["timesteps"] - m["learning_starts"])
    for t in range(m["timesteps"]):
        assert np.array_equal(orc.log[t]["reset"], d["reset"][t].astype(bool))
        assert np.abs(orc.log[t]["actions"] - d["actions"][t]).max() <= 1e-5
        for key in ("obs", "rew", "wheel_torqs"):
            ref = d[key][t]
            assert np.abs(orc.log[t][key] - ref).max() <= 2e-5 * max(1.0, np.abs(ref).max()), (t, key)
    assert orc.memory_index == m["memory_index"] and orc.filled == bool(m["filled"])
    assert np.array_equal(orc.memory["terminated"], d["mem_terminated"].astype(bool))
    for name in ("states", "actions", "rewards", "next_states"):
        ref = d[f"mem_{name}"]
        assert np.array_equal(np.isnan(ref), np.isnan(orc.memory[name]))
        assert np.nanmax(np.abs(ref - orc.memory[name])) <= 2e-5 * max(1.0, np.nanmax(np.abs(ref)))
    for name in ("policy", "critic", "target_policy", "target_critic"):
        flat = flat_params(getattr(orc.agent, name))
        assert np.abs(flat[::m["stride"]].numpy() - d[f"{name}_sub"]).max() <= 2e-5
        assert flat.double().sum().item() == pytest.approx(float(d[f"{name}_sum"]), abs=2e-3)
    assert np.allclose(orc.agent.scaler.running_mean.numpy(), d["scaler_mean"], rtol=1e-6, atol=1e-9)
    assert float(orc.agent.scaler.current_count) == float(d["scaler_count"])
