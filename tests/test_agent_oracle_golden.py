"""CPU: oracle/agent_oracle.py replayed against the vectors recorded from the reference's own
my_agents/ddpg.py (act with supplied OU normals / uniforms, _update with supplied sample indices)."""
import numpy as np
import pytest
import torch

from agent_cases import BATCH_NAMES, NETS, STAT_NAMES, load, make_oracle
from conftest import golden_files


@pytest.mark.parametrize("fname", golden_files("agent_"))
def test_agent_oracle_matches_reference_vectors(fname):
    from oracle.agent_oracle import flat_params

    d, meta = load(fname)
    o = make_oracle(meta)
    # the seeded default init is the reference's (same construction order): pinned by float64 checksums
    assert flat_params(o.policy).double().sum().item() == pytest.approx(meta["init_policy_sum"], abs=1e-9)
    assert flat_params(o.critic).double().sum().item() == pytest.approx(meta["init_critic_sum"], abs=1e-9)
    ra = o.act(d["rand_states"], 3, 1000, uniforms=d["rand_uniforms"])
    assert np.array_equal(ra.numpy(), d["rand_actions"])
    for k in range(meta["rounds"]):
        t = int(d[f"r{k}_timestep"])
        a = o.act(d[f"r{k}_states"], t, 1000, normals=d[f"r{k}_normals"])
        assert np.abs(a.numpy() - d[f"r{k}_actions"]).max() <= 1e-6
        assert np.abs(np.asarray(o.noise.state) - d[f"r{k}_ou_state"]).max() <= 1e-7
        idx = d[f"r{k}_indices"]
        st = o.update(*[d[f"mem_{n}"][idx] for n in BATCH_NAMES])
        for name, ref in zip(STAT_NAMES, d[f"r{k}_stats"]):
            assert abs(st[name] - ref) <= 1e-5 * max(1.0, abs(ref)), name
        for net in NETS:
            sub = flat_params(getattr(o, net))[:: meta["stride"]].numpy()
            assert np.abs(sub - d[f"r{k}_{net}_sub"]).max() <= 1e-6, net
        assert np.abs(o.scaler.running_mean.numpy() - d[f"r{k}_scaler_mean"]).max() <= 1e-12
        assert np.abs(o.scaler.running_variance.numpy() - d[f"r{k}_scaler_var"]).max() <= 1e-12
        assert float(o.scaler.current_count) == float(d[f"r{k}_scaler_count"])
