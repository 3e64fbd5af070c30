"""The env-variant table (options_in_simulation_b200/config.py: VARIANTS) against the oracle's (oracle/env_oracle.py) and,
where the reference checkout exists (the build container only), against the constants in src/my_envs/<variant>.py
(SURVEY.md §2.2).  Host logic only: no GPU, no compute calls."""
import os
import re

import pytest

REF_ENVS = "/root/reference/src/my_envs"


def _tables():
    from options_in_simulation_b200 import _lib as L
    from options_in_simulation_b200.config import VARIANTS as PRODUCT
    from oracle.env_oracle import VARIANTS as ORACLE

    return L, PRODUCT, ORACLE


def test_product_and_oracle_tables_agree():
    L, product, oracle = _tables()
    assert set(product) == set(oracle)
    blend = {"linear": L.BLEND_LINEAR, "switch": L.BLEND_SWITCH}
    command = {"rand": L.CMD_RAND, "gv": L.CMD_GV, "bd": L.CMD_BD, "eval": L.CMD_EVAL}
    gamma = {"decay": (L.GAMMA_DECAY, None), "one": (L.GAMMA_CONST, 1.0), "zero": (L.GAMMA_CONST, 0.0)}
    for name, o in oracle.items():
        p = product[name]
        assert (p.name, p.act_dim, p.max_sim_steps, bool(p.eval_sums), p.action_low) == \
               (o.name, o.act_dim, o.max_sim_steps, bool(o.eval_sums), o.action_low), name
        assert p.blend == blend[o.blend] and p.command_mode == command[o.command], name
        assert p.guiding_post_clip == (1 if o.guiding == "post" else 0), name
        for mode, const, kind in ((p.gamma1_mode, p.gamma1_const, o.gamma1), (p.gamma2_mode, p.gamma2_const, o.gamma2)):
            want_mode, want_const = gamma[kind]
            assert mode == want_mode and (want_const is None or const == want_const), name


@pytest.mark.skipif(not os.path.isdir(REF_ENVS), reason="reference checkout only exists in the build container")
def test_table_matches_the_reference_sources():
    """One row per src/my_envs/*.py env file: every file has a variant, and the constants a variant carries are the ones
    its file assigns (max_sim_steps, the eval sums, the command sampler that wins in sample_velocity_command)."""
    _, product, _ = _tables()
    files = sorted(f[:-3] for f in os.listdir(REF_ENVS) if f.endswith(".py") and f not in ("__init__.py", "rl_task.py"))
    assert files == sorted(product)
    for name in files:
        src = open(os.path.join(REF_ENVS, name + ".py")).read()
        v = product[name]
        m = re.search(r"^\s*self\.max_sim_steps\s*=\s*([0-9.eE+]+)", src, re.M)
        assert m and float(m.group(1)) == v.max_sim_steps, name
        assert ('"Tracking error": torch_zeros()' in src) == bool(v.eval_sums), name
        assert bool(re.search(r'"EvalSampling":\s*True', src)) == (v.command_mode == _tables()[0].CMD_EVAL), name
        act = {"KA_DDPG_1D": 1, "KA_DDPG_2D": 2}.get(name.replace("_eval", ""), 4)
        assert v.act_dim == act, name
