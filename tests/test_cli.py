"""options-sim-train / options-sim-eval on the B200 path (options_in_simulation_b200/cli.py; reference routers:
src/options_sim/cli/train.py, src/options_sim/cli/eval.py).  Router behaviour on CPU; short real runs on the GPU."""
import json

import pytest


def test_router_resolves_the_reference_scripts(capsys):
    from options_in_simulation_b200 import cli

    rc = cli.main(["train", "--algorithm", "kamma", "--action-dim", "4d", "--fifo", "fifo", "--curriculum", "gv",
                   "--learning-strategy", "rlil", "--dry-run", "--", "--seed", "7", "--linear_x", "1.0", "1.5"])
    d = json.loads(capsys.readouterr().out)
    assert rc == 0 and d["reference_script"] == "train/train_kamma_4d_fifo_gv_RLIL.py" and d["variant"] == "KAMMA_curr1"
    assert d["memory"] == ["fifo", 10000] and d["legacy_args"]["seed"] == 7 and d["legacy_args"]["linear_x"] == [1.0, 1.5]
    assert d["legacy_args"]["stiffness"] == 0.035 and d["timesteps"] == 600000


def test_router_lists_and_rejects(capsys):
    from options_in_simulation_b200 import cli

    assert cli.main(["train", "--algorithm", "kamma", "--action-dim", "4d", "--fifo", "nofifo", "--curriculum", "random",
                     "--learning-strategy", "rlil", "--list"]) == 0
    assert "train/train_kaddpg_1d_nofifo_random_RLIL.py" in capsys.readouterr().out
    assert cli.main(["train", "--algorithm", "kamma", "--action-dim", "1d", "--fifo", "nofifo", "--curriculum", "random",
                     "--learning-strategy", "rlil"]) == 2          # no such script in the reference either
    assert cli.main(["bogus"]) == 2


@pytest.mark.gpu
@pytest.mark.parametrize("extra", [["--precision", "fp32"], ["--precision", "bf16", "--graph-rollout"]],
                         ids=["fp32-call-by-call", "bf16-graph"])
def test_train_run(capsys, extra):
    from options_in_simulation_b200 import cli

    rc = cli.main(["train", "--algorithm", "kaddpg", "--action-dim", "2d", "--fifo", "nofifo", "--curriculum", "random",
                   "--learning-strategy", "rlil", "--num-envs", "256", "--timesteps", "120", "--write-interval", "0",
                   "--checkpoint-interval", "0", *extra, "--", "--seed", "5"])
    out = [l for l in capsys.readouterr().out.splitlines() if l.startswith("{")]
    d = json.loads(out[-1])
    assert rc == 0 and d["variant"] == "KA_DDPG_2D"
    assert d["updates"] == 2 * (120 - 80)              # learning_starts = 80, two updates per timestep afterwards
    assert d["transitions"] == 120 * 256               # 137-step ring (35 000 transitions at 256 envs) not yet full


@pytest.mark.gpu
def test_eval_run_writes_the_curve(tmp_path, capsys):
    from options_in_simulation_b200 import cli

    out_csv = tmp_path / "tracking_error_vs_speed.csv"
    rc = cli.main(["eval", "--algorithm", "kamma", "--action-dim", "4d", "--num-envs", "990", "--out", str(out_csv)])
    d = json.loads([l for l in capsys.readouterr().out.splitlines() if l.startswith("{")][-1])
    assert rc == 0 and len(d["curve"]) == 10
    rows = out_csv.read_text().strip().splitlines()
    assert rows[0] == "speed_mps,tracking_error_mps,episodes" and len(rows) == 11
    assert sum(r[2] for r in d["curve"]) >= 990        # every env finished at least one full episode after the skipped one
