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


def test_eval_router_covers_the_reference_eval_map(capsys):
    """src/options_sim/cli/eval.py:10-14: three (algorithm, action_dim) keys; the 1d / 2d scripts use the *_eval env twins and a
    sigmoid actor (eval/eval_kaddpg_1d.py:29-36,68), the 4d one KAMMA_eval and a tanh (eval/eval_kamma_4d.py:33-36,64)."""
    from options_in_simulation_b200 import cli
    from options_in_simulation_b200.config import VARIANTS

    want = {("kaddpg", "1d"): ("eval/eval_kaddpg_1d.py", "KA_DDPG_1D_eval", "sigmoid", 1),
            ("kaddpg", "2d"): ("eval/eval_kaddpg_2d.py", "KA_DDPG_2D_eval", "sigmoid", 2),
            ("kamma", "4d"): ("eval/eval_kamma_4d.py", "KAMMA_eval", "tanh", 4)}
    for (alg, dim), (script, variant, output, act_dim) in want.items():
        assert cli.main(["eval", "-a", alg, "-d", dim, "--dry-run", "--", "--seed", "3"]) == 0
        d = json.loads(capsys.readouterr().out)
        assert (d["reference_script"], d["variant"], d["output"]) == (script, variant, output)
        assert d["legacy_args"]["seed"] == 3
        v = VARIANTS[variant]
        assert v.act_dim == act_dim and v.eval_sums and v.max_sim_steps == 50000.0 and v.gamma1_const == 0.0
    assert cli.main(["eval", "-a", "kamma", "-d", "1d", "--dry-run"]) == 2     # no such script in the reference either


def test_unrouted_scripts_are_reachable_by_path(capsys):
    """train/train_kamma_controller.py and eval/eval_kamma_controller.py are in no router map of the reference."""
    from options_in_simulation_b200 import cli

    assert cli.main(["train", "--script", "train/train_kamma_controller.py", "--dry-run"]) == 0
    d = json.loads(capsys.readouterr().out)
    assert d["variant"] == "KAMMA_g11_g21" and d["memory"] == ["random", 1000000] and d["timesteps"] == 400000
    assert cli.main(["eval", "--script", "eval/eval_kamma_controller.py", "--dry-run"]) == 0
    d = json.loads(capsys.readouterr().out)
    assert d["variant"] == "KAMMA_eval" and d["output"] == "sigmoid"
    assert cli.main(["train", "--list"]) == 0
    assert "train/train_kamma_controller.py" in capsys.readouterr().out


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


@pytest.mark.gpu
@pytest.mark.parametrize("dim,variant", [("1d", "KA_DDPG_1D_eval"), ("2d", "KA_DDPG_2D_eval")])
def test_eval_run_of_the_kaddpg_scripts(capsys, dim, variant):
    """eval/eval_kaddpg_1d.py / eval_kaddpg_2d.py on the B200 path: the sweep over the *_eval env twins yields one row per
    speed bin, every env finishes at least one episode, and the binned speeds stay inside the ramp's range."""
    from options_in_simulation_b200 import cli

    rc = cli.main(["eval", "--algorithm", "kaddpg", "--action-dim", dim, "--num-envs", "495", "--bins", "5"])
    d = json.loads([l for l in capsys.readouterr().out.splitlines() if l.startswith("{")][-1])
    assert rc == 0 and d["reference_script"] == f"eval/eval_kaddpg_{dim}.py" and len(d["curve"]) == 5
    assert sum(r[2] for r in d["curve"]) >= 495
    assert all(1.0 <= r[0] <= 2.0 and r[1] >= 0.0 for r in d["curve"] if r[2] > 0)


@pytest.mark.gpu
def test_unrouted_scripts_run(capsys):
    """train/train_kamma_controller.py (KAMMA_g11_g21: the pure controller; random_timesteps = learning_starts = 1000, so a short
    run records transitions and makes no update) and eval/eval_kamma_controller.py (KAMMA_eval, sigmoid actor) by path."""
    from options_in_simulation_b200 import cli

    rc = cli.main(["train", "--script", "train/train_kamma_controller.py", "--num-envs", "128", "--timesteps", "60",
                   "--write-interval", "0", "--checkpoint-interval", "0", "--", "--seed", "3"])
    d = json.loads([l for l in capsys.readouterr().out.splitlines() if l.startswith("{")][-1])
    assert rc == 0 and d["variant"] == "KAMMA_g11_g21" and d["updates"] == 0 and d["transitions"] == 60 * 128
    rc = cli.main(["eval", "--script", "eval/eval_kamma_controller.py", "--num-envs", "495", "--bins", "5"])
    d = json.loads([l for l in capsys.readouterr().out.splitlines() if l.startswith("{")][-1])
    assert rc == 0 and d["reference_script"] == "eval/eval_kamma_controller.py" and len(d["curve"]) == 5
    assert sum(r[2] for r in d["curve"]) >= 495
