"""Needs >= 2 GPUs on one host (skipped on the single-GPU box): the NCCL data-parallel update path and bench.py's
multi-rank contract, each spawned with torchrun exactly as the driver does."""
import json
import os
import subprocess
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
pytestmark = pytest.mark.gpu


def _gpus():
    import torch

    return torch.cuda.device_count() if torch.cuda.is_available() else 0


def _torchrun(n, script, *args, port=29517):
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={n}", "--master-addr", "127.0.0.1",
           "--master-port", str(port), script, *args]
    return subprocess.run(cmd, cwd=ROOT, capture_output=True, text=True, timeout=600)


@pytest.mark.skipif(_gpus() < 2, reason="needs 2 GPUs")
def test_ddpg_update_is_data_parallel_on_nccl():
    r = _torchrun(2, os.path.join("tests", "scripts", "dist_update_check.py"))
    assert r.returncode == 0 and "dist_update_check ok" in r.stdout, r.stdout[-2000:] + r.stderr[-4000:]


@pytest.mark.skipif(_gpus() < 2, reason="needs 2 GPUs")
def test_bench_two_ranks_prints_one_aggregate_line():
    r = _torchrun(2, "bench.py", "--gpus", "2", "--steps", "30", "--warmup", "5", port=29519)
    assert r.returncode == 0, r.stderr[-4000:]
    lines = [l for l in r.stdout.splitlines() if l.startswith("{")]
    assert len(lines) == 1
    d = json.loads(lines[0])
    assert d["n_gpus"] == 2 and d["scaling"] == "weak" and d["value"] > 0 and d["e2e"]["value"] > 0
    # the data-parallel update leg: graph-captured, gradients exchanged over peer memory and (second leg) NCCL,
    # bit-identical weights on both ranks (bench.py exits non-zero otherwise)
    u = d["update"]
    assert u["ms_per_update"] > 0 and u["train_env_steps_per_s"] > 0
    c = u["collective"]
    assert c["weights_identical_across_ranks"] and c["in_update_graph"] and set(c["ms_per_update"]) >= {"local", c["exchange"]}


@pytest.mark.skipif(_gpus() < 2, reason="needs 2 GPUs")
def test_eval_sweep_shards_over_ranks_and_reduces_once():
    r = _torchrun(2, os.path.join("tests", "scripts", "eval_sweep_ranks.py"), port=29521)
    assert r.returncode == 0 and "eval_sweep_ranks ok" in r.stdout, r.stdout[-2000:] + r.stderr[-4000:]
