"""Host-side pieces of bench.py that run without a GPU: the clock sampler (against a stand-in `nvidia-smi` that
prints the recipe's CSV line) and the host-core count the CPU legs use."""
import importlib.util
import os
import stat
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _bench():
    spec = importlib.util.spec_from_file_location("ois_bench", os.path.join(ROOT, "bench.py"))
    mod = importlib.util.module_from_spec(spec)
    sys.modules["ois_bench"] = mod
    spec.loader.exec_module(mod)
    return mod


def test_clock_sampler_parses_the_recipe_line(tmp_path, monkeypatch):
    fake = tmp_path / "nvidia-smi"
    fake.write_text("#!/bin/bash\nwhile true; do echo '1965, 1980, Not Active, Not Active, Not Active, Active'; sleep 0.02; done\n")
    fake.chmod(fake.stat().st_mode | stat.S_IEXEC)
    monkeypatch.setenv("PATH", f"{tmp_path}:{os.environ['PATH']}")
    sampler = _bench().ClockSampler(0)
    sampler.start()
    time.sleep(0.3)
    got = sampler.stop()
    assert got["sm_mhz"] == 1965 and got["sm_max_mhz"] == 1980
    assert got["reasons"] == ["sw_power_cap"] and got["samples"] >= 3


def test_clock_sampler_without_nvidia_smi(tmp_path, monkeypatch):
    monkeypatch.setenv("PATH", str(tmp_path))
    sampler = _bench().ClockSampler(0)
    sampler.start()
    assert sampler.stop()["reasons"] == ["nvidia-smi unavailable"]


def test_host_cores_is_within_the_affinity_mask():
    n = _bench().host_cores()
    assert 1 <= n <= len(os.sched_getaffinity(0))
