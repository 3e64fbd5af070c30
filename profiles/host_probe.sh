#!/bin/bash
# what the box gives the CPU legs: core counts, cgroup quota, and the reference arm alone vs under torchrun
python - <<'PY'
import os
print("cpu_count", os.cpu_count(), "affinity", len(os.sched_getaffinity(0)))
for p in ("/sys/fs/cgroup/cpu.max", "/sys/fs/cgroup/cpu/cpu.cfs_quota_us", "/sys/fs/cgroup/cpu/cpu.cfs_period_us"):
    try: print(p, open(p).read().strip())
    except OSError as e: print(p, "-")
print("loadavg", open("/proc/loadavg").read().strip())
PY
python bench.py --impl reference --steps 10 --warmup 3 | cut -c1-230
python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29515 bench.py --impl reference --gpus 2 --steps 10 --warmup 3 2>/dev/null | cut -c1-230
