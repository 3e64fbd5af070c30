"""Time the dynamics-only kernel and the full env kernel at 1M envs with a library variant (diagnostic builds:
make EXTRA=-DOIS_DIAG_...).  usage: python profiles/variant_probe.py <lib.so> [<lib.so> ...]   (one process per library)"""
import os
import subprocess
import sys

import numpy as np


def worker(path):
    import torch

    sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
    from options_in_simulation_b200 import _lib as L
    L.LIB_PATH = os.path.abspath(path)
    import options_in_simulation_b200 as ois
    from options_in_simulation_b200.config import script_cfg
    from oracle.dynamics_twin import random_state

    n = 1 << 20
    rng = np.random.default_rng(7)
    st = {k: v.astype(np.float32) for k, v in random_state(n, rng).items()}
    tau = torch.as_tensor(rng.uniform(-4, 4, (n, 4)).astype(np.float32)).cuda()
    act = torch.rand(n, 4, device="cuda")
    task = ois.TorqueDistributionTask(sim_config=script_cfg(n), variant="KAMMA_medium", seed=1)
    task.load_state(px=st["px"], py=st["py"], yaw=st["yaw"], vx=st["vx"], vy=st["vy"], wz=st["wz"], w=st["w"])

    def timed(fn, reps=20):
        for _ in range(3):
            fn()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(reps):
            fn()
        e1.record()
        torch.cuda.synchronize()
        return e0.elapsed_time(e1) / reps * 1e3

    t_dyn = timed(lambda: task.dynamics_step(tau))
    t_env = timed(lambda: task.step(act))
    print(f"{os.path.basename(path)}: dynamics {t_dyn:.1f} us, env step {t_env:.1f} us (1M envs, 20 launches back to back)", flush=True)


if __name__ == "__main__":
    if os.environ.get("VARIANT_WORKER"):
        worker(sys.argv[1])
    else:
        for p in sys.argv[1:]:
            subprocess.run([sys.executable, __file__, p], check=False, env={**os.environ, "VARIANT_WORKER": "1"})
