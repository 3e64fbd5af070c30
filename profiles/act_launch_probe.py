import sys, os, torch
sys.path.insert(0, '/root/repo')
import bench
n = int(sys.argv[1]) if len(sys.argv) > 1 else 4096
w = bench.World(bench.WORKLOADS["kamma_4d"], n, "bf16", torch.device("cuda", 0), seed=1)
flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")
for cold in (False, True):
    ts = []
    for it in range(20):
        ev = [torch.cuda.Event(enable_timing=True) for _ in range(2)]
        if cold: flush.zero_()
        ev[0].record()
        actions = w.agent.act(w.obs, w.t, w.T)[0]
        ev[1].record()
        torch.cuda.synchronize()
        ts.append(ev[0].elapsed_time(ev[1]) * 1e3)
    ts.sort()
    print(f"n={n} debug={os.environ.get('OIS_TCT_DEBUG_MODE')} maxrows={os.environ.get('OIS_TCT_MAX_ROWS')} {'cold' if cold else 'warm'}: act events median {ts[10]:.2f} min {ts[0]:.2f} us")
