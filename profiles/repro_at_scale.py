import sys, torch
sys.path.insert(0, '/root/repo')
import bench
wl = bench.WORKLOADS["kamma_4d"]
flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")
for big in [int(x) for x in sys.argv[1:]]:
    w = bench.World(wl, big, "bf16", torch.device("cuda", 0), seed=99)
    small = big * 335 < (96 << 20)
    try:
        a_ms, e_ms = bench.kernel_split(w, 20, flush if small else None)
        print(big, "ok", a_ms, e_ms, flush=True)
    except Exception as e:
        print(big, "FAILED", str(e)[:100], flush=True)
        break
    del w
    torch.cuda.empty_cache()
