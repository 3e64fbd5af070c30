import sys, torch
sys.path.insert(0, '/root/repo')
import bench
wl = bench.WORKLOADS["kamma_4d"]
n = int(sys.argv[1]); mode = sys.argv[2]
w = bench.World(wl, n, "bf16", torch.device("cuda", 0), seed=99)
torch.cuda.synchronize(); print("world ok", flush=True)
for it in range(5):
    if mode == "act":
        a = w.agent.act(w.obs, w.t, w.T)[0]
    elif mode == "env":
        a = torch.rand(n, 4, device="cuda"); w.env.step(a)
    else:
        w.step()
    torch.cuda.synchronize(); print(mode, it, "ok", flush=True)
