import ctypes, sys, torch
sys.path.insert(0, "tests"); sys.path.insert(0, ".")
from test_rollout_graph_gpu import make_world, T_TOTAL
from options_in_simulation_b200.rollout import RolloutGraph
from options_in_simulation_b200 import _lib as L
rt = ctypes.CDLL("libcudart.so.12")
def status(tag):
    st = ctypes.c_int(-1)
    r = rt.cudaStreamIsCapturing(ctypes.c_void_p(torch.cuda.current_stream().cuda_stream), ctypes.byref(st))
    print(tag, "rc", r, "status", st.value, flush=True)
for variant, n in (("KAMMA_medium", 640), ("KAMMA_eval", 640)):
    w = make_world(variant, n, "bf16")
    g = RolloutGraph(w["env"], w["agent"], 1000, T_TOTAL)
    g.step()
    torch.cuda.synchronize()
    g._write_ctl()
    graph = torch.cuda.CUDAGraph()
    try:
        with torch.cuda.graph(graph):
            status("begin")
            ctl = g._ctl.data_ptr()
            g.agent.act(g.states, g.timestep, g.timesteps, _ctl=ctl, _out=g._actions)
            status("after act")
            g.task.step(g._actions, True, _ctl=ctl)
            status("after env")
            L.check(L.lib().ois_step_ctl_advance(ctl, L.current_stream()))
            status("after advance")
        print(variant, "captured ok")
    except Exception as e:
        print(variant, "FAILED", str(e)[:200])
        break
