"""Per-phase %globaltimer trace of CTA 0 of the grouped-GEMM kernel (csrc/update_tc.cu) for the update's shapes.
    python profiles/tcu_trace.py
Columns (ns from kernel entry): setup done | after griddepcontrol.wait | K stage 0..3 landed | accumulator complete |
epilogue done; plus the launch's duration from CUDA events (warm: second of two identical launches)."""
import ctypes as C
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests"))
import torch  # noqa: E402

from options_in_simulation_b200 import _lib as L  # noqa: E402
from test_update_tc_gpu import to_blocked  # noqa: E402

lib = L.lib()
fn = lib.ois_debug_tc_gemm
fn.restype = C.c_int
fn.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_void_p, C.c_int, C.c_int] + [C.c_int] * 5 + [C.c_void_p, C.c_int, C.c_int, C.c_void_p]
tr = lib.ois_debug_tcu_trace
tr.argtypes = [C.c_void_p, C.c_int]

for (M, N, K, tile_n, mode, name) in [(512, 512, 512, 32, 0, "fwd L2 t32"), (512, 512, 512, 64, 0, "fwd L2 t64"), (512, 128, 512, 32, 0, "fwd L3 t32"),
                                       (512, 512, 512, 128, 0, "fwd L2"), (512, 512, 512, 128, 2, "dgrad L2"),
                                       (512, 512, 512, 128, 3, "wgrad L2"), (512, 512, 16, 128, 0, "fwd L1"),
                                       (512, 16, 128, 16, 0, "fwd L4"), (512, 128, 512, 128, 0, "fwd L3")]:
    ks = min(K, 128)
    A, B = torch.randn(M, K), torch.randn(N, K)
    a_store, a_rb = (A.T.contiguous(), ks) if mode & 1 else (A, 128)
    b_store, b_rb = (B.T.contiguous(), ks) if mode & 2 else (B, tile_n)
    a_blk, b_blk = to_blocked(a_store, a_rb).cuda(), to_blocked(b_store, b_rb).cuda()
    out = torch.empty(M, N, device="cuda")
    e = [torch.cuda.Event(enable_timing=True) for _ in range(3)]
    for i in range(2):
        e[i].record()
        L.check(fn(a_blk.data_ptr(), a_store.shape[1] // 8, a_rb, b_blk.data_ptr(), b_store.shape[1] // 8, b_rb, M, N, K, tile_n,
                   mode, out.data_ptr(), N, 0, L.current_stream()))
    e[2].record()
    torch.cuda.synchronize()
    buf = (C.c_ulonglong * 16)()
    tr(buf, 16)
    t0 = buf[0]
    rel = [int(buf[i]) - int(t0) if buf[i] else None for i in range(10)]
    print(f"{name:9s} M{M} N{N} K{K} mode{mode}: event {1e3 * e[1].elapsed_time(e[2]):7.1f} us | setup {rel[1]} wait {rel[2]} "
          f"stages {rel[3:7]} acc {rel[8]} epi {rel[9]}")

# steady-state rate: 200 back-to-back programmatic-dependent launches of one GEMM
for (M, N, K, tile_n, mode, name) in [(512, 512, 512, 32, 0, "fwd L2 t32"), (512, 512, 512, 64, 0, "fwd L2 t64"), (512, 128, 512, 32, 0, "fwd L3 t32"),
                                       (512, 512, 128, 32, 0, "K128 t32"),
                                       (512, 512, 512, 128, 0, "fwd L2"), (512, 512, 512, 128, 3, "wgrad L2"), (512, 16, 128, 16, 0, "fwd L4")]:
    ks = min(K, 128)
    A, B = torch.randn(M, K), torch.randn(N, K)
    a_store, a_rb = (A.T.contiguous(), ks) if mode & 1 else (A, 128)
    b_store, b_rb = (B.T.contiguous(), ks) if mode & 2 else (B, tile_n)
    a_blk, b_blk = to_blocked(a_store, a_rb).cuda(), to_blocked(b_store, b_rb).cuda()
    out = torch.empty(M, N, device="cuda")
    call = lambda: L.check(fn(a_blk.data_ptr(), a_store.shape[1] // 8, a_rb, b_blk.data_ptr(), b_store.shape[1] // 8, b_rb, M, N, K,
                              tile_n, mode, out.data_ptr(), N, 0, L.current_stream()))
    for _ in range(10):
        call()
    g = torch.cuda.CUDAGraph()
    torch.cuda.synchronize()
    with torch.cuda.graph(g):
        for _ in range(50):
            call()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    g.replay()
    e0.record()
    for _ in range(4):
        g.replay()
    e1.record()
    torch.cuda.synchronize()
    print(f"{name:9s} back-to-back in a graph: {1e3 * e0.elapsed_time(e1) / 200:6.2f} us per launch")

# the same with the update's real epilogues (0 = bias + ELU -> bf16, 2 = ELU' -> bf16)
fe = lib.ois_debug_tc_gemm_epi
fe.restype = C.c_int
fe.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_void_p, C.c_int, C.c_int] + [C.c_int] * 6 + [C.c_void_p] * 4
for epi, name in ((0, "hidden"), (2, "dgrad")):
    M = N = K = 512
    a_blk, b_blk = to_blocked(torch.randn(M, K), 128).cuda(), to_blocked(torch.randn(N, K), 128).cuda()
    bias = torch.randn(N, device="cuda")
    h = to_blocked(torch.randn(M, N), 128).cuda()
    ob = torch.empty(M * N, dtype=torch.bfloat16, device="cuda")
    call = lambda: L.check(fe(a_blk.data_ptr(), K // 8, 128, b_blk.data_ptr(), K // 8, 128, M, N, K, 128, 0, epi, bias.data_ptr(),
                              h.data_ptr(), ob.data_ptr(), L.current_stream()))
    for _ in range(10):
        call()
    torch.cuda.synchronize()
    buf = (C.c_ulonglong * 16)()
    tr(buf, 16)
    rel = [int(buf[i]) - int(buf[0]) if buf[i] else None for i in range(10)]
    g = torch.cuda.CUDAGraph()
    with torch.cuda.graph(g):
        for _ in range(50):
            call()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    g.replay()
    e0.record()
    for _ in range(4):
        g.replay()
    e1.record()
    torch.cuda.synchronize()
    print(f"epilogue {name:7s}: {1e3 * e0.elapsed_time(e1) / 200:6.2f} us per launch back to back | acc {rel[8]} epi {rel[9]}")
