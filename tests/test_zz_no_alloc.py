"""ois_mlp_forward_static (csrc/no_alloc.cu): the fp32 MLP forward with its hidden activations in a static device array,
for callers that may not allocate.  No counterpart in the reference, so the
checker is the existing entry point ois_mlp_forward — same kernels, so the outputs must be bit-identical (the fp32 tests'
tolerance, 1e-5 relative, is asserted as well, against a plain torch fp32 restatement).  Collected last on
purpose."""
import ctypes as C

import pytest
import torch

import options_in_simulation_b200._lib as L

pytestmark = pytest.mark.gpu

HIDDEN = (512, 512, 128)
MAX_ROWS = 4096          # OIS_STATIC_MLP_MAX_ROWS (include/ois_b200.h)
ACT_SIGMOID, PRECISION_FP32 = 1, 0


def _flat_params(shape, gen):
    dims = (shape.in_dim, shape.h0, shape.h1, shape.h2, shape.out_dim)
    parts = []
    for l in range(4):
        bound = 1.0 / dims[l] ** 0.5
        parts.append((torch.rand(dims[l + 1] * dims[l], generator=gen) * 2 - 1) * bound)
        parts.append((torch.rand(dims[l + 1], generator=gen) * 2 - 1) * bound)
    return torch.cat(parts).float().cuda().contiguous(), dims


def _torch_mlp(flat, dims, x, sigmoid):
    off, h = 0, x
    for l in range(4):
        w = flat[off:off + dims[l + 1] * dims[l]].view(dims[l + 1], dims[l])
        off += dims[l + 1] * dims[l]
        b = flat[off:off + dims[l + 1]]
        off += dims[l + 1]
        h = h @ w.t() + b
        if l < 3:
            h = torch.nn.functional.elu(h)
    return torch.sigmoid(h) if sigmoid else h


@pytest.fixture
def handle(built_lib):
    lib = built_lib
    shape = L.MlpShape(6, *HIDDEN, 4, ACT_SIGMOID)
    critic = L.MlpShape(10, *HIDDEN, 1, 0)
    h = C.c_void_p()
    L.check(lib.ois_agent_create(C.byref(h), C.byref(shape), C.byref(critic), MAX_ROWS, 128, PRECISION_FP32, L.current_stream()))
    yield lib, h
    lib.ois_agent_destroy(h)


def test_static_scratch_forward_matches_the_handle_forward_on_two_calls_in_a_row(handle):
    lib, h = handle
    gen = torch.Generator().manual_seed(20261105)
    backend_prefer = torch.backends.cuda.matmul.allow_tf32
    torch.backends.cuda.matmul.allow_tf32 = False
    try:
        # second call: fewer rows, other shape / weights / inputs — the array still holds the first call's activations
        for rows, in_dim, out_dim, act in ((MAX_ROWS, 6, 4, ACT_SIGMOID), (1000, 10, 1, 0), (1, 6, 4, ACT_SIGMOID)):
            shape = L.MlpShape(in_dim, *HIDDEN, out_dim, act)
            flat, dims = _flat_params(shape, gen)
            x = torch.randn(rows, in_dim, generator=gen).cuda().contiguous()
            want = torch.full((rows, out_dim), float("nan"), device="cuda")
            got = torch.full((rows, out_dim), float("nan"), device="cuda")
            s = L.current_stream()
            L.check(lib.ois_mlp_forward(h, C.byref(shape), L.ptr(flat), L.ptr(x), L.ptr(want), rows, s))
            torch.cuda.synchronize()
            free_before, torch_before = torch.cuda.mem_get_info()[0], torch.cuda.memory_allocated()
            L.check(lib.ois_mlp_forward_static(C.byref(shape), L.ptr(flat), L.ptr(x), L.ptr(got), rows, s))
            torch.cuda.synchronize()
            if rows != MAX_ROWS:   # (the first call may load the module lazily; later ones must not move device memory)
                assert torch.cuda.mem_get_info()[0] == free_before and torch.cuda.memory_allocated() == torch_before
            assert torch.isfinite(got).all()
            assert torch.equal(got, want)
            ref = _torch_mlp(flat, dims, x, act == ACT_SIGMOID)
            assert ((got - ref).abs() / ref.abs().clamp(min=1.0)).max().item() <= 1e-5
    finally:
        torch.backends.cuda.matmul.allow_tf32 = backend_prefer


def test_static_scratch_forward_rejects_sizes_beyond_its_maximum(built_lib):
    lib = built_lib
    shape = L.MlpShape(6, *HIDDEN, 4, ACT_SIGMOID)
    flat, _ = _flat_params(shape, torch.Generator().manual_seed(1))
    x = torch.zeros(MAX_ROWS + 1, 6, device="cuda")
    y = torch.full((MAX_ROWS + 1, 4), 7.0, device="cuda")
    s = L.current_stream()
    assert lib.ois_mlp_forward_static(C.byref(shape), L.ptr(flat), L.ptr(x), L.ptr(y), MAX_ROWS + 1, s) != 0
    msg = lib.ois_last_error_string().decode()
    assert "OIS_STATIC_MLP_MAX_ROWS" in msg and "4096" in msg
    wide = L.MlpShape(6, 1024, 512, 128, 4, ACT_SIGMOID)
    assert lib.ois_mlp_forward_static(C.byref(wide), L.ptr(flat), L.ptr(x), L.ptr(y), 16, s) != 0
    assert "hidden sizes" in lib.ois_last_error_string().decode()
    assert lib.ois_mlp_forward_static(C.byref(shape), L.ptr(flat), L.ptr(x), L.ptr(y), 0, s) == 0   # empty input: a no-op
    torch.cuda.synchronize()
    assert (y == 7.0).all()
