"""rmem_groupnorm_tokens_static: the allocation-free variant of the token GroupNorm (its statistics scratch is a
file-scope ``static __device__`` array inside librmem_b200.so, see include/rmem_b200.h).  No counterpart in the
reference; the checker is the existing entry point rmem_groupnorm_tokens with a caller workspace.

This file is named so that pytest collects it LAST: a failure here must not
hide any other test.  It passes alone in a fresh process and after the rest of the suite."""
import ctypes

import pytest
import torch

pytestmark = pytest.mark.gpu


def _p(t):
    return None if t is None else ctypes.c_void_p(t.data_ptr())


def _call_both(lib, xa, xb, g, b, groups, out_f32):
    """One call through each C-ABI entry point on the current stream; returns (static, workspace) outputs."""
    L, Ca = xa.shape
    C = Ca + (xb.shape[1] if xb is not None else 0)
    dt = torch.float32 if out_f32 else torch.bfloat16
    o_static = torch.empty(L, C, device="cuda", dtype=dt)
    o_ws = torch.empty(L, C, device="cuda", dtype=dt)
    ws = torch.zeros(2 * groups + 1, device="cuda", dtype=torch.float64)
    stream = ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)
    before = torch.cuda.memory_allocated()
    rc = lib.rmem_groupnorm_tokens_static(_p(xa), Ca, _p(xb), _p(g), _p(b), _p(o_static), int(out_f32), L, C,
                                          groups, stream)
    assert rc == 0, lib.rmem_last_error()
    assert torch.cuda.memory_allocated() == before          # nothing went through torch's allocator either
    rc = lib.rmem_groupnorm_tokens(_p(xa), Ca, _p(xb), _p(g), _p(b), _p(o_ws), int(out_f32), L, C, groups,
                                   _p(ws), stream)
    assert rc == 0, lib.rmem_last_error()
    torch.cuda.synchronize()
    return o_static, o_ws


def test_static_scratch_groupnorm_matches_workspace_entry_point_twice(lib_built):
    lib = lib_built.lib()
    torch.manual_seed(11)
    # two calls in a row on different inputs and shapes: the shared scratch keeps the first call's state, so
    # the second one is only right if the entry point cleared it
    cases = [(1620, 256, 256, 2, True), (777, 192, 64, 8, False)]
    for (L, Ca, Cb, groups, out_f32) in cases:
        xa = torch.randn(L, Ca, device="cuda") * 2 + 0.5
        xb = torch.randn(L, Cb, device="cuda") * 0.3 - 1 if Cb else None
        g, b = torch.randn(Ca + Cb, device="cuda"), torch.randn(Ca + Cb, device="cuda")
        o_static, o_ws = _call_both(lib, xa, xb, g, b, groups, out_f32)
        ref = o_ws.float()
        rel = ((o_static.float() - ref).norm() / ref.norm()).item()
        # same kernels, same inputs: only the order of the fp64 atomics differs (f32 out: 1e-5 as test_gpu_ops;
        # bf16 out: one rounding step, 4e-3 as test_gpu_ops)
        assert rel < (1e-5 if out_f32 else 4e-3), (L, Ca, Cb, groups, rel)
        assert torch.isfinite(o_static.float()).all()
    # and the same call twice, back to back, gives the same answer both times
    xa = torch.randn(1705, 256, device="cuda")
    xb = torch.randn(1705, 256, device="cuda")
    g, b = torch.randn(512, device="cuda"), torch.randn(512, device="cuda")
    first, ref1 = _call_both(lib, xa, xb, g, b, 2, True)
    second, _ = _call_both(lib, xa, xb, g, b, 2, True)
    assert ((first - ref1).norm() / ref1.norm()).item() < 1e-5
    assert ((second - ref1).norm() / ref1.norm()).item() < 1e-5


def test_static_scratch_groupnorm_rejects_sizes_beyond_the_maximum(lib_built):
    lib = lib_built.lib()
    L, C, groups = 64, 1024, 16                              # 16 groups > RMEM_GN_STATIC_MAX_GROUPS (8)
    x = torch.randn(L, C, device="cuda")
    g, b = torch.ones(C, device="cuda"), torch.zeros(C, device="cuda")
    out = torch.empty(L, C, device="cuda")
    rc = lib.rmem_groupnorm_tokens_static(_p(x), C, None, _p(g), _p(b), _p(out), 1, L, C, groups,
                                          ctypes.c_void_p(torch.cuda.current_stream().cuda_stream))
    assert rc == -1
    assert b"RMEM_GN_STATIC_MAX_GROUPS" in lib.rmem_last_error()
    torch.cuda.synchronize()
