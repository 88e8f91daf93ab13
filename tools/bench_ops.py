"""Per-operator timings at the DeAOT-L 480p shapes (L = 1620): CUDA events around `iters` back-to-back launches
of each op on rotating buffers (so producer -> consumer L2 reuse is similar to the real frame, where every
activation was just written).  Development helper; the judged numbers come from bench.py."""
import sys
import torch
sys.path.insert(0, ".")
from rmem_ocu_b200 import ops

dev = "cuda"
L, h, w = 1620, 30, 54
bf = lambda *s: torch.randn(*s, device=dev).to(torch.bfloat16)
f32 = lambda *s: torch.randn(*s, device=dev)
NB = 4


def timed(name, fn, iters=20):
    """Device time per launch: `iters` launches captured in one CUDA graph (no host launch cost), replayed 5x."""
    for i in range(3):
        fn(i % NB)
    torch.cuda.synchronize()
    g = torch.cuda.CUDAGraph()
    with torch.cuda.graph(g):
        for i in range(iters):
            fn(i % NB)
    g.replay()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(5):
        g.replay()
    e1.record()
    torch.cuda.synchronize()
    print("%-44s %8.2f us" % (name, e0.elapsed_time(e1) / (5 * iters) * 1e3), flush=True)


def gemm_case(name, M, N, K, mode, act=ops.ACT_NONE):
    a = [bf(M, K) for _ in range(NB)]
    wt = [bf(N, K) for _ in range(NB)]
    b = f32(N)
    if mode == ops.OUT_BF16:
        d = [torch.zeros(M, N, device=dev, dtype=torch.bfloat16) for _ in range(NB)]
    else:
        d = [torch.zeros(M, N, device=dev) for _ in range(NB)]
    timed("gemm %-18s M%d N%d K%d" % (name, M, N, K),
          lambda i: ops.linear(a[i], wt[i], b, [(0, N, act, mode, d[i], 0)]))


gemm_case("qvu bf16+silu", L, 1152, 256, ops.OUT_BF16, ops.ACT_SILU)
gemm_case("idu bf16+silu", L, 512, 256, ops.OUT_BF16, ops.ACT_SILU)
gemm_case("rel f32", L, 256, 128, ops.OUT_F32)
gemm_case("proj_ls accum", L, 512, 2048, ops.OUT_F32_ACCUM)
gemm_case("sa in bf16+silu", L, 2176, 512, ops.OUT_BF16, ops.ACT_SILU)
gemm_case("proj_sa accum", L, 512, 1024, ops.OUT_F32_ACCUM)
gemm_case("idv bf16+silu", L, 512, 512, ops.OUT_BF16, ops.ACT_SILU)
gemm_case("aot 256 bf16", L, 256, 256, ops.OUT_BF16)
gemm_case("aot ff1 f32", L, 1024, 256, ops.OUT_F32)
gemm_case("aot ff2 accum", L, 256, 1024, ops.OUT_F32_ACCUM)

x = [bf(L, 1024) for _ in range(NB)]
y = [bf(L, 1024) for _ in range(NB)]
wt = f32(25, 1024)
timed("dwconv5 1024ch", lambda i: ops.dwconv5(x[i], 0, wt, y[i], 0, h, w, 1024))
t = [f32(L, 256) for _ in range(NB)]
o = [bf(L, 512) for _ in range(NB)]
g, b = f32(256), f32(256)
timed("layernorm dual", lambda i: ops.layernorm(t[i], g, b, o[i], 0, t[(i + 1) % NB], g, b, o[i], 256))
q = [bf(L, 128) for _ in range(NB)]
k = [bf(L, 128) for _ in range(NB)]
v = [bf(L, 1024) for _ in range(NB)]
u = [bf(L, 1024) for _ in range(NB)]
rel = [f32(L, 256) * 0.1 for _ in range(NB)]
probs = [torch.empty(L, 225, device=dev) for _ in range(NB)]
banded = [ops.local_banded_buffer(h, w, dev) for _ in range(NB)]
timed("local scores", lambda i: ops.local_attn_scores(q[i], k[i], rel[i], None, banded[i], h, w))
timed("local aggregate", lambda i: ops.local_attn_aggregate(banded[i], v[i], u[i], y[i], h, w))
for T in (1, 5):
    from rmem_ocu_b200.engine import _chunks_per_frame
    cpf = _chunks_per_frame(T, L)
    kb = [bf(T, L, 128) for _ in range(NB)]
    vb = [bf(T, L, 1024) for _ in range(NB)]
    po = torch.empty(T * cpf, L, 1024, device=dev)
    pm, pl = torch.empty(T * cpf, L, device=dev), torch.empty(T * cpf, L, device=dev)
    order = torch.arange(T, device=dev, dtype=torch.int32)
    mass = torch.empty(L, T, device=dev)
    timed("lt partial T=%d cpf=%d" % (T, cpf), lambda i: ops.lt_attn_partial(q[i], kb[i], vb[i], order, T, cpf, po, pm, pl))
    timed("lt combine T=%d cpf=%d" % (T, cpf), lambda i: ops.lt_attn_combine(po, pm, pl, None, None, T, L, cpf, u[i], y[i], mass))
    from rmem_ocu_b200.engine_aot import mha_chunks_per_frame
    cpf = mha_chunks_per_frame(T, L)
    kb2 = [bf(T, L, 256) for _ in range(NB)]
    vb2 = [bf(T, L, 256) for _ in range(NB)]
    q2 = [bf(L, 256) for _ in range(NB)]
    po2 = torch.empty(T * cpf, L, 256, device=dev)
    pm2, pl2 = torch.empty(T * cpf, 8, L, device=dev), torch.empty(T * cpf, 8, L, device=dev)
    timed("mha partial T=%d cpf=%d" % (T, cpf), lambda i: ops.mha_partial(q2[i], kb2[i], vb2[i], order, T, cpf, po2, pm2, pl2))
    timed("mha combine T=%d cpf=%d" % (T, cpf), lambda i: ops.mha_combine(po2, pm2, pl2, None, None, T, L, cpf, o[i], 0, mass))
xf = [f32(L, 1024) for _ in range(NB)]
ws = torch.zeros(65, device=dev, dtype=torch.float64)
g1, b1 = f32(1024), f32(1024)
timed("gn32+gelu tokens (2 kernels)", lambda i: ops.groupnorm_gelu_tokens(xf[i], g1, b1, y[i], 32, ws))
P = 120 * 216
xd = [f32(P, 256) for _ in range(2)]
gd, bd = f32(256), f32(256)
timed("gn nhwc 120x216x256 (2 kernels)", lambda i: ops.groupnorm_nhwc(xd[i % 2], gd, bd, xd[i % 2], 8, True, ws))
