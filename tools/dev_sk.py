"""Dev: time the stream-K long-term attention against the first generation (GPU box)."""
import math, sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from rmem_ocu_b200 import ops

def timed(fn, iters=30, reps=4):
    for i in range(3): fn(i)
    torch.cuda.synchronize()
    g = torch.cuda.CUDAGraph()
    with torch.cuda.graph(g):
        for i in range(iters): fn(i)
    g.replay(); torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(reps): g.replay()
    e1.record(); torch.cuda.synchronize()
    return e0.elapsed_time(e1) / (iters * reps) * 1e3

for (L, T) in ((1620, 5), (1620, 1), (1620, 3), (1620, 9), (3600, 5)):
    nc = 6
    mk = lambda *s: torch.randn(*s, device="cuda").to(torch.bfloat16)
    qs = [mk(L, 128) for _ in range(nc)]; kbs = [mk(T, L, 128) for _ in range(nc)]; vbs = [mk(T, L, 1024) for _ in range(nc)]
    gate = mk(L, 1024); out = torch.empty(L, 1024, device="cuda", dtype=torch.bfloat16)
    order = torch.arange(T, device="cuda", dtype=torch.int32)
    bias = torch.randn(L, 4, device="cuda"); rows = [min(t, 3) for t in range(T)]
    mass = torch.empty(L, T, device="cuda")
    ws = ops.lt_attn_sk_workspace(L, T, "cuda")
    from rmem_ocu_b200.engine import _chunks_per_frame
    cpf = _chunks_per_frame(T, L)
    po = torch.empty(T * cpf, L, 1024, device="cuda"); pm = torch.empty(T * cpf, L, device="cuda"); pl = torch.empty(T * cpf, L, device="cuda")
    t_old_p = timed(lambda i: ops.lt_attn_partial(qs[i % nc], kbs[i % nc], vbs[i % nc], order, T, cpf, po, pm, pl))
    t_old_c = timed(lambda i: ops.lt_attn_combine(po, pm, pl, bias, rows, T, L, cpf, gate, out, mass))
    t_new_p = timed(lambda i: ops.lt_attn_sk_partial(qs[i % nc], kbs[i % nc], vbs[i % nc], order, T, bias, rows, True, ws))
    t_new_c = timed(lambda i: ops.lt_attn_sk_combine(T, L, gate, out, mass, ws))
    fl = 2.0 * L * T * L * 1152
    print("L %d T %d: gen1 partial %.1f us (cpf %d) combine %.1f | stream-K partial %.1f us (%.0f TFLOP/s) combine %.1f"
          % (L, T, t_old_p, cpf, t_old_c, t_new_p, fl / t_new_p / 1e6, t_new_c), flush=True)
