"""Self-attention of one frame (L = 1620): one-launch kernel vs first-generation partial + combine vs stream-K at T = 1."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from rmem_ocu_b200 import ops
L = 1620
q = (torch.randn(L, 128, device="cuda")).bfloat16(); v = torch.randn(L, 1024, device="cuda").bfloat16()
g = torch.rand(L, 1024, device="cuda").bfloat16(); out = torch.empty(L, 1024, device="cuda", dtype=torch.bfloat16)
out2 = torch.empty_like(out)
order = torch.zeros(1, device="cuda", dtype=torch.int32)
po = torch.empty(3, L, 1024, device="cuda"); pm = torch.empty(3, L, device="cuda"); pl = torch.empty(3, L, device="cuda")
flush = torch.empty(256 << 20, device="cuda", dtype=torch.uint8)
def timeit(fn, n=30):
    ts = []
    for _ in range(n):
        flush.zero_(); a = torch.cuda.Event(enable_timing=True); b = torch.cuda.Event(enable_timing=True)
        a.record(); fn(); b.record(); torch.cuda.synchronize(); ts.append(a.elapsed_time(b) * 1e3)
    ts.sort(); return ts[len(ts) // 2]
def pair():
    ops.lt_attn_partial(q, q.view(1, L, 128), v.view(1, L, 1024), order, 1, 3, po, pm, pl)
    ops.lt_attn_combine(po, pm, pl, None, None, 1, L, 3, g, out, None)
res = ["single %.1f us" % timeit(lambda: ops.lt_attn_single(q, q, v, g, out)), "pair(cpf 3) %.1f us" % timeit(pair)]
for nc in (0, 56, 64):
    ws = ops.lt_attn_sk_workspace(L, 1, "cuda", nc)
    def sk():
        ops.lt_attn_sk_partial(q, q.view(1, L, 128), v.view(1, L, 1024), order, 1, None, None, False, ws, nc)
        ops.lt_attn_sk_combine(1, L, g, out2, None, ws, nc)
    res.append("stream-K(%d clusters) %.1f us" % (nc, timeit(sk)))
ops.lt_attn_single(q, q, v, g, out); sk(); torch.cuda.synchronize()
print("   ".join(res), "  rel diff %.2e" % ((out.float() - out2.float()).norm() / out.float().norm()).item())
