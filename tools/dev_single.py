"""Times the one-launch single-frame attention against the partial + combine pair it replaces (L = 1620)."""
import torch
from rmem_ocu_b200 import ops
L = 1620
q = (torch.randn(L, 128, device="cuda")).bfloat16(); v = torch.randn(L, 1024, device="cuda").bfloat16()
g = torch.rand(L, 1024, device="cuda").bfloat16(); out = torch.empty(L, 1024, device="cuda", dtype=torch.bfloat16)
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
print("single %.1f us   pair(cpf 3) %.1f us" % (timeit(lambda: ops.lt_attn_single(q, q, v, g, out)), timeit(pair)))
