"""Times the two local-attention kernels separately, S = 1, 2, 4, 8 stacked sequences at 30 x 54 tokens, rotating over
working sets larger than L2 (buffers x nbuf > 126 MB)."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from rmem_ocu_b200 import ops
h, w = 30, 54
L = h * w
dev = "cuda"
for S in (1, 2, 4, 8):
    nb = max(2, 160 // (11 * S))
    q = [torch.randn(S * L, 128, device=dev).bfloat16() for _ in range(nb)]
    k = [torch.randn(S * L, 128, device=dev).bfloat16() for _ in range(nb)]
    v = [torch.randn(S * L, 1024, device=dev).bfloat16() for _ in range(nb)]
    g = [torch.rand(S * L, 1024, device=dev).bfloat16() for _ in range(nb)]
    o = [torch.empty(S * L, 1024, device=dev, dtype=torch.bfloat16) for _ in range(nb)]
    rel = [torch.randn(S * L, 232, device=dev) for _ in range(nb)]
    band = [ops.local_banded_buffer(h, w, dev, S) for _ in range(nb)]
    def run(which, n):
        for i in range(n):
            j = i % nb
            if which != "agg":
                ops.local_attn_scores(q[j], k[j], rel[j], None, band[j], h, w, S)
            if which != "sc":
                ops.local_attn_aggregate(band[j], v[j], g[j], o[j], h, w, S)
    res = {}
    for which in ("sc", "agg", "pair"):
        run(which, nb); torch.cuda.synchronize()
        n = 4 * nb
        a = torch.cuda.Event(enable_timing=True); b = torch.cuda.Event(enable_timing=True)
        a.record(); run(which, n); b.record(); torch.cuda.synchronize()
        res[which] = a.elapsed_time(b) * 1e3 / n
    bytes_alg = S * (2 * L * 128 * 2 + 3 * L * 1024 * 2 + L * 225 * 4)
    print("S=%d  scores %.1f us  aggregate %.1f us  pair %.1f us  -> %.0f GB/s algorithmic" % (S, res["sc"], res["agg"], res["pair"], bytes_alg / res["pair"] / 1e3))
