"""Kernel durations of the decoder's last stage (statistics, head, resize + argmax) at 480p, from torch.profiler."""
import os, sys, json
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from rmem_ocu_b200 import ops
from torch.profiler import profile, ProfilerActivity
C, Hs, Ws, Ho, Wo, Co = 128, 120, 216, 480, 864, 11
x = torch.randn(1, C, Hs, Ws, device="cuda").half().contiguous(memory_format=torch.channels_last)
pb, g, b = (torch.randn(C, device="cuda") for _ in range(3))
w = torch.randn(Co, C, device="cuda") * 0.2; bo = torch.randn(Co, device="cuda")
ws = torch.zeros(1, 17, device="cuda", dtype=torch.float64)
pred = torch.empty(1, Co, Hs, Ws, device="cuda"); out = torch.empty(1, Co, Ho, Wo, device="cuda"); label = torch.empty(1, 1, Ho, Wo, device="cuda")
for _ in range(5): ops.logits_head(x, pb, g, b, w, bo, [6], pred, out, label, ws)
torch.cuda.synchronize()
with profile(activities=[ProfilerActivity.CUDA]) as prof:
    for _ in range(10): ops.logits_head(x, pb, g, b, w, bo, [6], pred, out, label, ws)
    torch.cuda.synchronize()
p = "/tmp/head_trace.json"; prof.export_chrome_trace(p)
ev = [e for e in json.load(open(p))["traceEvents"] if e.get("cat") == "kernel"]
by = {}
for e in ev: by.setdefault(e["name"][:50], []).append(e["dur"])
for k, v in by.items(): print("%-52s n=%d median %.2f us" % (k, len(v), sorted(v)[len(v) // 2]))
