"""Chain of the path's small kernels (LayerNorm -> GEMM -> dw conv -> GEMM) x10 replayed as one CUDA graph:
time per launch with programmatic dependent launch on (default) and off (RMEM_PDL=0)."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from rmem_ocu_b200 import ops
L, h, w = 1620, 30, 54
dev = "cuda"
tgt = torch.randn(L, 256, device=dev); tid = torch.randn(L, 256, device=dev)
g = torch.ones(256, device=dev); b = torch.zeros(256, device=dev)
xn = torch.empty(L, 512, device=dev, dtype=torch.bfloat16)
w1 = (torch.randn(2048, 512, device=dev) * 0.02).bfloat16(); b1 = torch.zeros(2048, device=dev)
y1 = torch.empty(L, 2048, device=dev, dtype=torch.bfloat16)
dwt = torch.randn(25, 2048, device=dev) * 0.1
y2 = torch.empty(L, 2048, device=dev, dtype=torch.bfloat16)
w2 = (torch.randn(512, 2048, device=dev) * 0.02).bfloat16(); b2 = torch.zeros(512, device=dev)
def chain():
    for _ in range(10):
        ops.layernorm(tgt, g, b, xn, 0, tid, g, b, xn, 256)
        ops.linear(xn, w1, b1, [(0, 2048, ops.ACT_SILU, ops.OUT_BF16, y1, 0)])
        ops.dwconv5(y1, 0, dwt, y2, 0, h, w, 2048)
        ops.linear(y2, w2, b2, [(0, 256, ops.ACT_NONE, ops.OUT_F32_ACCUM, tgt, 0), (256, 512, ops.ACT_NONE, ops.OUT_F32_ACCUM, tid, 0)])
chain(); torch.cuda.synchronize()
ref = tgt.clone()
s = torch.cuda.Stream()
with torch.cuda.stream(s):
    chain(); torch.cuda.synchronize()
    gr = torch.cuda.CUDAGraph()
    with torch.cuda.graph(gr, stream=s):
        chain()
torch.cuda.synchronize()
ts = []
for _ in range(20):
    a = torch.cuda.Event(enable_timing=True); e = torch.cuda.Event(enable_timing=True)
    a.record(); gr.replay(); e.record(); torch.cuda.synchronize(); ts.append(a.elapsed_time(e) * 1e3)
ts.sort()
print("RMEM_PDL=%s  graph of 40 launches: %.1f us  (%.2f us / launch)  finite=%s" % (os.environ.get("RMEM_PDL", "0"), ts[10], ts[10] / 40, bool(torch.isfinite(tgt).all())))
