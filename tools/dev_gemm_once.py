"""Run a few launches of one GEMM shape (for ncu captures): python tools/dev_gemm_once.py M N K mode"""
import sys
import torch
sys.path.insert(0, ".")
from rmem_ocu_b200 import ops
M, N, K, mode = int(sys.argv[1]), int(sys.argv[2]), int(sys.argv[3]), int(sys.argv[4])
a = torch.randn(M, K, device="cuda").bfloat16()
w = torch.randn(N, K, device="cuda").bfloat16()
b = torch.randn(N, device="cuda")
d = torch.zeros(M, N, device="cuda", dtype=torch.bfloat16 if mode == 0 else torch.float32)
for _ in range(6):
    ops.linear(a, w, b, [(0, N, 0, mode, d, 0)])
torch.cuda.synchronize()
ref = a.float() @ w.float().t() + b
print("ok", (d.float() - (ref * (6 if mode == 2 else 1))).abs().max().item())
