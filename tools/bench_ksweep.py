"""BASELINE.json config 2: long-term attention roofline study over the bank size, K = LATTER_MEM_LEN in {2, 4, 8}
(steady-state bank length T = K + 1 = 3 / 5 / 9) at 480p (L = 1620) and 720p (L = 3600), for both families:
AOT-L multi-head attention (8 x 32; csrc/mha.cu, SURVEY 8a row A1, FLOPs 4 L (T L) 256) and DeAOT-L gated attention
(d_att 128, value width 1024; csrc/lt_attn.cu, row A4, FLOPs 2 L (T L) (128 + 1024)).  CUDA events around
graph-captured launches on rotating (> L2) inputs; one JSON line per case on stdout."""
import json
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from rmem_ocu_b200 import ops  # noqa: E402
from rmem_ocu_b200.engine import _chunks_per_frame  # noqa: E402
from rmem_ocu_b200.engine_aot import mha_chunks_per_frame  # noqa: E402

dev = "cuda"
NB = 6
peak = 1590.0
pk = os.path.join(ROOT, "MEASURED_PEAKS.json")
if os.path.exists(pk):
    peak = json.load(open(pk))["bf16_tflops"]
bf = lambda *s: torch.randn(*s, device=dev).to(torch.bfloat16)


def timed(fn, iters=20):
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
    return e0.elapsed_time(e1) / (5 * iters) * 1e-3


for L in (1620, 3600):
    for K in (2, 4, 8):
        T = K + 1
        order = torch.arange(T, device=dev, dtype=torch.int32)
        # DeAOT-L
        cpf = _chunks_per_frame(T, L)
        q = [bf(L, 128) for _ in range(NB)]
        kb = [bf(T, L, 128) for _ in range(NB)]
        vb = [bf(T, L, 1024) for _ in range(NB)]
        po = torch.empty(T * cpf, L, 1024, device=dev)
        pm, pl = torch.empty(T * cpf, L, device=dev), torch.empty(T * cpf, L, device=dev)
        t = timed(lambda i: ops.lt_attn_partial(q[i], kb[i], vb[i], order, T, cpf, po, pm, pl))
        fl = 2.0 * L * T * L * (128 + 1024)
        print(json.dumps({"family": "deaot", "kernel": "lt_attn_partial_kernel", "L": L, "K": K, "T": T, "cpf": cpf,
                          "us": round(t * 1e6, 2), "tflops": round(fl / t / 1e12, 1),
                          "frac_of_peak": round(fl / t / 1e12 / peak, 3), "peak_tflops": peak}), flush=True)
        del q, kb, vb, po
        # AOT-L
        cpf = mha_chunks_per_frame(T, L)
        q2 = [bf(L, 256) for _ in range(NB)]
        kb2 = [bf(T, L, 256) for _ in range(NB)]
        vb2 = [bf(T, L, 256) for _ in range(NB)]
        po2 = torch.empty(T * cpf, L, 256, device=dev)
        pm2, pl2 = torch.empty(T * cpf, 8, L, device=dev), torch.empty(T * cpf, 8, L, device=dev)
        t = timed(lambda i: ops.mha_partial(q2[i], kb2[i], vb2[i], order, T, cpf, po2, pm2, pl2))
        fl = 4.0 * L * T * L * 256
        print(json.dumps({"family": "aot", "kernel": "mha_partial_kernel", "L": L, "K": K, "T": T, "cpf": cpf,
                          "us": round(t * 1e6, 2), "tflops": round(fl / t / 1e12, 1),
                          "frac_of_peak": round(fl / t / 1e12 / peak, 3), "peak_tflops": peak,
                          "note": "head dim 32: exp-bound (MUFU), 16 FLOP per exponential"}), flush=True)
        del q2, kb2, vb2, po2
