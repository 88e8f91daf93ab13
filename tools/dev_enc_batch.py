"""Encoder graph (ResNet-50 trunk + projector + decoder adapters) at batch 1, 2, 4 frames: time per frame."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from rmem_ocu_b200.engine import prepare
from rmem_ocu_b200.model import DeAOTModel, RMemConfig, deterministic_init_
m = DeAOTModel(RMemConfig()); deterministic_init_(m.state_dict(), 0)
dev = torch.device("cuda", 0)
prep = prepare(m, dev)
H, W = 480, 864
for B in (1, 2, 3, 4):
    img = torch.randint(0, 256, (B, H, W, 3), dtype=torch.uint8, device=dev)
    def body():
        embs = list(prep.encoder(img))
        return embs + prep.decoder.adapters(embs)
    with torch.no_grad():
        for _ in range(3): body()
        torch.cuda.synchronize()
        g = torch.cuda.CUDAGraph()
        with torch.cuda.graph(g):
            out = body()
        torch.cuda.synchronize()
        for _ in range(5): g.replay()
        a = torch.cuda.Event(enable_timing=True); b = torch.cuda.Event(enable_timing=True)
        a.record()
        for _ in range(50): g.replay()
        b.record(); torch.cuda.synchronize()
        t = a.elapsed_time(b) * 1e3 / 50
    print("batch %d: %.0f us per pass, %.0f us per frame" % (B, t, t / B))
