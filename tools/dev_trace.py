"""Dev tool (GPU): CUPTI timeline of steady-state frames through torch.profiler -- per-kernel start / duration /
gap to the previous kernel, to see how much of a frame is launch gaps between CUDA-graph nodes."""
import json
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench  # noqa: E402
from rmem_ocu_b200.engine import DeAOTInferEngine  # noqa: E402

dev = torch.device("cuda", 0)
torch.backends.cudnn.benchmark = True
model = bench.build_model()
n_frames = 40
imgs, lab = bench.make_clip(n_frames, 1000)
eng = DeAOTInferEngine(model, 0, long_term_mem_gap=5)
dimgs = [im.to(dev) for im in imgs]
with torch.no_grad():
    eng.add_reference_frame(dimgs[0], lab.to(dev), obj_nums=[bench.N_OBJ], frame_step=0)
    PF = "--prefetch" in sys.argv
    nxt = (lambda f: dimgs[f + 1]) if PF else (lambda f: None)
    for f in range(1, 30):
        bench.frame_step(eng, dimgs[f], (bench.H, bench.W), nxt(f))
        if f == 1:
            eng.precapture((bench.H, bench.W), prefetch=PF)
    torch.cuda.synchronize()
    from torch.profiler import profile, ProfilerActivity
    with profile(activities=[ProfilerActivity.CUDA, ProfilerActivity.CPU]) as prof:
        for f in range(30, 36):
            bench.frame_step(eng, dimgs[f], (bench.H, bench.W), nxt(f))
        torch.cuda.synchronize()
out = os.path.join(ROOT, "gpurun_out", "trace.json")
prof.export_chrome_trace(out)
ev = [e for e in json.load(open(out))["traceEvents"] if e.get("cat") == "kernel"]
ev.sort(key=lambda e: e["ts"])
# frames: split at pack_image
starts = [i for i, e in enumerate(ev) if "pack_image" in e["name"]]
print("kernels", len(ev), "frame starts", starts)
if len(starts) >= 4:
    a, b = starts[2], starts[3]
    fr = ev[a:b]
    t0 = fr[0]["ts"]
    busy = sum(e["dur"] for e in fr)
    span = ev[b]["ts"] - t0
    print("frame span %.1f us, kernel busy %.1f us, gaps %.1f us, kernels %d" % (span, busy, span - busy, len(fr)))
    prev_end = t0
    for e in fr:
        print("%8.1f %7.2f gap %6.2f  s%-3s %s" % (e["ts"] - t0, e["dur"], e["ts"] - prev_end,
                                                   e.get("args", {}).get("stream", "?"), e["name"][:70]))
        prev_end = e["ts"] + e["dur"]
os.remove(out)
