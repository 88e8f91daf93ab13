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
n_frames = 44
la = 2
imgs, lab = bench.make_clip(n_frames + la, 1000)
eng = DeAOTInferEngine(model, 0, long_term_mem_gap=5)
clip = torch.cat(imgs).to(dev)
with torch.no_grad():
    eng.add_reference_frame(clip[0:1], lab.to(dev), obj_nums=[bench.N_OBJ], frame_step=0)
    PF = "--no-prefetch" not in sys.argv
    nxt = (lambda f: clip[f + 1:f + 1 + la] if eng.frames_ahead() <= 1 else None) if PF else (lambda f: None)
    for f in range(1, 32):
        bench.frame_step(eng, clip[f:f + 1], (bench.H, bench.W), nxt(f))
        if f == 1:
            eng.precapture((bench.H, bench.W), prefetch=PF, lookahead=la)
    torch.cuda.synchronize()
    from torch.profiler import profile, ProfilerActivity
    with profile(activities=[ProfilerActivity.CUDA, ProfilerActivity.CPU]) as prof:
        for f in range(32, 40):
            bench.frame_step(eng, clip[f:f + 1], (bench.H, bench.W), nxt(f))
        torch.cuda.synchronize()
out = os.path.join(ROOT, "gpurun_out", "trace.json")
prof.export_chrome_trace(out)
ev = [e for e in json.load(open(out))["traceEvents"] if e.get("cat") == "kernel"]
ev.sort(key=lambda e: e["ts"])
# frames: split at the device frame counter's tick (first kernel of a frame's read path)
starts = [i for i, e in enumerate(ev) if "bank_tick" in e["name"]]
print("kernels", len(ev), "frame starts", starts)
for k in range(2, len(starts) - 1):
    a, b = starts[k], starts[k + 1]
    fr = ev[a:b]
    alone = not any("pack_image" in e["name"] for e in fr)
    t0 = fr[0]["ts"]
    span = ev[b]["ts"] - t0
    print("frame %d: span %.1f us, %d kernels, %s" % (k, span, len(fr), "alone" if alone else "beside an encoder pass"))
    if alone and "--all" not in sys.argv and k >= 3:
        busy = sum(e["dur"] for e in fr)
        print("kernel busy %.1f us (sum over streams)" % busy)
        prev_end = {}
        last_end = t0
        for e in fr:
            st = e.get("args", {}).get("stream", "?")
            print("%8.1f %7.2f gap %6.2f  s%-3s %s" % (e["ts"] - t0, e["dur"], e["ts"] - last_end, st, e["name"][:64]))
            last_end = max(last_end, e["ts"] + e["dur"])
        break
os.remove(out)
