"""Per-frame device time by position inside an encoder look-ahead stack (la = 6): the frames that run beside the
batched encoder pass against the frames that run alone (= the LSTT + decoder chain by itself)."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import bench as B
from rmem_ocu_b200.engine import DeAOTInferEngine
la = int(sys.argv[1]) if len(sys.argv) > 1 else 6
dev = torch.device("cuda", 0)
model = B.build_model()
gap = B.gap_rule(100)
head = gap * B.LATTER + 12
steps = 60
n_frames = 1 + head + steps
limit = 2 + la * (-(-(n_frames - 1) // la))
imgs, lab = B.make_clip(limit, 1000, pinned=False)
clip = torch.cat(imgs).to(dev)
eng = DeAOTInferEngine(model, 0, long_term_mem_gap=gap)
ev = []
with torch.no_grad():
    eng.add_reference_frame(clip[0:1], lab.to(dev), obj_nums=[B.N_OBJ], frame_step=0)
    for f in range(1, n_frames):
        nxt = clip[f + 1:f + 1 + la] if eng.frames_ahead() <= 1 and f + 1 < limit else None
        e = torch.cuda.Event(enable_timing=True); e.record(); ev.append((f, nxt is not None, e))
        B.frame_step(eng, clip[f:f + 1], (B.H, B.W), nxt)
        if f == 1:
            eng.precapture((B.H, B.W), prefetch=True, lookahead=la)
    e = torch.cuda.Event(enable_timing=True); e.record(); ev.append((n_frames, False, e))
torch.cuda.synchronize()
with_enc, alone = [], []
for (f, launched, a), (_, _, b) in zip(ev[head:-1], ev[head + 1:]):
    (with_enc if launched else alone).append(a.elapsed_time(b) * 1e3)
import statistics as st
print("la=%d  frames launching an encoder stack: %d, median %.0f us;  frames running alone: %d, median %.0f us" % (
    la, len(with_enc), st.median(with_enc), len(alone), st.median(alone) if alone else float("nan")))
