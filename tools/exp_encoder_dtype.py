"""Experiment (GPU box): mask agreement of the drop-in engine with the reference fixtures as a function of the
cuDNN backbone's arithmetic type.  python tools/exp_encoder_dtype.py [clip names...]"""
import os
import sys

import numpy as np
import torch
import torch.nn.functional as F

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from tests import _synth  # noqa: E402

GOLDEN = os.path.join(ROOT, "tests", "golden")


def run(name, dtype, extra_env=None):
    from rmem_ocu_b200.engine import build_engine
    os.environ["RMEM_ENCODER_DTYPE"] = dtype
    for k, v in (extra_env or {}).items():
        os.environ[k] = v
    g = np.load(os.path.join(GOLDEN, name + ".npz"))
    H, W, n, n_obj, latter, gap, repeat, cseed = [int(v) for v in g["meta"]]
    m, _ = _synth.models(latter, vos="deaot")
    imgs, lab = _synth.clip(H, W, n, n_obj, cseed, repeat)
    frames = [int(f) for f in g["mask_frames"]] if "mask_frames" in g.files else list(range(1, n))
    eng = build_engine("deaotengine", phase="eval", aot_model=m, gpu_id=0, long_term_mem_gap=gap)
    agree, idx_ok = [], True
    with torch.no_grad():
        eng.add_reference_frame(imgs[0].cuda(), lab.cuda(), obj_nums=[n_obj], frame_step=0)
        for f in range(1, n):
            lg = eng.match_propogate_one_frame(imgs[f].cuda(), output_size=(H, W))
            label = torch.argmax(torch.softmax(lg, 1), 1, keepdim=True).float()
            if f in frames:
                gm = g["masks"][frames.index(f)]
                agree.append(float((label[0, 0].to(torch.uint8).cpu().numpy() == gm).mean()))
            eng.update_memory(F.interpolate(label, size=eng.input_size_2d, mode="nearest"))
            if f % 50 == 0 or f == n - 1:
                want = [int(v) for v in g["indexes"][f - 1] if v >= 0]
                idx_ok = idx_ok and list(eng.aot_engines[0].long_memories_indexes) == want
    return min(agree), sum(agree) / len(agree), idx_ok


if __name__ == "__main__":
    names = sys.argv[1:] or ["clip_deaot_mid", "clip_deaot_480p", "clip_deaot_small"]
    for name in names:
        for dt in ("fp32", "fp16", "bf16"):
            mn, avg, ok = run(name, dt)
            print("%-18s encoder %-5s min agreement %.5f mean %.5f indexes_ok %s" % (name, dt, mn, avg, ok), flush=True)
