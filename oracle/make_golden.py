"""TEST INFRASTRUCTURE ONLY -- generates the committed golden fixtures under tests/golden/ by running
the UNMODIFIED reference (Bardli/RMem_ocu, imported from /root/reference through oracle/ref_loader.py)
on CPU in the build container.  The reference ships no golden vectors for this path (SURVEY.md
section 4), so these outputs of the reference itself are what pins oracle/restate.py and, on the GPU
box (where /root/reference does not exist), what the CUDA path is compared with.

    python -m oracle.make_golden [--only NAME ...]

Fixtures (all .npz, inputs included so nothing depends on RNG reproducibility across machines):
  evict_{aot,deaot}.npz   restrict_long_memories driven directly over long synthetic histories
  ops_deaot.npz           GatedPropagation / LocalGatedPropagation module outputs
  ops_aot.npz             MultiheadAttention module outputs (explicit + SDPA paths)
  layer_{aot,deaot}.npz   one full block forward per bank length T = 1..12 (TPE row selection,
                          record_attn_weight)
  clip_*.npz              whole-engine runs on synthetic clips: masks, long_memories_indexes per
                          frame, logits of two frames
Weights are always ``deterministic_init_`` (name-keyed, numpy legacy RNG) so every side can rebuild them.
"""
import argparse
import os
import sys

import numpy as np
import torch
import torch.nn.functional as F

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

from oracle import ref_loader as RL  # noqa: E402
from rmem_ocu_b200.model import deterministic_init_  # noqa: E402
from tests import _synth  # noqa: E402

GOLDEN = os.path.join(ROOT, "tests", "golden")

# name -> (model, H, W, n_frames, n_obj, latter, gap, repeat, clip_seed)
CLIPS = {
    "clip_deaot_small": ("r50_deaotl", 128, 192, 14, 3, 2, 1, 0, 1000),
    "clip_deaot_repeat": ("r50_deaotl", 128, 192, 16, 2, 3, 1, 4, 1001),
    "clip_deaot_gap": ("r50_deaotl", 128, 192, 16, 2, 2, 3, 0, 1002),
    "clip_deaot_mid": ("r50_deaotl", 240, 432, 9, 3, 2, 1, 0, 1003),
    "clip_deaot_480p": ("r50_deaotl", 480, 864, 8, 5, 4, 1, 0, 1000),
    "clip_aot_small": ("r50_aotl", 128, 192, 14, 1, 3, 1, 0, 1012),
    "clip_aot_repeat": ("r50_aotl", 128, 192, 16, 2, 2, 1, 4, 1005),
    "clip_aot_480p": ("r50_aotl", 480, 864, 10, 1, 3, 1, 0, 1000),
}


# free-running long clips (BASELINE.json configs 1 and 3 regimes): masks kept for every `mask_every`-th frame,
# bank indexes after EVERY frame.  name -> (model, H, W, n_frames, n_obj, latter, gap, repeat, clip_seed, mask_every)
#   clip_deaot_cfg2   : config 2 as written -- 480x864, 5 objects, K = 4, 100 frames, evaluator gap rule (5)
#   clip_deaot_long   : config 4's regime -- K = 8 (eval_vost.sh:28), gap 67, 1150 frames at reduced resolution
#                       (17 appends, 9 evictions)
LONG_CLIPS = {
    "clip_deaot_cfg2": ("r50_deaotl", 480, 864, 100, 5, 4, 5, 0, 1000, 10),
    "clip_deaot_long": ("r50_deaotl", 160, 256, 1150, 3, 8, 67, 0, 1007, 50),
}


def _np(t):
    return t.detach().cpu().numpy()


def ref_model(name, latter, seed=0):
    cfg = RL.make_cfg(name, 1, latter)
    model = RL.build_model(cfg, seed)
    deterministic_init_(model.state_dict(), seed)
    return cfg, model


# ------------------------------------------------------------------------------------------------
def gen_clip(name):
    mask_every = 1
    if name in LONG_CLIPS:
        mname, H, W, n, n_obj, latter, gap, repeat, cseed, mask_every = LONG_CLIPS[name]
    else:
        mname, H, W, n, n_obj, latter, gap, repeat, cseed = CLIPS[name]
    cfg, model = ref_model(mname, latter)
    imgs, lab = _synth.clip(H, W, n, n_obj, cseed, repeat)
    eng = RL.build_engine(model, cfg, long_term_mem_gap=gap)
    masks, idx_log, logits = [], [], {}
    with torch.no_grad(), RL.quiet():
        eng.add_reference_frame(imgs[0], lab, obj_nums=[n_obj], frame_step=0)
        logits[0] = _np(eng.aot_engines[0].pred_id_logits[0])
        for f in range(1, n):
            lg = eng.match_propogate_one_frame(imgs[f], output_size=(H, W))
            prob = torch.softmax(lg, dim=1)
            label = torch.argmax(prob, dim=1, keepdim=True).float()
            if f % mask_every == 0 or f == n - 1:
                masks.append(_np(label[0, 0].to(torch.uint8)))
            if f in (1, n - 1):
                logits[f] = _np(eng.aot_engines[0].pred_id_logits[0])
            eng.update_memory(F.interpolate(label, size=eng.input_size_2d, mode="nearest"))
            idx_log.append(list(eng.aot_engines[0].long_memories_indexes))
    width = max(len(i) for i in idx_log)
    idx_arr = np.full((len(idx_log), width), -1, np.int32)
    for r, i in enumerate(idx_log):
        idx_arr[r, :len(i)] = i
    out = dict(meta=np.array([H, W, n, n_obj, latter, gap, repeat, cseed], np.int64),
               model=np.array(mname), masks=np.stack(masks), indexes=idx_arr,
               logits_ref_frame=logits[0], logits_first=logits[1], logits_last=logits[n - 1])
    if name in LONG_CLIPS:
        out["mask_frames"] = np.array([f for f in range(1, n) if f % mask_every == 0 or f == n - 1], np.int32)
    out["margins"] = oracle_margins(name, idx_log)
    return out


def oracle_margins(name, ref_idx_log):
    """Decision margin (runner-up score minus the evicted frame's score) of every eviction of the clip, from the
    oracle restatement run on the same inputs (the reference does not expose it); inf where nothing was scored.
    Doubles as a cross-check: the oracle's bank indexes must equal the reference's after every frame.  The GPU
    tests use it to assert that a fixture has no near-tie a bf16 kernel could not resolve."""
    from oracle import restate as R
    mname, H, W, n, n_obj, latter, gap, repeat, cseed = (LONG_CLIPS[name][:9] if name in LONG_CLIPS else CLIPS[name])
    kind = "deaot" if "deaot" in mname else "aot"
    _, sd = _synth.models(latter, vos=kind)
    imgs, lab = _synth.clip(H, W, n, n_obj, cseed, repeat)
    eng = R.Engine(sd, kind == "deaot", 1, latter, long_term_mem_gap=gap)
    margins = np.full(n - 1, np.inf, np.float32)
    with torch.no_grad():
        eng.add_reference_frame(imgs[0], lab, obj_nums=[10], frame_step=0)
        for f in range(1, n):
            lg = eng.match_propogate_one_frame(imgs[f], output_size=(H, W))
            label = torch.argmax(torch.softmax(lg, 1), 1, keepdim=True).float()
            n_log = len(eng.evict_log)
            eng.update_short_term_memory(F.interpolate(label, size=(H, W), mode="nearest"))
            if len(eng.evict_log) > n_log and eng.evict_log[-1]["dropped"]:
                margins[f - 1] = eng.evict_log[-1]["margin"]
            assert list(eng.long_memories_indexes) == list(ref_idx_log[f - 1]), (name, f)
    return margins


# ------------------------------------------------------------------------------------------------
def gen_evict(kind):
    """Drive the reference restrict_long_memories directly (transformer.py:324-436 / 880-991)."""
    mname = "r50_aotl" if kind == "aot" else "r50_deaotl"
    out = {}
    for case, (cap_latter, steps, L) in enumerate(_synth.EVICT_CASES):
        cfg, model = ref_model(mname, cap_latter)
        lstt = model.LSTT
        n_t = 2 if kind == "aot" else 4
        lstt.long_term_memories = [[torch.zeros(1, 1, 1, 1) if not (kind == "deaot" and j == 2) else None
                                    for j in range(n_t)] for _ in lstt.layers]
        lstt.stored_attn_weight_dict = {}
        lstt.stored_frame_times = {}
        idx = [0]
        idx_after, T_olds = [], []
        for step in range(1, steps + 1):
            T_old = len(idx)
            W, fg = _synth.evict_inputs(kind, case, step, T_old)
            # append (update_long_term_memory) then restrict, as aot_engine.py:350-369
            lstt.long_term_memories = [[None if m is None else torch.cat([m, torch.zeros(1, 1, 1, 1)], 0)
                                        for m in layer] for layer in lstt.long_term_memories]
            idx.append(step * 3)
            lstt.layers[0].record_attn_weight = W.clone()
            with RL.quiet():
                lstt.restrict_long_memories(former_memory_len=1, latter_memory_len=cap_latter,
                                            use_atten_weight=True, long_memories_indexes=idx,
                                            foreground_proba=fg)
            T_olds.append(T_old)
            ia = np.full(16, -1, np.int32)
            ia[:len(idx)] = idx
            idx_after.append(ia)
        out["c%d_T_old" % case] = np.array(T_olds, np.int32)
        out["c%d_idx_after" % case] = np.stack(idx_after)
    return out


# ------------------------------------------------------------------------------------------------
def gen_ops_deaot():
    cfg, model = ref_model("r50_deaotl", 4)
    layer = model.LSTT.layers[1]
    out = {}
    rs = _synth.ROW_STRIDE
    for tag in _synth.OPS_DEAOT_CASES:
        i = _synth.ops_deaot_inputs(tag)
        h, w, T, L = i["h"], i["w"], i["T"], i["L"]
        with torch.no_grad():
            o, attn = layer.long_term_attn(i["Q"], i["K"], i["V"], i["U"], (h, w), is_return_attn_weight=True)
            o2, _ = layer.self_attn(i["X"], i["X"], i["X"], i["X"], (h, w))      # use_linear=True, SDPA path
            o3, la = layer.short_term_attn(i["q"], i["k"], i["v"], i["U"], (h, w))
        out["lt_%s_out" % tag] = _np(o[::rs])
        out["lt_%s_mass" % tag] = _np(attn.view(1, 1, L, T, L).sum(-1)[0, 0])     # [L, T]
        out["sa_%s_out" % tag] = _np(o2[::rs])
        out["st_%s_out" % tag] = _np(o3[::rs])
        out["st_%s_attn" % tag] = _np(la[0, 0].t()[::rs])                         # [L/rs, 225]
    return out


def gen_ops_aot():
    cfg, model = ref_model("r50_aotl", 3)
    layer = model.LSTT.layers[0]
    out = {}
    rs = _synth.ROW_STRIDE
    for tag in _synth.OPS_AOT_CASES:
        i = _synth.ops_aot_inputs(tag)
        L, T = i["L"], i["T"]
        with torch.no_grad():
            o, attn = layer.long_term_attn(i["Q"], i["K"], i["V"], is_return_attn_weight=True)   # explicit path
            o2, _ = layer.self_attn(i["Q"], i["Q"], i["K"][:L])                                  # linear + SDPA
        out["lt_%s_out" % tag] = _np(o[::rs])
        out["lt_%s_mass" % tag] = _np(attn.mean(1).view(1, L, T, L).sum(-1)[0])                   # [L, T]
        out["sa_%s_out" % tag] = _np(o2[::rs])
    return out


def gen_layer(kind):
    """One block forward per bank length T (covers TPE resampling for T = 1..12)."""
    mname = "r50_aotl" if kind == "aot" else "r50_deaotl"
    cfg, model = ref_model(mname, 11)
    layer = model.LSTT.layers[1]
    h, w = _synth.LAYER_HW
    tpe = torch.cat((model.cur_pos_emb, model.mem_pos_emb), dim=0).detach() * 20.0   # make the rows matter
    out = {}
    for T in range(1, 13):
        i = _synth.layer_inputs(kind, T)
        with torch.no_grad():
            if kind == "deaot":
                o, o_id, _ = layer(i["tgt"], i["tgt_id"], [i["gK"], i["gV"], None, i["gI"]],
                                   [i["lK"], i["lV"], None, i["lI"]], curr_id_emb=None, self_pos=None,
                                   size_2d=(h, w), temporal_encoding=tpe, save_atten_weights=True)
                res = dict(out=o, out_id=o_id, W=layer.record_attn_weight)
            else:
                o, mems = layer(i["tgt"], [i["gK"], i["gV"]], [i["lK"], i["lV"]], curr_id_emb=None,
                                self_pos=i["pos"], size_2d=(h, w), temporal_encoding=tpe, save_atten_weights=True)
                res = dict(out=o, W=layer.record_attn_weight, new_local_K=mems[2][0], new_local_V=mems[2][1])
        for nme, t in res.items():
            out["T%d_%s" % (T, nme)] = _np(t)
    return out


GENERATORS = {"evict_aot": lambda: gen_evict("aot"), "evict_deaot": lambda: gen_evict("deaot"),
              "ops_deaot": gen_ops_deaot, "ops_aot": gen_ops_aot,
              "layer_aot": lambda: gen_layer("aot"), "layer_deaot": lambda: gen_layer("deaot")}
GENERATORS.update({k: (lambda k=k: gen_clip(k)) for k in CLIPS})
GENERATORS.update({k: (lambda k=k: gen_clip(k)) for k in LONG_CLIPS})


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--only", nargs="*")
    args = ap.parse_args()
    if not RL.available():
        raise SystemExit("the reference is not present at %s" % RL.REFERENCE_ROOT)
    os.makedirs(GOLDEN, exist_ok=True)
    torch.set_num_threads(os.cpu_count() or 1)
    for name, fn in GENERATORS.items():
        if args.only and name not in args.only:
            continue
        data = fn()
        path = os.path.join(GOLDEN, name + ".npz")
        np.savez_compressed(path, **data)
        print("%-22s %8.1f KB" % (name, os.path.getsize(path) / 1024), flush=True)


if __name__ == "__main__":
    main()
