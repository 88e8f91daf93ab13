"""Pins the CPU oracle (oracle/restate.py) against the golden vectors under tests/golden/, which are
outputs of the UNMODIFIED reference run on CPU by oracle/make_golden.py (the reference has no golden
vectors of its own for this path).  CPU only; no CUDA code is touched here."""
import os

import numpy as np
import pytest
import torch
import torch.nn.functional as F

from oracle import restate as R
from tests import _synth

GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")
RS = _synth.ROW_STRIDE


def _load(name):
    return np.load(os.path.join(GOLDEN, name + ".npz"))


def _t(a):
    return torch.from_numpy(np.asarray(a))


def _close(a, b, tol, what=""):
    a, b = a.float(), _t(b).float()
    err = (a - b).abs().max().item() / max(b.abs().max().item(), 1e-12)
    assert err < tol, (what, err)


# ------------------------------------------------------------------------------------------------
@pytest.mark.parametrize("kind", ["aot", "deaot"])
def test_eviction_rule_matches_reference(kind):
    """restrict_long_memories (transformer.py:324-436 / 880-991): every slot choice over 380 appends."""
    g = _load("evict_" + kind)
    _, sd = _synth.models(4, vos=kind)
    for case, (latter, steps, L) in enumerate(_synth.EVICT_CASES):
        lstt = R.LSTT(sd, kind == "deaot")
        n_t = 2 if kind == "aot" else 4
        lstt.long_term_memories = [[torch.zeros(1, 1, 1, 1) if not (kind == "deaot" and j == 2) else None
                                    for j in range(n_t)] for _ in range(3)]
        idx = [0]
        for step in range(1, steps + 1):
            T_old = len(idx)
            assert T_old == int(g["c%d_T_old" % case][step - 1])
            W, fg = _synth.evict_inputs(kind, case, step, T_old)
            lstt.long_term_memories = [[None if m is None else torch.cat([m, torch.zeros(1, 1, 1, 1)], 0)
                                        for m in layer] for layer in lstt.long_term_memories]
            idx.append(step * 3)
            lstt.record[0] = W
            lstt.restrict_long_memories(1, latter, idx, fg)
            want = [int(v) for v in g["c%d_idx_after" % case][step - 1] if v >= 0]
            assert idx == want, (kind, case, step)


def test_deaot_operators_match_reference():
    """GatedPropagation (explicit + SDPA paths) and LocalGatedPropagation module outputs."""
    g = _load("ops_deaot")
    _, sd = _synth.models(4, vos="deaot")
    p = "LSTT.layers.1"
    for tag in _synth.OPS_DEAOT_CASES:
        i = _synth.ops_deaot_inputs(tag)
        h, w, T, L = i["h"], i["w"], i["T"], i["L"]
        with torch.no_grad():
            o, attn, _ = R.gated_propagation(i["Q"], i["K"], i["V"], i["U"], (h, w), sd, p + ".long_term_attn")
            o2, _, _ = R.gated_propagation(i["X"], i["X"], i["X"], i["X"], (h, w), sd, p + ".self_attn",
                                           use_linear=True)
            o3, la, _ = R.local_gated_propagation(i["q"], i["k"], i["v"], i["U"], (h, w), sd,
                                                  p + ".short_term_attn")
        _close(o[::RS], g["lt_%s_out" % tag], 2e-5, "lt out " + tag)
        _close(R.record_attn_weight(attn, L), g["lt_%s_mass" % tag], 2e-5, "lt mass " + tag)
        _close(o2[::RS], g["sa_%s_out" % tag], 2e-5, "sa out " + tag)
        _close(o3[::RS], g["st_%s_out" % tag], 2e-5, "st out " + tag)
        _close(la[0, 0].t()[::RS], g["st_%s_attn" % tag], 2e-5, "st attn " + tag)


def test_aot_operators_match_reference():
    """MultiheadAttention: explicit path (long-term) and linear + SDPA path (self-attention)."""
    g = _load("ops_aot")
    _, sd = _synth.models(3, vos="aot")
    p = "LSTT.layers.0"
    for tag in _synth.OPS_AOT_CASES:
        i = _synth.ops_aot_inputs(tag)
        L = i["L"]
        with torch.no_grad():
            o, attn = R.multihead_attention(i["Q"], i["K"], i["V"], sd, p + ".long_term_attn", 8, False)
            o2, _ = R.multihead_attention(i["Q"], i["Q"], i["K"][:L], sd, p + ".self_attn", 8, True)
        _close(o[::RS], g["lt_%s_out" % tag], 2e-5, "lt out " + tag)
        _close(R.record_attn_weight(attn, L), g["lt_%s_mass" % tag], 2e-5, "lt mass " + tag)
        _close(o2[::RS], g["sa_%s_out" % tag], 2e-5, "sa out " + tag)


@pytest.mark.parametrize("kind", ["aot", "deaot"])
def test_block_forward_per_bank_length(kind):
    """One GatedPropagationModule / SimplifiedTransformerBlock forward for T = 1..12: pins the temporal
    positional-embedding row selection (transformer.py:594-629 / 1140-1175) and record_attn_weight."""
    g = _load("layer_" + kind)
    _, sd = _synth.models(11, vos=kind)
    lstt = R.LSTT(sd, kind == "deaot")
    tpe = torch.cat((sd["cur_pos_emb"], sd["mem_pos_emb"]), 0) * 20.0
    hw = _synth.LAYER_HW
    for T in range(1, 13):
        i = _synth.layer_inputs(kind, T)
        with torch.no_grad():
            if kind == "deaot":
                o, o_id, _ = lstt._deaot_block(1, i["tgt"], i["tgt_id"], [i["gK"], i["gV"], None, i["gI"]],
                                               [i["lK"], i["lV"], None, i["lI"]], None, hw, tpe, True)
                _close(o_id, g["T%d_out_id" % T], 5e-5, "out_id T=%d" % T)
            else:
                o, mems = lstt._aot_block(1, i["tgt"], [i["gK"], i["gV"]], [i["lK"], i["lV"]], None, i["pos"],
                                          hw, tpe, True)
                _close(mems[2][0], g["T%d_new_local_K" % T], 5e-5, "new_local_K T=%d" % T)
                _close(mems[2][1], g["T%d_new_local_V" % T], 5e-5, "new_local_V T=%d" % T)
        _close(o, g["T%d_out" % T], 5e-5, "out T=%d" % T)
        _close(lstt.record[1], g["T%d_W" % T], 5e-5, "W T=%d" % T)


def test_tpe_row_table():
    """Host-side TPE row table == the table the reference's F.interpolate chain produces."""
    from rmem_ocu_b200.tpe import tpe_row_indices
    want = {1: [0], 2: [0, 1], 3: [0, 1, 2], 4: [0, 1, 2, 3], 5: [0, 1, 2, 3, 3], 6: [0, 1, 1, 2, 3, 3],
            7: [0, 1, 1, 2, 2, 3, 3], 8: [0, 0, 1, 1, 2, 2, 3, 3], 9: [0, 0, 1, 1, 2, 2, 3, 3, 3],
            10: [0, 0, 1, 1, 1, 2, 2, 3, 3, 3], 11: [0, 0, 1, 1, 1, 2, 2, 2, 3, 3, 3],
            12: [0, 0, 0, 1, 1, 1, 2, 2, 2, 3, 3, 3]}
    for T in range(1, 16):
        assert tpe_row_indices(T) == R.tpe_row_indices(T), T
        if T in want:
            assert tpe_row_indices(T) == want[T], T


# ------------------------------------------------------------------------------------------------
def run_oracle_clip(name, g):
    H, W, n, n_obj, latter, gap, repeat, cseed = [int(v) for v in g["meta"]]
    kind = "deaot" if "deaot" in str(g["model"]) else "aot"
    _, sd = _synth.models(latter, vos=kind)
    imgs, lab = _synth.clip(H, W, n, n_obj, cseed, repeat)
    eng = R.Engine(sd, kind == "deaot", 1, latter, long_term_mem_gap=gap)
    masks, idx_log, logits = [], [], {}
    with torch.no_grad():
        eng.add_reference_frame(imgs[0], lab, obj_nums=[10], frame_step=0)
        logits[0] = eng.pred_id_logits[0].clone()
        for f in range(1, n):
            lg = eng.match_propogate_one_frame(imgs[f], output_size=(H, W))
            label = torch.argmax(torch.softmax(lg, 1), 1, keepdim=True).float()
            masks.append(label[0, 0].to(torch.uint8))
            logits[f] = eng.pred_id_logits[0].clone()
            eng.update_short_term_memory(F.interpolate(label, size=(H, W), mode="nearest"))
            idx_log.append(list(eng.long_memories_indexes))
    return masks, idx_log, logits, n


@pytest.mark.parametrize("name", ["clip_deaot_small", "clip_deaot_repeat", "clip_deaot_gap", "clip_deaot_mid",
                                  "clip_deaot_480p", "clip_aot_small", "clip_aot_repeat", "clip_aot_480p"])
def test_engine_clip_matches_reference(name):
    """Whole-engine run (add_reference_frame / match_propogate_one_frame / update_short_term_memory):
    long_memories_indexes identical after every frame, masks >= 99.9 % equal, logits within 1e-3."""
    g = _load(name)
    masks, idx_log, logits, n = run_oracle_clip(name, g)
    for f, idx in enumerate(idx_log):
        want = [int(v) for v in g["indexes"][f] if v >= 0]
        assert idx == want, (name, f + 1, idx, want)
    agree = [(m.numpy() == gm).mean() for m, gm in zip(masks, g["masks"])]
    assert min(agree) >= 0.999, (name, agree)
    _close(logits[0], g["logits_ref_frame"], 1e-3, "logits frame 0")
    _close(logits[1], g["logits_first"], 1e-3, "logits frame 1")
    _close(logits[n - 1], g["logits_last"], 1e-3, "logits last frame")
