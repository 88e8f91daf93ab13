"""The callers' side of the path: checkpoint ingestion with load_network semantics, the evaluator's
size / gap rules and work queue, the palette PNG writer.  The comparisons with the live reference run in the build
container only; the rest is plain CPU logic."""
import io
import os
import sys

import numpy as np
import pytest
import torch

from oracle import ref_loader as RL
from rmem_ocu_b200 import checkpoint as C
from rmem_ocu_b200 import evaluation as E

needs_ref = pytest.mark.skipif(not RL.available(), reason="reference not present at /root/reference")


# ------------------------------------------------------------------------------------------------ checkpoint
def _toy():
    m = torch.nn.Sequential()
    m.add_module("bank", torch.nn.Conv2d(12, 4, 3))
    m.add_module("lin", torch.nn.Linear(5, 3))
    return m


def test_merge_rule_of_load_network():
    net = _toy()
    init_last = net.state_dict()["bank.weight"][:, -1].clone()
    ck = {"bank.weight": torch.full((4, 11, 3, 3), 2.0),        # one input channel fewer: fills [:, :-1]
          "bank.bias": torch.full((4,), 3.0),                    # equal shape: taken
          "module.lin.weight": torch.full((3, 5), 4.0),          # DataParallel prefix: taken under the stripped name
          "module.lin.bias": torch.zeros(7),                     # prefix but wrong shape: dropped silently
          "other.weight": torch.zeros(2)}                        # unknown: reported
    merged, removed = C.merge_state_dict(net.state_dict(), ck)
    assert removed == ["other.weight"]
    assert torch.all(merged["bank.weight"][:, :-1] == 2.0) and torch.equal(merged["bank.weight"][:, -1], init_last)
    assert torch.all(merged["bank.bias"] == 3.0) and torch.all(merged["lin.weight"] == 4.0)
    assert C.select_state_dict({"state_dict": 1, "model": 2}) == 1 and C.select_state_dict({"model": 2}) == 2
    d = {"a": torch.zeros(1)}
    assert C.select_state_dict(d) is d


@needs_ref
@pytest.mark.parametrize("wrap", ["state_dict", "model", None])
def test_load_network_matches_reference(tmp_path, wrap):
    """A checkpoint written from a reference-built DeAOT model (11-class id bank, module. prefixes, a stray key)
    loads into this package's container exactly as the reference's load_network loads it into the reference's model."""
    sys.path.insert(0, os.path.join(RL.REFERENCE_ROOT, "aot_plus"))
    RL.load()
    from utils.checkpoint import load_network as ref_load
    from rmem_ocu_b200.model import DeAOTModel, RMemConfig, deterministic_init_
    src = RL.build_model(RL.make_cfg("r50_deaotl", 1, 4), seed=5)
    sd = {}
    for i, (k, v) in enumerate(src.state_dict().items()):
        v = v.clone() + 0.25
        if k == "patch_wise_id_bank.weight":
            v = v[:, :-1].contiguous()                           # trained without the ignore channel
        sd[("module." + k) if i % 3 == 0 and k != "patch_wise_id_bank.weight" else k] = v
    sd["optimizer_junk"] = torch.zeros(3)
    path = str(tmp_path / "ck.pth")
    torch.save({wrap: sd} if wrap else sd, path)
    ref_net = RL.build_model(RL.make_cfg("r50_deaotl", 1, 4), seed=9)
    mine = DeAOTModel(RMemConfig(model_vos="deaot"))
    mine.load_state_dict(ref_net.state_dict())                   # same initial values on both sides
    ref_net, ref_removed = ref_load(ref_net, path, -1)
    mine, removed = C.load_network(mine, path, -1)
    assert removed == ref_removed == ["optimizer_junk"]
    a, b = ref_net.state_dict(), mine.state_dict()
    assert a.keys() == b.keys()
    assert all(torch.equal(a[k], b[k]) for k in a)


# ------------------------------------------------------------------------------------------------ sizes, gaps, queue
def test_gap_and_queue():
    assert [E.memory_gap(n) for n in (10, 100, 150, 2000)] == [5, 5, 5, 67]
    assert E.memory_gap(2000, no_memory_gap=True) == 17
    q = E.SequenceQueue(n_workers=1)
    q.fill(5)
    assert list(q.worker(5)) == [0, 1, 2, 3, 4]
    # two workers sharing one queue, stepping alternately: a partition of the indices
    q = E.SequenceQueue(n_workers=2)
    q.fill(7)
    w = [q.worker(7), q.worker(7)]
    got, alive, turn = [[], []], [True, True], 0
    while any(alive):
        if alive[turn]:
            try:
                got[turn].append(next(w[turn]))
            except StopIteration:
                alive[turn] = False
        turn = 1 - turn
    assert sorted(got[0] + got[1]) == list(range(7)) and not set(got[0]) & set(got[1])


@needs_ref
def test_restrict_size_matches_reference():
    sys.path.insert(0, os.path.join(RL.REFERENCE_ROOT, "aot_plus"))
    RL.load()
    import dataloaders.video_transforms as tr
    for (h, w) in [(480, 864), (720, 1280), (1080, 1920), (481, 865), (360, 640), (1000, 30), (853, 480), (1041, 577)]:
        for kw, rkw in (({}, dict(min_size=None, max_size=1040)),
                        (dict(max_long_edge=None, min_short_edge=480), dict(min_size=480, max_size=None)),
                        (dict(align_corners=False), dict(min_size=None, max_size=1040, align_corners=False))):
            t = tr.MultiRestrictSize(flip=False, multi_scale=[1.0], **rkw)
            out = t({"current_img": np.zeros((h, w, 3), np.float32), "meta": {}})[0]["current_img"]
            assert E.restrict_size(h, w, **kw) == tuple(out.shape[:2]), (h, w, kw)


# ------------------------------------------------------------------------------------------------ palette PNGs
def test_palette_structure():
    pal = E.palette()
    assert len(pal) == 768 and pal[:12] == [0, 0, 0, 128, 0, 0, 0, 128, 0, 128, 128, 0]
    assert pal[27:30] == [191, 0, 0] and pal[66:69] == [22, 22, 22] and pal[-3:] == [255, 255, 255]


@needs_ref
def test_mask_png_bytes_match_reference(tmp_path):
    sys.path.insert(0, os.path.join(RL.REFERENCE_ROOT, "aot_plus"))
    RL.load()
    import utils.image as UI
    assert E.palette() == list(UI._palette)
    rs = np.random.RandomState(0)
    for k, (shape, squeeze) in enumerate([((480, 864), None), ((37, 53), [0, 3, 7, 200]), ((8, 8), None)]):
        n_lab = 12 if squeeze is None else len(squeeze)
        mask = rs.randint(0, n_lab, size=shape).astype(np.uint8)
        mask[: shape[0] // 2, : shape[1] // 2] = 0
        a, b = str(tmp_path / ("a%d.png" % k)), str(tmp_path / ("b%d.png" % k))
        UI._save_mask(mask.copy(), a, squeeze)
        E.write_mask_png(mask.copy(), b, squeeze)
        assert open(a, "rb").read() == open(b, "rb").read()
    # the threaded wrapper writes the same file
    t = E.save_mask(torch.from_numpy(mask), str(tmp_path / "c.png"), squeeze)
    t.join()
    assert open(str(tmp_path / "c.png"), "rb").read() == open(b, "rb").read()
