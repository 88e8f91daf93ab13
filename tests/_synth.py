"""Synthetic inputs shared by the golden-vector generator (oracle/make_golden.py) and the tests
(SURVEY.md section 8d): seeded random frames, optionally the "repeat" family in which frames cycle
through a few base images + noise so the relevance term of the eviction rule matters, and a frame-0
label of axis-aligned rectangles."""
import zlib

import numpy as np
import torch


def clip(H, W, n, n_obj, seed=1000, repeat=0):
    rs = np.random.RandomState(seed)
    rn = lambda: torch.from_numpy(rs.standard_normal(size=(1, 3, H, W)).astype(np.float32))
    if repeat:
        base = [rn() for _ in range(repeat)]
        imgs = [base[f % repeat] + 0.1 * rn() for f in range(n)]
    else:
        imgs = [rn() for _ in range(n)]
    lab = torch.zeros(1, 1, H, W)
    s = H / 480.0
    for o in range(1, n_obj + 1):
        r0, r1 = int((40 + 60 * (o - 1)) * s), int((120 + 60 * (o - 1)) * s)
        c0, c1 = int((100 + 100 * (o - 1)) * s), int((300 + 100 * (o - 1)) * s)
        lab[:, :, r0:r1, c0:c1] = o
    return imgs, lab.int()


def models(latter, seed=0, vos="deaot"):
    """(product model container, plain state_dict copy) with deterministic, name-keyed weights."""
    from rmem_ocu_b200.model import AOTModel, DeAOTModel, RMemConfig, deterministic_init_
    cls = DeAOTModel if vos == "deaot" else AOTModel
    m = cls(RMemConfig(model_vos=vos, latter_mem_len=latter))
    deterministic_init_(m.state_dict(), seed)
    sd = {k: v.clone() for k, v in m.state_dict().items()}
    return m, sd


def nrand(name, *shape, scale=1.0, uniform=False):
    """Deterministic tensor keyed by a name (numpy legacy RandomState: identical on every machine), so
    fixtures store only the reference's OUTPUTS and every side regenerates the inputs."""
    rs = np.random.RandomState(zlib.crc32(name.encode()) % (2 ** 31))
    v = rs.random_sample(size=shape) if uniform else rs.standard_normal(size=shape)
    return torch.from_numpy((v * scale).astype(np.float32))


# ---- inputs of the operator / layer / eviction fixtures (tests/golden/*.npz hold the reference outputs)
OPS_DEAOT_CASES = {"a": (17, 20, 3), "b": (6, 9, 1), "c": (31, 8, 2)}       # tag -> (h, w, T)
OPS_AOT_CASES = {"a": (340, 3), "b": (54, 1), "c": (200, 5)}                # tag -> (L, T)
EVICT_CASES = [(4, 160, 24), (8, 160, 24), (2, 60, 24)]                     # (latter, steps, L)
LAYER_HW = (6, 7)      # L = 42 >= 32: the reference block runs a debug topk(32) over the keys (transformer.py:645)
ROW_STRIDE = 5         # large operator outputs are stored for every 5th token only


def ops_deaot_inputs(tag):
    h, w, T = OPS_DEAOT_CASES[tag]
    L = h * w
    n = "ops_deaot_%s_" % tag
    return dict(h=h, w=w, T=T, L=L,
                Q=nrand(n + "Q", L, 1, 128, scale=1.5), K=nrand(n + "K", T * L, 1, 128),
                V=nrand(n + "V", T * L, 1, 1024), U=nrand(n + "U", L, 1, 1024), X=nrand(n + "X", L, 1, 512),
                q=nrand(n + "q2d", 1, 128, h, w, scale=1.5), k=nrand(n + "k2d", 1, 128, h, w),
                v=nrand(n + "v2d", 1, 1024, h, w))


def ops_aot_inputs(tag):
    L, T = OPS_AOT_CASES[tag]
    n = "ops_aot_%s_" % tag
    return dict(L=L, T=T, Q=nrand(n + "Q", L, 1, 256, scale=1.5), K=nrand(n + "K", T * L, 1, 256),
                V=nrand(n + "V", T * L, 1, 256))


def layer_inputs(kind, T):
    h, w = LAYER_HW
    L = h * w
    n = "layer_%s_T%d_" % (kind, T)
    if kind == "deaot":
        return dict(tgt=nrand(n + "tgt", L, 1, 256), tgt_id=nrand(n + "tgt_id", L, 1, 256),
                    gK=nrand(n + "gK", T, L, 1, 128), gV=nrand(n + "gV", T, L, 1, 512),
                    gI=nrand(n + "gI", T, L, 1, 512), lK=nrand(n + "lK", 1, 128, h, w),
                    lV=nrand(n + "lV", 1, 512, h, w), lI=nrand(n + "lI", 1, 512, h, w))
    return dict(tgt=nrand(n + "tgt", L, 1, 256), pos=nrand(n + "pos", L, 1, 256),
                gK=nrand(n + "gK", T, L, 1, 256), gV=nrand(n + "gV", T, L, 1, 256),
                lK=nrand(n + "lK", L, 1, 256), lV=nrand(n + "lV", L, 1, 256))


def evict_inputs(kind, case, step, T_old):
    """Concentrated, non-uniform per-frame masses + a foreground map for one eviction step."""
    L = EVICT_CASES[case][2]
    n = "evict_%s_c%d_s%d_" % (kind, case, step)
    W = nrand(n + "W", L, T_old, uniform=True) ** 3 + 1e-3
    W = W / W.sum(1, keepdim=True)
    fg = nrand(n + "fg", 1, 1, 4, L // 4, uniform=True)
    return W, fg
