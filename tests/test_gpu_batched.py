"""Sequence-batched launches (grid dimension = sequence): every batched operator on a stack of S independent
sequences must give what S single-sequence launches give -- bit-exact where the arithmetic order is the same
(depth-wise conv, local attention), within the operator's usual tolerance where the work partition changes with the
batch (stream-K attention: other segment boundaries; GroupNorm: order of the fp64 atomics)."""
import math

import pytest
import torch

pytestmark = pytest.mark.gpu


def _bf(x):
    return x.to(torch.bfloat16)


def _rel(a, b):
    return ((a.float() - b.float()).norm() / b.float().norm().clamp_min(1e-12)).item()


@pytest.mark.parametrize("h,w,S", [(30, 54, 3), (9, 20, 2), (45, 80, 2)])
def test_batched_dwconv_and_local_attention(lib_built, h, w, S):
    from rmem_ocu_b200 import ops
    torch.manual_seed(h * w + S)
    L = h * w
    dev = "cuda"
    # depth-wise 5x5
    x = _bf(torch.randn(S * L, 1024, device=dev))
    wt = torch.randn(25, 1024, device=dev)
    y = torch.zeros(S * L, 2048, device=dev, dtype=torch.bfloat16)
    ops.dwconv5(x, 0, wt, y, 1024, h, w, 1024, S=S)
    for s in range(S):
        y1 = torch.zeros(L, 2048, device=dev, dtype=torch.bfloat16)
        ops.dwconv5(x[s * L:(s + 1) * L], 0, wt, y1, 1024, h, w, 1024)
        assert torch.equal(y[s * L:(s + 1) * L], y1), s
    # local attention: scores (+ probabilities) and aggregation
    q, k = _bf(torch.randn(S * L, 128, device=dev) * 1.5), _bf(torch.randn(S * L, 128, device=dev))
    v, u = _bf(torch.randn(S * L, 1024, device=dev)), _bf(torch.randn(S * L, 1024, device=dev))
    rel = torch.randn(S * L, 256, device=dev) * 0.3
    probs = torch.empty(S * L, 225, device=dev)
    banded = ops.local_banded_buffer(h, w, dev, S=S)
    out = torch.empty(S * L, 1024, device=dev, dtype=torch.bfloat16)
    ops.local_attn_scores(q, k, rel, probs, banded, h, w, S=S)
    ops.local_attn_aggregate(banded, v, u, out, h, w, S=S)
    for s in range(S):
        sl = slice(s * L, (s + 1) * L)
        p1 = torch.empty(L, 225, device=dev)
        b1 = ops.local_banded_buffer(h, w, dev)
        o1 = torch.empty(L, 1024, device=dev, dtype=torch.bfloat16)
        ops.local_attn_scores(q[sl], k[sl], rel[sl], p1, b1, h, w)
        ops.local_attn_aggregate(b1, v[sl], u[sl], o1, h, w)
        assert torch.equal(probs[sl], p1), s
        assert torch.equal(out[sl], o1), s


@pytest.mark.parametrize("L,T,S", [(1620, 5, 2), (1620, 1, 3), (333, 9, 3), (3600, 2, 2)])
def test_batched_streamk_attention(lib_built, L, T, S):
    """S sequences with their own banks and their own slot orders in one launch, against the fp32 evaluation of each
    sequence on its own (the oracle of test_gpu_attention.py): 2e-2 on the output, 5e-3 on the masses."""
    from rmem_ocu_b200 import ops
    torch.manual_seed(L + T + S)
    dev = "cuda"
    n_slots = T + 1
    q = _bf(torch.randn(S, L, 128, device=dev))
    kb = _bf(torch.randn(S, n_slots, L, 128, device=dev))
    vb = _bf(torch.randn(S, n_slots, L, 1024, device=dev))
    gate = _bf(torch.randn(S, L, 1024, device=dev))
    stride = 24
    order = torch.zeros(S, stride, device=dev, dtype=torch.int32)
    perms = [torch.randperm(n_slots)[:T] for _ in range(S)]
    for s in range(S):
        order[s, :T] = perms[s].to(dev)
    rows = [min(t, 3) for t in range(T)]
    bias_nat = torch.randn(S, L, 4, device=dev) * 0.5
    ws = ops.lt_attn_sk_workspace(L, T, dev, S=S)
    out = torch.empty(S, L, 1024, device=dev, dtype=torch.bfloat16)
    mass = torch.empty(S, L, T, device=dev)
    ops.lt_attn_sk_partial(q, kb, vb, order, T, (bias_nat * math.log2(math.e)).contiguous(), rows, True, ws, S=S,
                           order_stride=stride)
    ops.lt_attn_sk_combine(T, L, gate, out, mass, ws, S=S)
    torch.cuda.synchronize()
    for s in range(S):
        sc = torch.cat([q[s].float() @ kb[s, perms[s][t]].float().t() / math.sqrt(128.0) + bias_nat[s, :, rows[t]][:, None]
                        for t in range(T)], 1)
        a = torch.softmax(sc, -1)
        o_ref = (a @ torch.cat([vb[s, perms[s][t]].float() for t in range(T)], 0)) * gate[s].float()
        assert _rel(out[s], o_ref) < 2e-2, s
        assert (mass[s] - a.view(L, T, L).sum(-1)).abs().max().item() < 5e-3, s


def test_batched_token_groupnorm(lib_built):
    from rmem_ocu_b200 import ops
    import torch.nn.functional as F
    torch.manual_seed(3)
    S, L = 3, 1620
    xa = torch.randn(S * L, 256, device="cuda") * 2 + 0.5
    xb = torch.randn(S * L, 256, device="cuda") * 0.3 - 1
    xa[L:2 * L] *= 3.0                                           # sequences with different statistics
    g, b = torch.randn(512, device="cuda"), torch.randn(512, device="cuda")
    out = torch.empty(S * L, 512, device="cuda")
    ws = torch.zeros(S * 5, device="cuda", dtype=torch.float64)
    for _ in range(2):                                           # the workspaces clean themselves
        ops.groupnorm_tokens(xa, xb, g, b, out, 2, ws, S=S)
    torch.cuda.synchronize()
    assert (ws == 0).all()
    for s in range(S):
        x = torch.cat([xa[s * L:(s + 1) * L], xb[s * L:(s + 1) * L]], 1)
        ref = F.group_norm(x.t().unsqueeze(0), 2, g, b, 1e-5)[0].t()
        assert _rel(out[s * L:(s + 1) * L], ref) < 1e-5, s


def _clip(H, W, n, n_obj, seed):
    g = torch.Generator().manual_seed(seed)
    imgs = [torch.randn(1, 3, H, W, generator=g) for _ in range(n)]
    lab = torch.zeros(1, 1, H, W)
    s = H / 480.0
    for o in range(1, n_obj + 1):
        lab[:, :, int((40 + 60 * (o - 1)) * s):int((120 + 60 * (o - 1)) * s),
            int((100 + 100 * (o - 1)) * s):int((300 + 100 * (o - 1)) * s)] = o
    return imgs, lab.int()


@pytest.mark.parametrize("prefetch", [0, 1, 2])          # 0: no look-ahead; 1: next frame; 2: stacks of two lock-steps
def test_lockstep_engine_vs_single_engines_and_oracle(lib_built, monkeypatch, prefetch):
    """DeAOTEngine(n_sequences = 3): three different clips (different object counts) through one set of launches,
    against (a) three single-sequence engines -- same banks, masks >= 99.9 %, logits to 2e-2 (the stream-K partition
    and the batched cuDNN algorithms differ) -- and (b) the oracle on every clip: bank indexes equal, masks >= 99.9 %
    (fp32 cuDNN backbone, so that the comparison measures the batched kernels of the path)."""
    monkeypatch.setenv("RMEM_ENCODER_DTYPE", "fp32")
    import torch.nn.functional as F
    from oracle import restate as R
    from rmem_ocu_b200.engine import DeAOTEngine
    from rmem_ocu_b200.model import DeAOTModel, RMemConfig, deterministic_init_
    H, W, n, latter, S = 240, 432, 9, 2, 3
    m = DeAOTModel(RMemConfig(latter_mem_len=latter))
    deterministic_init_(m.state_dict(), 0)
    sd = {k: v.clone() for k, v in m.state_dict().items()}
    n_objs = [3, 1, 2]
    clips = [_clip(H, W, n, n_objs[s], 500 + s) for s in range(S)]

    def run(engine, idxs):
        imgs = lambda f: torch.cat([clips[s][0][f] for s in idxs]).cuda()
        labs = torch.cat([clips[s][1] for s in idxs]).cuda()
        masks, logits = [], []
        with torch.no_grad():
            engine.add_reference_frame(imgs(0), labs, obj_nums=[n_objs[s] for s in idxs], frame_step=0)
            for f in range(1, n):
                kw = {}
                if prefetch and f + 1 < n and engine.frames_ahead() <= 1:
                    # frame-major stack [B * S, ...]: the frames of the next B lock-steps, sequences adjacent
                    kw = {"next_img": torch.cat([imgs(k) for k in range(f + 1, min(f + 1 + prefetch, n))])}
                lg = engine.match_propogate_one_frame(imgs(f), output_size=(H, W), **kw)
                logits.append(lg.clone())
                label = torch.argmax(lg, 1, keepdim=True).float()
                masks.append(label[:, 0].to(torch.uint8).cpu())
                engine.update_short_term_memory(label)
        return masks, logits, engine.long_memories_indexes

    bm, bl, bidx = run(DeAOTEngine(m, 0, long_term_mem_gap=2, n_sequences=S), list(range(S)))
    assert len(bidx) == S and bl[0].shape == (S, 11, H, W)
    for s in range(S):
        sm, sl, sidx = run(DeAOTEngine(m, 0, long_term_mem_gap=2), [s])
        assert bidx[s] == sidx, s
        for f in range(n - 1):
            assert ((bl[f][s] - sl[f][0]).norm() / sl[f][0].norm()).item() < 2e-2, (s, f)
            assert (bm[f][s] == sm[f][0]).float().mean().item() >= 0.999, (s, f)
        o = R.Engine(sd, True, 1, latter, long_term_mem_gap=2)
        o_masks = []
        with torch.no_grad():
            o.add_reference_frame(clips[s][0][0], clips[s][1], obj_nums=[n_objs[s]], frame_step=0)
            for f in range(1, n):
                lg = o.match_propogate_one_frame(clips[s][0][f], output_size=(H, W))
                label = torch.argmax(lg, 1, keepdim=True).float()
                o_masks.append(label[0, 0].to(torch.uint8))
                o.update_short_term_memory(label)
        assert bidx[s] == o.long_memories_indexes, s
        agree = [(bm[f][s] == o_masks[f]).float().mean().item() for f in range(n - 1)]
        print("sequence", s, "lock-step engine vs oracle", [round(a, 5) for a in agree])
        assert min(agree) >= 0.999, (s, agree)
