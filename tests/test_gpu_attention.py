"""Parity of the attention kernels and the bank kernels against the oracle (oracle/restate.py)."""
import math

import pytest
import torch

pytestmark = pytest.mark.gpu


def _bf(x):
    return x.to(torch.bfloat16)


def _rel(a, b):
    return ((a.float() - b.float()).norm() / b.float().norm().clamp_min(1e-12)).item()


def _lt_run(q, kb, vb, order, T, cpf, bias=None, rows=None, gate=None):
    from rmem_ocu_b200 import ops
    L = q.shape[0]
    nch = T * cpf
    po = torch.empty(nch, L, 1024, device="cuda")
    pm = torch.empty(nch, L, device="cuda")
    pl = torch.empty(nch, L, device="cuda")
    out = torch.empty(L, 1024, device="cuda", dtype=torch.bfloat16)
    mass = torch.empty(L, T, device="cuda")
    ops.lt_attn_partial(q, kb, vb, order, T, cpf, po, pm, pl)
    ops.lt_attn_combine(po, pm, pl, bias, rows, T, L, cpf, gate, out, mass)
    torch.cuda.synchronize()
    return out, mass


@pytest.mark.parametrize("L,T,cpf,qscale", [(1705, 5, 1, 1.0), (1620, 3, 2, 3.0), (333, 9, 1, 6.0),
                                            (1705, 1, 3, 1.0), (64, 2, 1, 1.0), (3600, 2, 1, 2.0)])
def test_long_term_attention_vs_oracle(lib_built, L, T, cpf, qscale):
    """GatedPropagation explicit path + TPE + record_attn_weight + gate (oracle: restate.gated_propagation
    semantics, evaluated here in fp32 on the bf16-rounded inputs). Tolerance: bf16 kernel vs fp32,
    relative Frobenius error <= 2e-2 on the output, <= 5e-3 absolute on the per-frame masses."""
    torch.manual_seed(L + T)
    n_slots = T + 1
    q = _bf(torch.randn(L, 128, device="cuda") * qscale)
    kb = _bf(torch.randn(n_slots, L, 128, device="cuda"))
    vb = _bf(torch.randn(n_slots, L, 1024, device="cuda"))
    gate = _bf(torch.randn(L, 1024, device="cuda"))
    perm = torch.randperm(n_slots)[:T]
    order = perm.to(device="cuda", dtype=torch.int32)
    rows = [min(t, 3) for t in range(T)]
    bias_nat = torch.randn(L, 4, device="cuda") * 0.5
    out, mass = _lt_run(q, kb, vb, order, T, cpf, (bias_nat * math.log2(math.e)).contiguous(), rows, gate)
    s = torch.cat([q.float() @ kb[perm[t]].float().t() / math.sqrt(128.0) + bias_nat[:, rows[t]][:, None]
                   for t in range(T)], 1)
    a = torch.softmax(s, -1)
    o_ref = (a @ torch.cat([vb[perm[t]].float() for t in range(T)], 0)) * gate.float()
    m_ref = a.view(L, T, L).sum(-1)
    assert torch.isfinite(out.float()).all()
    assert _rel(out, o_ref) < 2e-2
    assert (mass - m_ref).abs().max().item() < 5e-3
    assert (mass.sum(1) - 1).abs().max().item() < 1e-4


@pytest.mark.parametrize("L,T,qscale,use_tpe,use_gate,record", [
    (1620, 5, 1.0, True, True, True),        # BASELINE config 2, steady state
    (1705, 5, 3.0, True, True, True), (1620, 1, 1.0, False, True, False),   # self-attention form (T = 1, no TPE)
    (1620, 9, 2.0, True, True, True),        # K = 8
    (333, 9, 6.0, True, False, True), (64, 2, 1.0, True, True, True), (42, 12, 1.0, True, True, True),
    (3600, 2, 2.0, True, True, True), (3600, 11, 1.0, True, True, True), (129, 1, 1.0, True, True, True)])
def test_long_term_attention_streamk_vs_oracle(lib_built, L, T, qscale, use_tpe, use_gate, record):
    """Second-generation long-term attention (persistent stream-K kernel on 2-CTA clusters with multicast K / V,
    csrc/lt_attn_sk.cu): the same oracle and tolerances as the first generation -- relative Frobenius error <= 2e-2
    on the output, <= 5e-3 absolute on the per-frame masses -- over shapes whose segments start / end inside
    frames, span several frames, and leave clusters without work."""
    from rmem_ocu_b200 import ops
    torch.manual_seed(L + T)
    n_slots = T + 1
    q = _bf(torch.randn(L, 128, device="cuda") * qscale)
    kb = _bf(torch.randn(n_slots, L, 128, device="cuda"))
    vb = _bf(torch.randn(n_slots, L, 1024, device="cuda"))
    gate = _bf(torch.randn(L, 1024, device="cuda")) if use_gate else None
    perm = torch.randperm(n_slots)[:T]
    order = perm.to(device="cuda", dtype=torch.int32)
    rows = [min(t, 3) for t in range(T)]
    bias_nat = torch.randn(L, 4, device="cuda") * 0.5 if use_tpe else torch.zeros(L, 4, device="cuda")
    ws = ops.lt_attn_sk_workspace(L, T, "cuda")
    ws.fill_(0xff)                                          # NaN patterns: nothing may be read before it is written
    out = torch.empty(L, 1024, device="cuda", dtype=torch.bfloat16)
    mass = torch.empty(L, T, device="cuda") if record else None
    for _ in range(2):                                      # twice: the kernel keeps no state between launches
        ops.lt_attn_sk_partial(q, kb, vb, order, T, (bias_nat * math.log2(math.e)).contiguous() if use_tpe else None,
                               rows if use_tpe else None, record, ws)
        ops.lt_attn_sk_combine(T, L, gate, out, mass, ws)
    torch.cuda.synchronize()
    s = torch.cat([q.float() @ kb[perm[t]].float().t() / math.sqrt(128.0) + bias_nat[:, rows[t]][:, None]
                   for t in range(T)], 1)
    a = torch.softmax(s, -1)
    o_ref = a @ torch.cat([vb[perm[t]].float() for t in range(T)], 0)
    if use_gate:
        o_ref = o_ref * gate.float()
    assert torch.isfinite(out.float()).all()
    assert _rel(out, o_ref) < 2e-2
    if record:
        m_ref = a.view(L, T, L).sum(-1)
        assert (mass - m_ref).abs().max().item() < 5e-3
        assert (mass.sum(1) - 1).abs().max().item() < 1e-4


@pytest.mark.parametrize("L,qscale,use_gate", [(1620, 1.0, True), (1705, 4.0, True), (128, 1.0, False), (77, 2.0, True),
                                               (3600, 3.0, True)])
def test_single_frame_attention_vs_oracle(lib_built, L, qscale, use_gate):
    """rmem_lt_attn_deaot_single (one launch, no partials): gated self-attention of one frame against fp32 on the
    bf16-rounded inputs, and against the partial + combine pair it replaces on the path (same kernel body, the
    pair merges three key chunks: rounding differs, so 1e-2 relative rather than equality)."""
    from rmem_ocu_b200 import ops
    torch.manual_seed(L)
    q = _bf(torch.randn(L, 128, device="cuda") * qscale)
    v = _bf(torch.randn(L, 1024, device="cuda"))
    gate = _bf(torch.rand(L, 1024, device="cuda")) if use_gate else None
    out = torch.full((L, 1024), float("nan"), device="cuda", dtype=torch.bfloat16)
    ops.lt_attn_single(q, q, v, gate, out)
    torch.cuda.synchronize()
    p = torch.softmax(q.float() @ q.float().t() / math.sqrt(128), -1)
    ref = p @ v.float()
    if use_gate:
        ref = ref * gate.float()
    assert torch.isfinite(out.float()).all()
    assert _rel(out, ref) <= 2e-2
    order = torch.zeros(1, device="cuda", dtype=torch.int32)
    pair, _ = _lt_run(q, q.view(1, L, 128), v.view(1, L, 1024), order, 1, 3, gate=gate)
    assert _rel(out, pair) <= 1e-2


@pytest.mark.parametrize("h,w", [(31, 55), (30, 54), (9, 20), (45, 80)])
def test_local_attention_vs_oracle(lib_built, h, w):
    """LocalGatedPropagation (15x15 window) against the oracle's direct windowed restatement."""
    from oracle import restate as R
    from rmem_ocu_b200 import ops
    torch.manual_seed(h * w)
    L = h * w
    q = _bf(torch.randn(L, 128) * 1.5)
    k = _bf(torch.randn(L, 128))
    v = _bf(torch.randn(L, 1024))
    u = _bf(torch.randn(L, 1024))
    rel_w = torch.randn(225, 128, 1, 1) * 0.1
    rel_b = torch.randn(225) * 0.1
    sd = {"p.relative_emb_k.weight": _bf(rel_w).float(), "p.relative_emb_k.bias": rel_b,
          "p.dw_conv.conv.weight": torch.zeros(1024, 1, 5, 5), "p.projection.weight": torch.zeros(8, 1024),
          "p.projection.bias": torch.zeros(8)}
    to2d = lambda t: t.float().view(h, w, 1, -1).permute(2, 3, 0, 1).contiguous()
    _, attn_ref, agg_ref = R.local_gated_propagation(to2d(q), to2d(k), to2d(v), u.float().view(L, 1, 1024),
                                                     (h, w), sd, "p")
    gated_ref = agg_ref.view(L, 1024) * u.float()
    probs_ref = attn_ref[0, 0].t()                                  # [L, 225]

    qd, kd, vd, ud = q.cuda(), k.cuda(), v.cuda(), u.cuda()
    rel = torch.zeros(L, 256, device="cuda")
    ops.linear(qd, _bf(rel_w.view(225, 128)).cuda(), rel_b.cuda(), [(0, 225, ops.ACT_NONE, ops.OUT_F32, rel, 0)])
    probs = torch.empty(L, 225, device="cuda")
    out = torch.empty(L, 1024, device="cuda", dtype=torch.bfloat16)
    banded = ops.local_banded_buffer(h, w, "cuda")
    ops.local_attn_scores(qd, kd, rel, probs, banded, h, w)
    ops.local_attn_aggregate(banded, vd, ud, out, h, w)
    torch.cuda.synchronize()
    assert (probs.cpu() - probs_ref).abs().max().item() < 2e-3
    assert (probs.sum(1) - 1).abs().max().item() < 1e-4
    assert _rel(out.cpu(), gated_ref) < 2e-2
    # the fused single-kernel form (what the engines call) -- also without a gate
    out_f = torch.empty_like(out)
    ops.local_attn(qd, kd, vd, rel, ud, out_f, h, w)
    out_n = torch.empty_like(out)
    ops.local_attn(qd, kd, vd, rel, None, out_n, h, w)
    torch.cuda.synchronize()
    assert torch.isfinite(out_f.float()).all()
    assert _rel(out_f.cpu(), gated_ref) < 2e-2
    assert _rel(out_n.cpu(), agg_ref.view(L, 1024)) < 2e-2


def test_bank_score_evict_bit_exact_vs_oracle(lib_built):
    """Slot choices of the device rule == oracle rule (restate.eviction_scores) over long random
    histories, for the DeAOT call pattern (statistics on every append) and the AOT one (only when full)."""
    from oracle import restate as R
    from rmem_ocu_b200 import ops, _lib
    for variant, seed in [("deaot", 0), ("aot", 1), ("deaot", 2)]:
        g = torch.Generator().manual_seed(seed)
        L, cap = 420, 5
        state = torch.zeros(_lib.BANK_STATE_WORDS, device="cuda", dtype=torch.int32)
        ops.bank_init(state, 0)
        idx = [0]
        ev = R.EvictionState()
        T = 1
        for step in range(1, 120):
            # concentrated, non-uniform masses so the relevance term matters
            W = torch.rand(L, T, generator=g) ** 3 + 1e-3
            W = W / W.sum(1, keepdim=True)
            fg = torch.rand(L, generator=g)
            idx.append(step * 3)
            src = torch.zeros(L, 8, device="cuda")
            dst = torch.zeros(_lib.MAX_BANK, L, 8, device="cuda")
            ops.bank_append(state, T, step * 3, [src], [dst])
            T_new = T + 1
            if variant == "aot" and T_new <= cap:
                T = T_new
                continue
            drop, detail = R.eviction_scores(W, fg, idx, ev)
            ops.bank_score_evict(state, W.cuda().contiguous(), fg.cuda(), L, T, cap)
            st = state.cpu()
            assert int(st[5 * _lib.MAX_BANK + 4]) == drop, (variant, step, detail["margin"])
            if T_new > cap:
                idx.remove(idx[drop])
                T = cap
            else:
                T = T_new
            order = st[:T].tolist()
            frame_ids = [int(st[_lib.MAX_BANK + s]) for s in order]
            assert frame_ids == idx, (variant, step)


def _mha_run(q, k, v, order, T, cpf, bias=None, rows=None):
    from rmem_ocu_b200 import ops
    L = q.shape[0]
    nch = T * cpf
    po = torch.empty(nch, L, 256, device="cuda")
    pm = torch.empty(nch, 8, L, device="cuda")
    pl = torch.empty(nch, 8, L, device="cuda")
    out = torch.zeros(L, 512, device="cuda", dtype=torch.bfloat16)
    mass = torch.empty(L, T, device="cuda")
    ops.mha_partial(q, k, v, order, T, cpf, po, pm, pl)
    ops.mha_combine(po, pm, pl, bias, rows, T, L, cpf, out, 256, mass)
    torch.cuda.synchronize()
    return out, mass


@pytest.mark.parametrize("L,T,cpf,qscale", [(1705, 5, 1, 1.0), (1620, 3, 2, 3.0), (333, 9, 1, 6.0),
                                            (1620, 1, 3, 1.0), (64, 2, 1, 1.0), (3600, 2, 1, 2.0)])
def test_multihead_attention_vs_oracle(lib_built, L, T, cpf, qscale):
    """MultiheadAttention explicit path (8 heads x 32, attention.py:45-64) + per-head TPE bias + head-averaged
    record_attn_weight, evaluated in fp32 on the bf16-rounded inputs.  Tolerance: relative Frobenius error
    <= 2e-2 on the output, <= 5e-3 absolute on the per-frame masses."""
    torch.manual_seed(L + T + 7)
    n_slots = T + 1
    qbuf = _bf(torch.randn(L, 512, device="cuda") * qscale)         # queries live in columns 256.. (ld 512)
    q = qbuf[:, 256:]
    kb = _bf(torch.randn(n_slots, L, 256, device="cuda"))
    vb = _bf(torch.randn(n_slots, L, 256, device="cuda"))
    perm = torch.randperm(n_slots)[:T]
    order = perm.to(device="cuda", dtype=torch.int32)
    rows = [min(t, 3) for t in range(T)]
    bias_nat = torch.randn(L, 8, 4, device="cuda") * 0.5
    out, mass = _mha_run(q, kb, vb, order, T, cpf, (bias_nat * math.log2(math.e)).contiguous(), rows)
    qh = q.float().view(L, 8, 32).permute(1, 0, 2)                                  # [8, L, 32]
    s = torch.cat([qh @ kb[perm[t]].float().view(L, 8, 32).permute(1, 2, 0) / math.sqrt(32.0)
                   + bias_nat[:, :, rows[t]].t()[:, :, None] for t in range(T)], 2)  # [8, L, T*L]
    a = torch.softmax(s, -1)
    vcat = torch.cat([vb[perm[t]].float() for t in range(T)], 0).view(T * L, 8, 32).permute(1, 0, 2)
    o_ref = (a @ vcat).permute(1, 0, 2).reshape(L, 256)
    m_ref = a.mean(0).view(L, T, L).sum(-1)
    assert torch.isfinite(out.float()).all()
    assert (out[:, :256] == 0).all()                                # column window respected
    assert _rel(out[:, 256:], o_ref) < 2e-2
    assert (mass - m_ref).abs().max().item() < 5e-3
    assert (mass.sum(1) - 1).abs().max().item() < 1e-4


def test_multihead_attention_single_slot_strided(lib_built):
    """Self / short-term attention form: K and V are column windows of wider [L, ld] buffers, T = 1, no TPE."""
    from rmem_ocu_b200 import ops
    torch.manual_seed(5)
    L = 1620
    buf = _bf(torch.randn(L, 768, device="cuda"))
    q, k, v = buf[:, :256], buf[:, 256:512], buf[:, 512:]
    order = torch.zeros(1, device="cuda", dtype=torch.int32)
    po = torch.empty(3, L, 256, device="cuda"); pm = torch.empty(3, 8, L, device="cuda"); pl = torch.empty(3, 8, L, device="cuda")
    out = torch.empty(L, 256, device="cuda", dtype=torch.bfloat16)
    ops.mha_partial(q, k, v, order, 1, 3, po, pm, pl)
    ops.mha_combine(po, pm, pl, None, None, 1, L, 3, out, 0, None)
    torch.cuda.synchronize()
    h = lambda t: t.float().view(L, 8, 32).permute(1, 0, 2)
    ref = (torch.softmax(h(q) @ h(k).transpose(1, 2) / math.sqrt(32.0), -1) @ h(v)).permute(1, 0, 2).reshape(L, 256)
    assert _rel(out, ref) < 2e-2
