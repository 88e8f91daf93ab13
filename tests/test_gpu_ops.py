"""Parity of the element-wise / GEMM / bank kernels against plain PyTorch fp32 and the oracle."""
import math

import pytest
import torch
import torch.nn.functional as F

pytestmark = pytest.mark.gpu


def _bf(x):
    return x.to(torch.bfloat16)


def _rel(a, b):
    return ((a.float() - b.float()).norm() / b.float().norm().clamp_min(1e-12)).item()


@pytest.mark.parametrize("M,K,N", [(1705, 256, 1152), (1620, 2048, 512), (100, 512, 2176), (3600, 1024, 512)])
def test_linear_segments(lib_built, M, K, N):
    from rmem_ocu_b200 import ops
    torch.manual_seed(0)
    dev = "cuda"
    a = _bf(torch.randn(M, K, device=dev))
    w = _bf(torch.randn(N, K, device=dev) / math.sqrt(K))
    bias = torch.randn(N, device=dev)
    ref = a.float() @ w.float().t() + bias
    n0 = 128 if N > 256 else 64
    n1 = (N - n0) // 2 // 32 * 32 + n0
    o0 = torch.zeros(M, n0 + 64, device=dev, dtype=torch.bfloat16)          # ld > width, col offset 32
    o1 = torch.zeros(M, n1 - n0, device=dev, dtype=torch.float32)
    o2 = torch.randn(M, N - n1, device=dev, dtype=torch.float32)
    o2_init = o2.clone()
    ops.linear(a, w, bias, [(0, n0, ops.ACT_NONE, ops.OUT_BF16, o0, 32),
                            (n0, n1, ops.ACT_SILU, ops.OUT_F32, o1, 0),
                            (n1, N, ops.ACT_NONE, ops.OUT_F32_ACCUM, o2, 0)])
    torch.cuda.synchronize()
    assert _rel(o0[:, 32:32 + n0], ref[:, :n0]) < 6e-3
    assert (o0[:, :32] == 0).all() and (o0[:, 32 + n0:] == 0).all()
    assert _rel(o1, F.silu(ref[:, n0:n1])) < 2e-3
    assert _rel(o2 - o2_init, ref[:, n1:]) < 2e-3


def test_linear_ragged_columns(lib_built):
    """N not a multiple of 32 (the 225 relative-position taps)."""
    from rmem_ocu_b200 import ops
    torch.manual_seed(1)
    a = _bf(torch.randn(333, 128, device="cuda"))
    w = _bf(torch.randn(225, 128, device="cuda") * 0.1)
    bias = torch.randn(225, device="cuda")
    out = torch.full((333, 256), 7.0, device="cuda")
    ops.linear(a, w, bias, [(0, 225, ops.ACT_NONE, ops.OUT_F32, out, 0)])
    torch.cuda.synchronize()
    ref = a.float() @ w.float().t() + bias
    assert _rel(out[:, :225], ref) < 2e-3
    assert (out[:, 225:] == 7.0).all()


@pytest.mark.parametrize("in_bf16", [False, True])
def test_layernorm_dual(lib_built, in_bf16):
    from rmem_ocu_b200 import ops
    torch.manual_seed(2)
    L = 1705
    x0 = torch.randn(L, 256, device="cuda") * 3 + 1
    x1 = torch.randn(L, 256, device="cuda")
    if in_bf16:
        x0, x1 = _bf(x0), _bf(x1)
    g0, b0, g1, b1 = (torch.randn(256, device="cuda") for _ in range(4))
    out = torch.zeros(L, 512, device="cuda", dtype=torch.bfloat16)
    ops.layernorm(x0, g0, b0, out, 0, x1, g1, b1, out, 256)
    torch.cuda.synchronize()
    r0 = F.layer_norm(x0.float(), (256,), g0, b0, 1e-5)
    r1 = F.layer_norm(x1.float(), (256,), g1, b1, 1e-5)
    assert _rel(out[:, :256], r0) < 4e-3 and _rel(out[:, 256:], r1) < 4e-3


@pytest.mark.parametrize("dt", [torch.float16, torch.bfloat16, torch.float32])
def test_layernorm_copy(lib_built, dt):
    """norm1 of layer 0 straight off the encoder's embedding (rmem_layernorm_copy): LN as bf16 at a column offset and
    the input itself as the fp32 residual stream, exactly (a widening conversion)."""
    from rmem_ocu_b200 import ops
    torch.manual_seed(12)
    L = 1705
    x = (torch.randn(L, 256, device="cuda") * 3 + 1).to(dt)
    g, b = torch.randn(256, device="cuda"), torch.randn(256, device="cuda")
    out = torch.zeros(L, 512, device="cuda", dtype=torch.bfloat16)
    tgt = torch.full((L, 256), float("nan"), device="cuda")
    ops.layernorm_copy(x, g, b, out, 256, tgt)
    torch.cuda.synchronize()
    assert torch.equal(tgt, x.float())
    assert _rel(out[:, 256:], F.layer_norm(x.float(), (256,), g, b, 1e-5)) < 4e-3 and (out[:, :256] == 0).all()


@pytest.mark.parametrize("h,w", [(31, 55), (30, 54), (5, 3)])
def test_dwconv5(lib_built, h, w):
    from rmem_ocu_b200 import ops
    torch.manual_seed(3)
    C = 1024
    x = _bf(torch.randn(h * w, C, device="cuda"))
    wt = torch.randn(C, 1, 5, 5, device="cuda") * 0.2
    y = torch.zeros(h * w, 2 * C, device="cuda", dtype=torch.bfloat16)
    ops.dwconv5(x, 0, wt.view(C, 25).t().contiguous(), y, C, h, w, C)
    torch.cuda.synchronize()
    ref = F.conv2d(x.float().view(1, h, w, C).permute(0, 3, 1, 2), wt, None, 1, 2, 1, C)
    ref = ref.permute(0, 2, 3, 1).reshape(h * w, C)
    assert _rel(y[:, C:], ref) < 4e-3 and (y[:, :C] == 0).all()


def test_groupnorm_tokens(lib_built):
    from rmem_ocu_b200 import ops
    torch.manual_seed(4)
    L = 1705
    xa = torch.randn(L, 256, device="cuda") * 2 + 0.5
    xb = torch.randn(L, 256, device="cuda") * 0.3 - 1
    g, b = torch.randn(512, device="cuda"), torch.randn(512, device="cuda")
    out = torch.empty(L, 512, device="cuda", dtype=torch.bfloat16)
    ws = torch.zeros(5, device="cuda", dtype=torch.float64)
    ops.groupnorm_tokens(xa, xb, g, b, out, 2, ws)
    out32 = torch.empty(L, 512, device="cuda")
    ops.groupnorm_tokens(xa, xb, g, b, out32, 2, ws)             # the workspace cleans itself
    torch.cuda.synchronize()
    x = torch.cat([xa, xb], 1)
    ref = F.group_norm(x.t().unsqueeze(0), 2, g, b, 1e-5)[0].t()
    assert _rel(out, ref) < 4e-3
    assert _rel(out32, ref) < 1e-5
    # other shapes: one input, groups that split a warp's channels, rows that do not fill the last CTA
    for (Ca, Cb, groups, L2) in ((256, 0, 8, 333), (192, 64, 2, 77), (1024, 0, 4, 1620)):
        xa2 = torch.randn(L2, Ca, device="cuda") + 0.3
        xb2 = torch.randn(L2, Cb, device="cuda") * 2 if Cb else None
        g2, b2 = torch.randn(Ca + Cb, device="cuda"), torch.randn(Ca + Cb, device="cuda")
        o2 = torch.empty(L2, Ca + Cb, device="cuda")
        ops.groupnorm_tokens(xa2, xb2, g2, b2, o2, groups, ws)
        x2 = torch.cat([xa2, xb2], 1) if Cb else xa2
        ref2 = F.group_norm(x2.t().unsqueeze(0), groups, g2, b2, 1e-5)[0].t()
        assert _rel(o2, ref2) < 1e-5, (Ca, Cb, groups, L2)


def test_tpe_prepare(lib_built):
    from rmem_ocu_b200 import ops
    torch.manual_seed(5)
    L = 1705
    q = _bf(torch.randn(L, 128, device="cuda"))
    cur = torch.randn(128, device="cuda") * 0.05
    mem = torch.randn(4, 128, device="cuda") * 0.05
    qt = torch.empty_like(q)
    bias = torch.empty(L, 4, device="cuda")
    ops.tpe_prepare(q, cur, mem, qt, bias)
    torch.cuda.synchronize()
    qref = q.float() + cur
    assert _rel(qt, qref) < 4e-3
    bref = (qref @ mem.t()) * (math.log2(math.e) / math.sqrt(128))   # f32 q + cur: the bias is not re-rounded
    assert (bias - bref).abs().max().item() < 1e-4


def test_foreground_prob(lib_built):
    from rmem_ocu_b200 import ops
    torch.manual_seed(6)
    logits = torch.randn(1, 11, 121, 217, device="cuda") * 3
    h, w = 31, 55
    fg = torch.empty(h * w, device="cuda")
    ops.foreground_prob(logits[0], fg, h, w)
    torch.cuda.synchronize()
    pl = F.interpolate(logits, size=(h, w), mode="bilinear", align_corners=True)
    ref = (1 - torch.softmax(pl, 1)[:, 0]).flatten()
    assert (fg - ref).abs().max().item() < 1e-5


@pytest.mark.parametrize("H,W", [(480, 864), (481, 865), (240, 432)])
def test_id_embed_vs_oracle(lib_built, H, W):
    """one-hot + ignore channel + 17x17 s16 conv + id_norm (oracle: restate.Engine.assign_identity)."""
    from oracle import restate as R
    from rmem_ocu_b200 import ops
    torch.manual_seed(7)
    wt = torch.randn(256, 12, 17, 17) * 0.05
    bias = torch.randn(256) * 0.1
    g, b = torch.randn(256), torch.randn(256)
    lab = torch.zeros(1, 1, H, W)
    lab[:, :, 40:200, 100:300] = 3
    lab[:, :, 150:260, 250:500] = 10
    lab[:, :, 5:30, 0:40] = 255                       # ignore region
    lab[:, :, 300:310, 600:650] = 77                  # label outside 0..10 / 255 -> no class at all
    lab[:, :, ::37, ::29] = 1                         # isolated pixels (runs of length 1)
    one_hot, ignore = R.one_hot_mask(lab, 10)
    one_hot[:, 0] = one_hot[:, 0] * (ignore == 0).float().squeeze()
    sd = {"patch_wise_id_bank.weight": wt, "patch_wise_id_bank.bias": bias, "id_norm.weight": g, "id_norm.bias": b}
    ref = R.get_id_emb(torch.cat((one_hot, ignore), 1), sd, True)          # [1, 256, h, w]
    h, w = ref.shape[-2:]
    ref = ref.permute(0, 2, 3, 1).reshape(h * w, 256)
    pref = ops.id_prefix_table(wt)
    out = torch.zeros(h * w, 512, device="cuda", dtype=torch.bfloat16)
    out2 = torch.zeros(h * w, 256, device="cuda", dtype=torch.bfloat16)
    ops.id_embed(lab[0, 0].cuda(), pref.cuda(), bias.cuda(), g.cuda(), b.cuda(), 12, [(out, 256), (out2, 0)], h, w)
    torch.cuda.synchronize()
    assert _rel(out[:, 256:].cpu(), ref) < 5e-3
    assert torch.equal(out[:, 256:], out2) and (out[:, :256] == 0).all()


@pytest.mark.parametrize("H,W,S,with_ignore,norm", [(480, 864, 1, True, True), (241, 433, 2, False, True),
                                                    (113, 97, 1, True, False)])
def test_id_embed_soft_vs_oracle(lib_built, H, W, S, with_ignore, norm):
    """Probability / one-hot masks through rmem_id_embed_soft (update_short_term_memory's second branch,
    aot_engine.py:330-336) against the oracle's assign_identity (fp32 conv + id_norm).  Output is bf16: relative
    Frobenius error <= 5e-3; a hard one-hot mask must also give what the label-map kernel gives."""
    from oracle import restate as R
    from rmem_ocu_b200 import ops
    torch.manual_seed(11)
    wt = torch.randn(256, 12, 17, 17) * 0.05
    bias = torch.randn(256) * 0.1
    g, b = torch.randn(256), torch.randn(256)
    sd = {"patch_wise_id_bank.weight": wt, "patch_wise_id_bank.bias": bias, "id_norm.weight": g, "id_norm.bias": b}
    prob = torch.softmax(torch.randn(S, 11, H, W) * 3, 1)
    prob[:, :, : H // 3] = F.one_hot(prob[:, :, : H // 3].argmax(1), 11).permute(0, 3, 1, 2).float()   # saturated part
    ignore = (torch.rand(S, 1, H, W) > 0.97).float() if with_ignore else None
    h, w = (H + 16 - 17) // 16 + 1, (W + 16 - 17) // 16 + 1
    refs = []
    for s in range(S):
        oh = prob[s:s + 1].clone()
        ig = ignore[s:s + 1] if with_ignore else torch.zeros(1, 1, H, W)
        oh[:, 0] = oh[:, 0] * (ig == 0).float().squeeze()
        r = R.get_id_emb(torch.cat((oh, ig), 1), sd, norm)
        assert r.shape[-2:] == (h, w)
        refs.append(r.permute(0, 2, 3, 1).reshape(h * w, 256))
    ref = torch.cat(refs, 0)
    taps = ops.id_tap_table(wt).cuda()
    out = torch.zeros(S * h * w, 512, device="cuda", dtype=torch.bfloat16)
    out2 = torch.zeros(S * h * w, 256, device="cuda", dtype=torch.bfloat16)
    gg, bb = (g.cuda(), b.cuda()) if norm else (None, None)
    ops.id_embed_soft(prob.cuda(), ignore[:, 0].contiguous().cuda() if with_ignore else None, taps, bias.cuda(), gg, bb,
                      12, [(out, 256), (out2, 0)], h, w)
    torch.cuda.synchronize()
    assert _rel(out[:, 256:].cpu(), ref) < 5e-3
    assert torch.equal(out[:, 256:], out2) and (out[:, :256] == 0).all()
    if S == 1:
        # hard masks: the label kernel's table look-up and the weighted sum agree to bf16 rounding of fp32 sums
        lab = prob.argmax(1).float()
        hard = F.one_hot(lab.long(), 11).permute(0, 3, 1, 2).float().contiguous()
        a = torch.zeros(h * w, 256, device="cuda", dtype=torch.bfloat16)
        c = torch.zeros(h * w, 256, device="cuda", dtype=torch.bfloat16)
        ops.id_embed_soft(hard.cuda(), None, taps, bias.cuda(), gg, bb, 12, [(a, 0)], h, w)
        ops.id_embed(lab[0].cuda(), ops.id_prefix_table(wt).cuda(), bias.cuda(), gg, bb, 12, [(c, 0)], h, w)
        torch.cuda.synchronize()
        assert _rel(a, c) < 3e-3


@pytest.mark.parametrize("C,hi,wi,ho,wo,S", [(256, 30, 54, 30, 54, 1), (128, 30, 54, 60, 108, 2), (128, 31, 55, 61, 109, 1)])
def test_fp16_decoder_glue_vs_torch(lib_built, C, hi, wi, ho, wo, S):
    """The fp16-activation forms of the decoder glue (rmem_groupnorm_nhwc_ex / rmem_upsample_add_nhwc_ex, f16 = 1):
    GroupNorm(8) + ReLU and adapter + bilinear resize computed in fp32 from fp16 storage, against torch in fp32 on
    the same rounded inputs; outputs are fp16: 1e-3 relative."""
    from rmem_ocu_b200 import ops
    torch.manual_seed(31)
    x = (torch.randn(S, C, hi, wi, device="cuda") * 2 + 0.3).half().contiguous(memory_format=torch.channels_last)
    pb, g, b = (torch.randn(C, device="cuda") for _ in range(3))
    ref = F.relu(F.group_norm(x.float() + pb.view(1, C, 1, 1), 8, g, b, 1e-5))
    tok = lambda t: t.permute(0, 2, 3, 1).reshape(-1, t.shape[1])
    y = torch.empty_like(x)
    ws = torch.zeros(S, 17, device="cuda", dtype=torch.float64)
    ops.groupnorm_nhwc(tok(x), g, b, tok(y), 8, True, ws, pre_bias=pb, S=S)
    torch.cuda.synchronize()
    assert y.dtype == torch.float16 and _rel(y, ref) < 1e-3 and (ws == 0).all()
    a = torch.randn(S, C, ho, wo, device="cuda").contiguous(memory_format=torch.channels_last)
    ab = torch.randn(C, device="cuda")
    out = torch.empty(S, C, ho, wo, device="cuda", dtype=torch.float16).contiguous(memory_format=torch.channels_last)
    ops.upsample_add_nhwc(tok(y), hi, wi, tok(a), tok(out), ho, wo, a_bias=ab, S=S)
    torch.cuda.synchronize()
    ref2 = a + ab.view(1, C, 1, 1) + F.interpolate(y.float(), size=(ho, wo), mode="bilinear", align_corners=True)
    assert _rel(out, ref2) < 1e-3
    # the two steps as one stage (rmem_gn_upsample_add_nhwc), fp16 and fp32 storage; the normalised map is never stored
    ref3 = a + ab.view(1, C, 1, 1) + F.interpolate(ref, size=(ho, wo), mode="bilinear", align_corners=True)
    for xin, tol in ((x, 1e-3), (x.float(), 1e-5)):
        o3 = torch.empty(S, C, ho, wo, device="cuda", dtype=xin.dtype).contiguous(memory_format=torch.channels_last)
        for _ in range(2):                                       # twice: the workspace cleans itself
            ops.gn_upsample_add_nhwc(tok(xin), pb, g, b, hi, wi, tok(a), tok(o3), ho, wo, ws, a_bias=ab, S=S)
        torch.cuda.synchronize()
        assert _rel(o3, ref3) < tol and (ws == 0).all()


@pytest.mark.parametrize("f16", [False, True])
@pytest.mark.parametrize("C,Hs,Ws,Ho,Wo,S", [(128, 120, 216, 480, 864, 1), (128, 61, 109, 241, 433, 2), (64, 9, 7, 0, 0, 1)])
def test_logits_head_vs_torch(lib_built, C, Hs, Ws, Ho, Wo, S, f16):
    """rmem_logits_head_batched -- GroupNorm(8) + ReLU + 1x1 conv_out + bias + class mask (+ bilinear resize + argmax)
    -- against F.group_norm / F.conv2d / F.interpolate in fp32 (TF32 off): pred to 1e-5 relative, resized logits to
    1e-5, the label map equal to torch.argmax of the kernel's own resized logits."""
    from rmem_ocu_b200 import ops
    torch.manual_seed(21)
    Co = 11
    x = (torch.randn(S, C, Hs, Ws, device="cuda") * 2 + 0.3).contiguous(memory_format=torch.channels_last)
    if f16:                                    # fp16 storage of the conv output; the reference sees the same rounded values
        xk = x.half()
        x = xk.float()
    else:
        xk = x
    pb, g, b = (torch.randn(C, device="cuda") for _ in range(3))
    w = torch.randn(Co, C, device="cuda") * 0.2
    bo = torch.randn(Co, device="cuda")
    n_valid = [4, 11][:S]
    old = torch.backends.cudnn.allow_tf32
    torch.backends.cudnn.allow_tf32 = False
    try:
        y = F.relu(F.group_norm(x + pb.view(1, C, 1, 1), 8, g, b, 1e-5))
        ref = F.conv2d(y, w.view(Co, C, 1, 1), bo)
    finally:
        torch.backends.cudnn.allow_tf32 = old
    for s, nv in enumerate(n_valid):
        ref[s, nv:] = -1e10
    ws = torch.zeros(S, 17, device="cuda", dtype=torch.float64)
    pred = torch.empty(S, Co, Hs, Ws, device="cuda")
    out = torch.empty(S, Co, Ho, Wo, device="cuda") if Ho else None
    label = torch.empty(S, 1, Ho, Wo, device="cuda") if Ho else None
    for _ in range(2):                                          # twice: the workspace must clean itself
        ops.logits_head(xk, pb, g, b, w, bo, n_valid, pred, out, label, ws)
    torch.cuda.synchronize()
    assert (ws == 0).all()
    assert _rel(pred, ref) < 1e-5
    if Ho:
        ref_out = F.interpolate(pred, size=(Ho, Wo), mode="bilinear", align_corners=True)
        valid = ref_out > -1e9
        assert torch.allclose(out[valid], ref_out[valid], rtol=1e-4, atol=1e-4)
        assert torch.equal(label, out.argmax(1, keepdim=True).float())


@pytest.mark.parametrize("C,H,W", [(256, 30, 54), (128, 60, 108), (128, 120, 216), (64, 7, 5)])
def test_groupnorm_nhwc(lib_built, C, H, W):
    """GroupNorm(8) + ReLU of the decoder's ConvGN blocks (basic.py:60-70, fpn.py:44-63) vs F.group_norm."""
    from rmem_ocu_b200 import ops
    torch.manual_seed(8)
    x = (torch.randn(1, C, H, W, device="cuda") * 2 + 0.3).contiguous(memory_format=torch.channels_last)
    g, b = torch.randn(C, device="cuda"), torch.randn(C, device="cuda")
    pb = torch.randn(C, device="cuda")                       # the conv bias, folded into the norm
    ref = F.relu(F.group_norm(x + pb.view(1, C, 1, 1), 8, g, b, 1e-5))
    ref0 = F.relu(F.group_norm(x, 8, g, b, 1e-5))
    t = x.permute(0, 2, 3, 1).reshape(H * W, C)
    y = torch.empty_like(t)
    ws = torch.zeros(17, device="cuda", dtype=torch.float64)
    ops.groupnorm_nhwc(t, g, b, y, 8, True, ws, pre_bias=pb)
    torch.cuda.synchronize()
    assert (y.view(1, H, W, C).permute(0, 3, 1, 2) - ref).abs().max().item() < 2e-4
    ops.groupnorm_nhwc(t, g, b, t, 8, True, ws)             # in place, no bias
    torch.cuda.synchronize()
    assert (x - ref0).abs().max().item() < 2e-4


@pytest.mark.parametrize("C,hi,wi,ho,wo", [(256, 30, 54, 60, 108), (128, 60, 108, 120, 216), (128, 31, 55, 61, 109)])
def test_upsample_add_nhwc(lib_built, C, hi, wi, ho, wo):
    """adapter(shortcut) + bilinear(align_corners=True) upsample of the FPN head (fpn.py:49-60)."""
    from rmem_ocu_b200 import ops
    torch.manual_seed(9)
    x = torch.randn(1, C, hi, wi, device="cuda").contiguous(memory_format=torch.channels_last)
    a = torch.randn(1, C, ho, wo, device="cuda").contiguous(memory_format=torch.channels_last)
    ab = torch.randn(C, device="cuda")
    ref = a + ab.view(1, C, 1, 1) + F.interpolate(x, size=(ho, wo), mode="bilinear", align_corners=True)
    ta = a.permute(0, 2, 3, 1).reshape(ho * wo, C)
    ops.upsample_add_nhwc(x.permute(0, 2, 3, 1).reshape(hi * wi, C), hi, wi, ta, ta, ho, wo, a_bias=ab)
    torch.cuda.synchronize()
    assert (a - ref).abs().max().item() < 1e-4


@pytest.mark.parametrize("Hs,Ws,Ho,Wo,n_valid", [(120, 216, 480, 864, 6), (32, 48, 128, 192, 2), (31, 55, 121, 217, 11)])
def test_finish_logits_and_argmax(lib_built, Hs, Ws, Ho, Wo, n_valid):
    """Tail of decode_current_logits (aot_engine.py:451-463): bias, -1e10 masking of unused object channels,
    bilinear align_corners resize; and the evaluator's argmax over channels (evaluator.py:430-441)."""
    from rmem_ocu_b200 import ops
    torch.manual_seed(12)
    C = 11
    conv = torch.randn(1, C, Hs, Ws, device="cuda").contiguous(memory_format=torch.channels_last)
    bias = torch.randn(C, device="cuda")
    ref = conv + bias.view(1, C, 1, 1)
    ref[:, n_valid:] = -1e10
    ref_up = F.interpolate(ref, size=(Ho, Wo), mode="bilinear", align_corners=True)
    pred = torch.empty(1, C, Hs, Ws, device="cuda")
    out = torch.empty(1, C, Ho, Wo, device="cuda")
    ops.finish_logits(conv, bias, pred, n_valid, out)
    label = ops.argmax_label(out)
    torch.cuda.synchronize()
    assert torch.equal(pred, ref.contiguous())
    assert (out[:, :n_valid] - ref_up[:, :n_valid]).abs().max().item() < 1e-4
    assert (out[:, n_valid:] < -9e9).all()
    assert torch.equal(label, torch.argmax(out, 1, keepdim=True).float())
    assert (label == torch.argmax(ref_up, 1, keepdim=True).float()).float().mean().item() > 0.9999


@pytest.mark.parametrize("dec_dtype,tol", [("fp32", 1e-4), ("fp16", 3e-3)])
def test_decoder_head_vs_oracle(lib_built, monkeypatch, dec_dtype, tol):
    """InferenceFPN (cuDNN convs + csrc/decoder.cu) against the oracle's restatement of fpn.py:36-68: everything in
    fp32 (RMEM_DECODER_DTYPE=fp32) to 1e-4, the default fp16-activation form (fp32 accumulation, statistics,
    normalisation, interpolation and logits) to 3e-3."""
    monkeypatch.setenv("RMEM_DECODER_DTYPE", dec_dtype)
    from oracle import restate as R
    from rmem_ocu_b200.engine import InferenceFPN
    from tests import _synth
    _, sd = _synth.models(2)
    torch.manual_seed(10)
    h, w = 15, 27
    embs = [torch.randn(1, 256, 4 * h, 4 * w), torch.randn(1, 512, 2 * h, 2 * w), torch.randn(1, 1024, h, w),
            torch.randn(1, 256, h, w)]
    lstt = torch.randn(1, 512, h, w)
    with torch.no_grad():
        ref = R.fpn_decoder([embs[-1], lstt], embs, sd, intermediate=False)
    fpn = InferenceFPN(sd, torch.device("cuda"), intermediate=False)
    cl = lambda t: t.cuda().contiguous(memory_format=torch.channels_last)
    with torch.no_grad(), torch.backends.cudnn.flags(enabled=True, allow_tf32=False):
        conv, bias = fpn([cl(embs[-1]), cl(lstt)], [cl(e) for e in embs])
    out = conv + bias.view(1, -1, 1, 1)
    torch.cuda.synchronize()
    assert _rel(out.cpu(), ref) < tol


def test_decoder_head_bf16_shortcuts(lib_built):
    """Production form of the head: the encoder's feature maps arrive in bf16 and the three 1x1 adapters run on
    the tcgen05 GEMM (bf16 weights, f32 accumulate).  Against the oracle on the same bf16-rounded inputs."""
    from oracle import restate as R
    from rmem_ocu_b200.engine import InferenceFPN
    from tests import _synth
    _, sd = _synth.models(2)
    torch.manual_seed(11)
    h, w = 15, 27
    embs = [_bf(torch.randn(1, 256, 4 * h, 4 * w)), _bf(torch.randn(1, 512, 2 * h, 2 * w)),
            _bf(torch.randn(1, 1024, h, w)), _bf(torch.randn(1, 256, h, w))]
    lstt = _bf(torch.randn(1, 512, h, w))
    with torch.no_grad():
        ref = R.fpn_decoder([embs[-1].float(), lstt.float()], [e.float() for e in embs], sd, intermediate=False)
    fpn = InferenceFPN(sd, torch.device("cuda"), intermediate=False)
    cl = lambda t: t.cuda().contiguous(memory_format=torch.channels_last)
    with torch.no_grad():
        conv, bias = fpn([cl(embs[-1]), cl(lstt)], [cl(e) for e in embs])
    out = conv + bias.view(1, -1, 1, 1)
    torch.cuda.synchronize()
    assert _rel(out.cpu(), ref) < 5e-3          # bf16 adapter weights + TF32 cuDNN convolutions


def test_pack_image(lib_built):
    from rmem_ocu_b200 import ops
    torch.manual_seed(13)
    img = torch.randn(1, 3, 37, 53, device="cuda")
    out = torch.empty(1, 8, 37, 53, device="cuda", dtype=torch.bfloat16).contiguous(memory_format=torch.channels_last)
    ops.pack_image(img, out)
    torch.cuda.synchronize()
    assert torch.equal(out[:, :3], img.to(torch.bfloat16))
    assert (out[:, 3:] == 0).all()


@pytest.mark.parametrize("dt", [torch.bfloat16, torch.float16])
@pytest.mark.parametrize("C,H,W", [(64, 240, 432), (64, 241, 433), (8, 5, 7)])
def test_maxpool3s2_nhwc(lib_built, C, H, W, dt):
    """Encoder stem MaxPool2d(3, 2, 1) on bf16 / fp16 channels-last: bit-exact against torch (a max is exact)."""
    from rmem_ocu_b200 import ops
    torch.manual_seed(5)
    x = torch.randn(1, C, H, W, device="cuda").to(dt).contiguous(memory_format=torch.channels_last)
    ref = F.max_pool2d(x.float(), 3, 2, 1).to(dt)
    out = torch.empty_like(ref).contiguous(memory_format=torch.channels_last)
    ops.maxpool3s2_nhwc(x, out)
    assert torch.equal(out, ref)
    # fused form: relu(maxpool(x) + bias) == maxpool(relu(x + bias)) up to the final 16-bit rounding
    bias = torch.randn(C, device="cuda")
    ops.maxpool3s2_nhwc(x, out, bias, True)
    ref2 = F.max_pool2d(F.relu(x.float() + bias.view(1, -1, 1, 1)), 3, 2, 1)
    assert torch.equal(out, ref2.to(dt))


@pytest.mark.parametrize("dt", [torch.bfloat16, torch.float16])
def test_pack_image_s2d_u8_frames(lib_built, dt):
    """Decoded u8 frames [H, W, 3]: the dataloader's normalisation (video_transforms.py:537-539) happens inside the
    packing kernel; equal to packing the normalised f32 planes up to one rounding of the 16-bit output."""
    from rmem_ocu_b200 import ops
    from rmem_ocu_b200.engine import InferenceResNet
    torch.manual_seed(19)
    H, W = 121, 75
    u8 = torch.randint(0, 256, (1, H, W, 3), device="cuda", dtype=torch.uint8)
    mean, std = InferenceResNet.U8_MEAN, InferenceResNet.U8_STD
    planes = (u8.permute(0, 3, 1, 2).float() / 255.0 - torch.tensor(mean, device="cuda").view(1, 3, 1, 1)) \
        / torch.tensor(std, device="cuda").view(1, 3, 1, 1)
    Hs, Ws = (H + 1) // 2, (W + 1) // 2
    a = torch.empty(1, Hs + 3, Ws + 3, 16, device="cuda", dtype=dt).permute(0, 3, 1, 2)
    b = torch.empty_like(a)
    ops.pack_image_s2d(u8, a, mean, std)
    ops.pack_image_s2d(planes.contiguous(), b)
    torch.cuda.synchronize()
    eps = 2.0 ** (-8 if dt == torch.bfloat16 else -11)
    assert (a.float() - b.float()).abs().max().item() <= eps * b.float().abs().max().item()
    assert (a[:, 12:] == 0).all() and (a[:, :, :2] == 0).all() and (a[:, :, :, :2] == 0).all()


def test_linear_fp16_operands(lib_built):
    """rmem_linear_ex with IEEE fp16 A / W (the decoder adapters read the fp16 backbone's feature maps)."""
    from rmem_ocu_b200 import ops
    torch.manual_seed(23)
    M, K, N = 1620, 1024, 256
    a = torch.randn(M, K, device="cuda").half()
    w = (torch.randn(N, K, device="cuda") * 0.05).half()
    bias = torch.randn(N, device="cuda")
    out = torch.empty(M, N, device="cuda")
    ops.linear(a, w, bias, [(0, N, ops.ACT_NONE, ops.OUT_F32, out, 0)])
    torch.cuda.synchronize()
    ref = a.float() @ w.float().t() + bias
    assert _rel(out, ref) < 1e-5                    # exact products, f32 accumulation
    with pytest.raises(Exception):
        ops.linear(a, w.to(torch.bfloat16), bias, [(0, N, ops.ACT_NONE, ops.OUT_F32, out, 0)])


@pytest.mark.parametrize("H,W", [(480, 864), (481, 865), (37, 54)])
def test_pack_image_s2d_and_stem_equivalence(lib_built, H, W):
    """Space-to-depth packing is bit-exact against a torch construction, and the 4x4 stride-1 convolution on it
    equals the reference's 7x7 stride-2 stem convolution (encoders/resnet.py:150-151) on the same bf16 inputs."""
    from rmem_ocu_b200 import ops
    from rmem_ocu_b200.engine import InferenceResNet
    from rmem_ocu_b200.model import DeAOTModel, RMemConfig, deterministic_init_
    torch.manual_seed(17)
    img = torch.randn(1, 3, H, W, device="cuda")
    Hs, Ws = (H + 1) // 2, (W + 1) // 2
    out = torch.empty(1, Hs + 3, Ws + 3, 16, device="cuda", dtype=torch.bfloat16).permute(0, 3, 1, 2)
    ops.pack_image_s2d(img, out)
    ref = torch.zeros(1, 16, Hs + 3, Ws + 3, device="cuda")
    for dy in range(2):
        for dx in range(2):
            sub = img[:, :, dy::2, dx::2]
            ch = (dy * 2 + dx) * 3
            ref[:, ch:ch + 3, 2:2 + sub.shape[2], 2:2 + sub.shape[3]] = sub
    assert torch.equal(out.float(), ref.to(torch.bfloat16).float())
    m = DeAOTModel(RMemConfig(latter_mem_len=1))
    deterministic_init_(m.state_dict(), 3)
    enc = InferenceResNet(m.encoder, m.encoder_projector, torch.device("cuda"), torch.bfloat16)
    assert enc.conv1_s2d is not None
    w8, b8, stride, pad = enc.conv1
    with torch.no_grad():
        y_s2d = F.relu(F.conv2d(out.float(), enc.conv1_s2d[0].float(), enc.conv1_s2d[1].float()))
        x8 = torch.zeros(1, 8, H, W, device="cuda")
        x8[:, :3] = img.to(torch.bfloat16).float()
        y_ref = F.relu(F.conv2d(x8, w8.float(), b8.float(), stride, pad))
    assert y_s2d.shape == y_ref.shape
    assert _rel(y_s2d, y_ref) < 1e-5
