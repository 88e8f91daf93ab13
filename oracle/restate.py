"""TEST INFRASTRUCTURE ONLY -- CPU restatement (PyTorch fp32, eager) of the reference algorithm for
the RMem per-frame memory-reading hot path of AOT / DeAOT (reference: Bardli/RMem_ocu).

It is the oracle the CUDA path is checked against and the `cpu_baseline` ("port") that bench.py
times.  Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs
may import it; nothing under rmem_ocu_b200/ does.

Parity status: the reference ships no golden vectors / known-answer tests for this path
(SURVEY.md section 4), so by the reference's own tests parity is UNPINNED.  This restatement is
instead pinned against outputs of the reference itself: `oracle/make_golden.py` imports the
unmodified reference (Python, CPU) in the build container and commits its outputs as fixtures
under tests/golden/, and tests/test_oracle_vs_reference.py re-checks live whenever
/root/reference is present.

Everything is functional over a plain ``state_dict`` (reference key names);
file:line citations are relative to /root/reference/aot_plus/.
"""
import math

import torch
import torch.nn.functional as F

# ------------------------------------------------------------------------------------------------
# small pieces


def silu(x):
    # networks/layers/attention.py:89-90
    return x * torch.sigmoid(x)


def linear(x, sd, name):
    return F.linear(x, sd[name + ".weight"], sd.get(name + ".bias"))


def layer_norm(x, sd, name):
    w = sd[name + ".weight"]
    return F.layer_norm(x, (w.numel(),), w, sd[name + ".bias"], 1e-5)


def group_norm_1d(x, sd, name, groups):
    # networks/layers/basic.py:6-12 (GroupNorm over [B, C, L] of an [L, B, C] tensor)
    y = F.group_norm(x.permute(1, 2, 0), groups, sd[name + ".gn.weight"], sd[name + ".gn.bias"], 1e-5)
    return y.permute(2, 0, 1)


def seq_to_2d(x, size_2d):
    # networks/layers/basic.py:73-77
    h, w = size_2d
    _, n, c = x.shape
    return x.view(h, w, n, c).permute(2, 3, 0, 1).contiguous()


def dw_conv5(x, weight, size_2d):
    # networks/layers/basic.py:38-57 (DWConv2d; dropout is identity in eval)
    h, w = size_2d
    _, bs, c = x.shape
    y = x.view(h, w, bs, c).permute(2, 3, 0, 1)
    y = F.conv2d(y, weight, None, 1, 2, 1, c)
    return y.reshape(bs, c, h * w).permute(2, 0, 1)


def one_hot_mask(mask, cls_num):
    # utils/image.py:69-74
    if mask.dim() == 3:
        mask = mask.unsqueeze(1)
    idx = torch.arange(0, cls_num + 1, device=mask.device).view(1, -1, 1, 1)
    return (mask == idx).float(), (mask == 255).float()


def position_embedding_sine(h, w, num_pos_feats=128, temperature=10000.0, device=None):
    # networks/layers/position.py:35-79 with normalize=True, scale=2*pi (models/aot.py:87-90)
    gy, gx = torch.meshgrid(torch.arange(h, device=device), torch.arange(w, device=device), indexing="ij")
    y_embed = gy.unsqueeze(0).float()
    x_embed = gx.unsqueeze(0).float()
    eps = 1e-6
    scale = 2 * math.pi
    y_embed = y_embed / (y_embed[:, -1:, :] + eps) * scale
    x_embed = x_embed / (x_embed[:, :, -1:] + eps) * scale
    dim_t = torch.arange(num_pos_feats, dtype=torch.float32, device=device)
    dim_t = temperature ** (2 * torch.div(dim_t, 2, rounding_mode="trunc") / num_pos_feats)
    pos_x = x_embed[:, :, :, None] / dim_t
    pos_y = y_embed[:, :, :, None] / dim_t
    pos_x = torch.stack((pos_x[:, :, :, 0::2].sin(), pos_x[:, :, :, 1::2].cos()), dim=4).flatten(3)
    pos_y = torch.stack((pos_y[:, :, :, 0::2].sin(), pos_y[:, :, :, 1::2].cos()), dim=4).flatten(3)
    return torch.cat((pos_y, pos_x), dim=3).permute(0, 3, 1, 2)  # [1, 2*npf, h, w]


# ------------------------------------------------------------------------------------------------
# temporal positional embedding (A7)


def tpe_table(mem_pos_emb, T):
    """Rows added to bank position t = 0..T-1 (networks/layers/transformer.py:594-623 / 1140-1169).

    mem_pos_emb: [R, C] (R = 4 with TEMPORAL_POSITIONAL_EMBEDDING_SLOT_4).  Returns [T, C].
    """
    R, C = mem_pos_emb.shape
    max_T = 4
    pe = mem_pos_emb
    if T <= R:
        pe = pe[:T]
    if T == 1:
        return pe[0].view(1, C)
    x = pe.clone().permute(1, 0).view(1, C, -1)
    if T <= max_T:
        x = F.interpolate(x, size=T, mode="linear", align_corners=True)
    else:
        x = F.interpolate(x, size=max_T, mode="linear", align_corners=True)
        x = torch.flip(x, dims=(-1,))
        x = F.interpolate(x, size=T, mode="nearest")
        x = torch.flip(x, dims=(-1,))
    return x.view(C, T).permute(1, 0).contiguous()


def tpe_row_indices(T, R=4):
    """Which mem_pos_emb row each bank position uses (valid because for T <= 4 the linear resample
    is an identity and for T > 4 the path is 4 rows -> nearest).  Derived by pushing a row-index
    ramp through tpe_table, so it follows whatever F.interpolate does."""
    ramp = torch.arange(R, dtype=torch.float32).view(R, 1)
    return [int(round(v)) for v in tpe_table(ramp, T).view(-1).tolist()]


# ------------------------------------------------------------------------------------------------
# attention operators


def multihead_attention(Q, K, V, sd, prefix, num_head=8, use_linear=True):
    """MultiheadAttention.forward (networks/layers/attention.py:28-81); returns (out, attn)."""
    if use_linear:
        Q = linear(Q, sd, prefix + ".linear_Q")
        K = linear(K, sd, prefix + ".linear_K")
        V = linear(V, sd, prefix + ".linear_V")
    bs = Q.shape[1]
    d_model = Q.shape[2]
    hd = d_model // num_head
    T_ = (d_model / num_head) ** 0.5
    Q = Q / T_
    Qh = Q.view(-1, bs, num_head, hd).permute(1, 2, 0, 3)
    Kh = K.view(-1, bs, num_head, hd).permute(1, 2, 3, 0)
    Vh = V.view(-1, bs, num_head, hd).permute(1, 2, 0, 3)
    attn = torch.softmax(Qh @ Kh, dim=-1)
    out = (attn @ Vh).permute(2, 0, 1, 3).reshape(-1, bs, d_model)
    out = linear(out, sd, prefix + ".projection")
    return out, attn


def gated_propagation(Q, K, V, U, size_2d, sd, prefix, d_att=128, use_linear=False):
    """GatedPropagation.forward with num_head = 1 (networks/layers/attention.py:139-211).

    Returns (out [L,B,d_vu], attn [B,1,Lq,Lk], pre_gate [L,B,expand_d_vu])."""
    l, bs, _ = Q.shape
    if use_linear:
        Q = K = linear(Q, sd, prefix + ".linear_QK")
        half = V.shape[-1] // 2
        V = silu(torch.cat([linear(V[..., :half], sd, prefix + ".linear_V1"),
                            linear(V[..., half:], sd, prefix + ".linear_V2")], dim=-1))
        U = silu(torch.cat([linear(U[..., :half], sd, prefix + ".linear_U1"),
                            linear(U[..., half:], sd, prefix + ".linear_U2")], dim=-1))
    Qs = Q / (d_att ** 0.5)
    Qh = Qs.view(-1, bs, 1, d_att).permute(1, 2, 0, 3)
    Kh = K.view(-1, bs, 1, d_att).permute(1, 2, 3, 0)
    Vh = V.view(-1, bs, 1, V.shape[-1]).permute(1, 2, 0, 3)
    attn = torch.softmax(Qh @ Kh, dim=-1)
    agg = (attn @ Vh).permute(2, 0, 1, 3).reshape(l, bs, -1)
    gated = agg * U
    out = dw_conv5(gated, sd[prefix + ".dw_conv.conv.weight"], size_2d)
    out = linear(out, sd, prefix + ".projection")
    return out, attn, agg


def local_gated_propagation(q, k, v, u, size_2d, sd, prefix, d_att=128, max_dis=7):
    """LocalGatedPropagation.forward, num_head = 1, dilation = 1, use_linear = False
    (networks/layers/attention.py:289-361), restated as a direct windowed gather instead of
    unfold + local2global scatter + dense matmul (attention.py:331-335, 363-402).

    q, k: [1, d_att, h, w]; v: [1, C, h, w]; u: [L, 1, C].  Returns (out [L,1,d_vu],
    local_attn [1,1,225,L], agg [L,1,C])."""
    n, c, h, w = v.shape
    win = 2 * max_dis + 1
    rel = F.conv2d(q, sd[prefix + ".relative_emb_k.weight"], sd[prefix + ".relative_emb_k.bias"])
    rel = rel.view(n, win * win, h * w)                       # bias from the UNSCALED q (:314)
    qs = q / (d_att ** 0.5)                                   # (:317)
    kp = F.pad(k, (max_dis,) * 4)
    vp = F.pad(v, (max_dis,) * 4)
    ones = F.pad(torch.ones(1, 1, h, w, device=q.device), (max_dis,) * 4)
    qk = torch.empty(n, win * win, h, w, device=q.device)
    inside = torch.empty(1, win * win, h, w, device=q.device)
    for dy in range(win):
        for dx in range(win):
            t = dy * win + dx                                 # F.unfold tap order (:409-412)
            qk[:, t] = (qs * kp[:, :, dy:dy + h, dx:dx + w]).sum(1)
            inside[:, t] = ones[:, 0, dy:dy + h, dx:dx + w]
    qk = qk.view(n, win * win, h * w) + rel
    qk = qk - (1 - inside.view(1, win * win, h * w)) * 1e8    # (:344), fp32 constant
    attn = torch.softmax(qk, dim=1)                           # over the 225 taps (:346)
    a4 = attn.view(n, win * win, h, w)
    agg = torch.zeros(n, c, h, w, device=q.device)
    for dy in range(win):
        for dx in range(win):
            t = dy * win + dx
            agg = agg + a4[:, t:t + 1] * vp[:, :, dy:dy + h, dx:dx + w]
    agg = agg.view(n, c, h * w).permute(2, 0, 1)              # [L, n, C]
    gated = agg * u
    out = dw_conv5(gated, sd[prefix + ".dw_conv.conv.weight"], size_2d)
    out = linear(out, sd, prefix + ".projection")
    return out, attn.view(n, 1, win * win, h * w), agg


def record_attn_weight(attn, L):
    # networks/layers/transformer.py:636-643 / 1185-1192: mean over heads, sum over each frame's keys
    bs, head, hw, thw = attn.shape
    T = thw // hw
    a = attn.view(bs, head, hw, T, hw).mean(dim=1).squeeze(0)
    return a.sum(dim=2)  # [L, T]


# ------------------------------------------------------------------------------------------------
# eviction rule (A9 / "R")


class EvictionState:
    """stored_attn_weight_dict / stored_frame_times of transformer.py:438-443, 993-998."""

    def __init__(self):
        self.ema = {}
        self.times = {}


def eviction_scores(W, fg, idx, state):
    """Steps 1-6 of restrict_long_memories (transformer.py:338-411 AOT, 891-964 DeAOT).

    W: [L, T_old] fp32 layer-0 record_attn_weight; fg: [L] foreground probability or None;
    idx: long_memories_indexes (len T_old + 1, newest last).  Mutates `state`.
    Returns (drop_idx, detail dict)."""
    T_old = W.shape[1]
    W = W.float().cpu()                       # the reference synchronises here too (transformer.py:353 / 906)
    if fg is not None:
        fg = fg.float().cpu()
    attn_weight = torch.stack([W, W]).mean(dim=0)
    if fg is not None:
        attn_weight = attn_weight * fg.reshape(-1, 1)
    attn_weight = attn_weight.sum(dim=0)
    attn_weight = attn_weight / attn_weight.sum()
    r_raw = attn_weight.clone()
    cur = {idx[i]: attn_weight[i] for i in range(T_old)}
    f = 0.8
    cur = {k: ((1 - f) * state.ema[k] + f * a) if k in state.ema else a for k, a in cur.items()}
    state.ema = cur
    for i in range(T_old):
        attn_weight[i] = cur[idx[i]]
    r_ema = attn_weight.clone()
    times = {k: 1 for k in idx}
    times = {k: (t + state.times[k]) if k in state.times else t for k, t in times.items()}
    state.times = times
    n = torch.Tensor([times[k] for k in idx[:-1]])
    n[0] = len(n)
    bonus = 1.5 * torch.sqrt(torch.log(n.sum()) / (n + 8))
    o = attn_weight + bonus
    drop = 1
    rest = o[1:]
    if rest.numel() > 0:
        drop = int(torch.argmin(rest).item()) + 1
    margin = float("inf")
    if rest.numel() > 1:
        srt = torch.sort(rest).values
        margin = float(srt[1] - srt[0])
    return drop, dict(r_raw=r_raw, r_ema=r_ema, n=n, bonus=bonus, o=o, margin=margin)


# ------------------------------------------------------------------------------------------------
# encoder / decoder / id bank (black boxes of the path; restated so the oracle is self-contained)


def _frozen_bn(x, sd, p):
    # networks/layers/normalization.py:6-43 (FrozenBatchNorm2d, eval branch, eps 1e-5)
    return F.batch_norm(x, sd[p + ".running_mean"], sd[p + ".running_var"], sd[p + ".weight"],
                        sd[p + ".bias"], False, 0.0, 1e-5)


def _bottleneck(x, sd, p, stride, dilation, has_down):
    # networks/encoders/resnet.py:10-69
    out = F.relu(_frozen_bn(F.conv2d(x, sd[p + ".conv1.weight"]), sd, p + ".bn1"))
    out = F.conv2d(out, sd[p + ".conv2.weight"], None, stride, dilation, dilation)
    out = F.relu(_frozen_bn(out, sd, p + ".bn2"))
    out = _frozen_bn(F.conv2d(out, sd[p + ".conv3.weight"]), sd, p + ".bn3")
    if has_down:
        x = _frozen_bn(F.conv2d(x, sd[p + ".downsample.0.weight"], None, stride), sd, p + ".downsample.1")
    return F.relu(out + x)


def resnet50_s16(img, sd, prefix="encoder"):
    # networks/encoders/resnet.py:71-194: stride 16, layer4 dropped, strides [1,2,2], dilation 1
    x = F.conv2d(img, sd[prefix + ".conv1.weight"], None, 2, 3)
    x = F.relu(_frozen_bn(x, sd, prefix + ".bn1"))
    x = F.max_pool2d(x, 3, 2, 1)
    xs = []
    for li, (blocks, stride) in enumerate([(3, 1), (4, 2), (6, 2)], start=1):
        for b in range(blocks):
            p = "%s.layer%d.%d" % (prefix, li, b)
            x = _bottleneck(x, sd, p, stride if b == 0 else 1, 1, b == 0)
        xs.append(x)
    xs.append(x)
    return xs


def encode_image(img, sd):
    # networks/models/aot.py:116-134
    xs = resnet50_s16(img, sd)
    xs[-1] = F.conv2d(xs[-1], sd["encoder_projector.weight"], sd["encoder_projector.bias"])
    return xs


def _conv_gn(x, sd, p, k, groups=8):
    # networks/layers/basic.py:60-70
    y = F.conv2d(x, sd[p + ".conv.weight"], sd[p + ".conv.bias"], 1, k // 2)
    return F.group_norm(y, groups, sd[p + ".gn.weight"], sd[p + ".gn.bias"], 1e-5)


def fpn_decoder(inputs, shortcuts, sd, intermediate, prefix="decoder"):
    # networks/decoders/fpn.py:36-68 (align_corners=True)
    x = torch.cat(inputs, dim=1) if intermediate else inputs[-1]
    x = F.relu(_conv_gn(x, sd, prefix + ".conv_in", 1))
    a16 = F.conv2d(shortcuts[-2], sd[prefix + ".adapter_16x.weight"], sd[prefix + ".adapter_16x.bias"])
    x = F.relu(_conv_gn(a16 + x, sd, prefix + ".conv_16x", 3))
    x = F.interpolate(x, size=shortcuts[-3].shape[-2:], mode="bilinear", align_corners=True)
    a8 = F.conv2d(shortcuts[-3], sd[prefix + ".adapter_8x.weight"], sd[prefix + ".adapter_8x.bias"])
    x = F.relu(_conv_gn(a8 + x, sd, prefix + ".conv_8x", 3))
    x = F.interpolate(x, size=shortcuts[-4].shape[-2:], mode="bilinear", align_corners=True)
    a4 = F.conv2d(shortcuts[-4], sd[prefix + ".adapter_4x.weight"], sd[prefix + ".adapter_4x.bias"])
    x = F.relu(_conv_gn(a4 + x, sd, prefix + ".conv_4x", 3))
    return F.conv2d(x, sd[prefix + ".conv_out.weight"], sd[prefix + ".conv_out.bias"])


def get_id_emb(one_hot12, sd, deaot):
    # networks/models/aot.py:111-114, deaot.py:65-69 (17x17 stride-16 conv; DeAOT adds LayerNorm)
    e = F.conv2d(one_hot12, sd["patch_wise_id_bank.weight"], sd["patch_wise_id_bank.bias"], 16, 8)
    if deaot:
        e = F.layer_norm(e.permute(2, 3, 0, 1), (e.shape[1],), sd["id_norm.weight"], sd["id_norm.bias"],
                         1e-5).permute(2, 3, 0, 1)
    return e


# ------------------------------------------------------------------------------------------------
# LSTT (DeAOT: DualBranchGPM / GatedPropagationModule; AOT: LongShortTermTransformer / block)


class LSTT:
    """State + forward of the LSTT, both model families.  Mirrors transformer.py:133-464 (AOT) and
    :700-1007 (DeAOT) including the bank lists (torch.cat semantics) and the eviction statistics."""

    def __init__(self, sd, deaot, num_layers=3, d_model=256):
        self.sd = sd
        self.deaot = deaot
        self.num_layers = num_layers
        self.d_model = d_model
        self.clear_memory()
        self.record = [None] * num_layers   # per-layer record_attn_weight
        self.trace = None                   # optional dict for per-layer intermediates

    # -- memory management ---------------------------------------------------------------
    def clear_memory(self):
        self.curr = self.long = self.short = None
        self.long_term_memories = None
        self.short_term_memories = None
        self.evict = EvictionState()

    def init_memory(self):
        # transformer.py:438-443 / 993-998
        self.long_term_memories = [list(m) for m in self.long]
        self.short_term_memories = [list(m) for m in self.short]
        self.evict = EvictionState()

    # -- blocks ----------------------------------------------------------------------------
    def _tpe(self, global_K, curr_Q, temporal_encoding):
        # transformer.py:594-629 / 1140-1175
        T = global_K.shape[0]
        if temporal_encoding is None:
            return global_K.flatten(0, 1), curr_Q
        C = global_K.shape[-1]
        cur, mem = temporal_encoding[0:1], temporal_encoding[1:]
        pe = tpe_table(mem, T)
        Kf = (global_K + pe.view(T, 1, 1, C)).flatten(0, 1)
        return Kf, curr_Q + cur.view(1, 1, C)

    def _deaot_block(self, i, tgt, tgt_id, long_mem, short_mem, curr_id_emb, size_2d, tpe, save):
        sd = self.sd
        p = "LSTT.layers.%d" % i
        d_att, d_model = 128, self.d_model
        _tgt = layer_norm(tgt, sd, p + ".norm1")
        curr_QV = linear(_tgt, sd, p + ".linear_QV")
        curr_Q = curr_K = curr_QV[..., :d_att]
        local_Q = seq_to_2d(curr_Q, size_2d)
        curr_V = silu(curr_QV[..., d_att:])
        curr_U = linear(_tgt, sd, p + ".linear_U")
        if tgt_id is None:
            tgt_id = 0
            cat_curr_U = torch.cat([silu(curr_U), torch.ones_like(curr_U)], dim=-1)
            curr_ID_V = None
        else:
            _tgt_id = layer_norm(tgt_id, sd, p + ".id_norm1")
            curr_ID_V = _tgt_id
            curr_ID_U = linear(_tgt_id, sd, p + ".linear_ID_U")
            cat_curr_U = silu(torch.cat([curr_U, curr_ID_U], dim=-1))
        if curr_id_emb is not None:
            global_K, global_V = curr_K, curr_V
            local_K = seq_to_2d(global_K, size_2d)
            local_V = seq_to_2d(global_V, size_2d)
            global_ID_V = self.fuse_key_value_id(i, curr_ID_V, curr_id_emb)
            local_ID_V = seq_to_2d(global_ID_V, size_2d)
            global_K, global_V, global_ID_V = global_K[None], global_V[None], global_ID_V[None]
        else:
            global_K, global_V, _, global_ID_V = long_mem
            local_K, local_V, _, local_ID_V = short_mem
        Kf, Qt = self._tpe(global_K, curr_Q, tpe)
        cat_global_V = torch.cat([global_V.flatten(0, 1), global_ID_V.flatten(0, 1)], dim=-1)
        cat_local_V = torch.cat([local_V, local_ID_V], dim=1)
        cat_tgt2, attn, lt_agg = gated_propagation(Qt, Kf, cat_global_V, cat_curr_U, size_2d, sd,
                                                   p + ".long_term_attn", d_att)
        if save:
            self.record[i] = record_attn_weight(attn, tgt.shape[0])
        cat_tgt3, _, st_agg = local_gated_propagation(local_Q, local_K, cat_local_V, cat_curr_U, size_2d,
                                                      sd, p + ".short_term_attn", d_att)
        tgt2, tgt_id2 = cat_tgt2[..., :d_model], cat_tgt2[..., d_model:]
        tgt3, tgt_id3 = cat_tgt3[..., :d_model], cat_tgt3[..., d_model:]
        tgt = tgt + (tgt2 + tgt3)
        tgt_id = tgt_id + (tgt_id2 + tgt_id3)
        _tgt = layer_norm(tgt, sd, p + ".norm2")
        _tgt_id = layer_norm(tgt_id, sd, p + ".id_norm2")
        x = torch.cat([_tgt, _tgt_id], dim=-1)
        cat_sa, _, _ = gated_propagation(x, x, x, x, size_2d, sd, p + ".self_attn", d_att, use_linear=True)
        tgt = tgt + cat_sa[..., :d_model]
        tgt_id = tgt_id + cat_sa[..., d_model:]
        if self.trace is not None:
            self.trace.setdefault("layers", []).append(dict(
                curr_Q=curr_Q, curr_V=curr_V, cat_curr_U=cat_curr_U, lt_agg=lt_agg, st_agg=st_agg,
                cat_tgt2=cat_tgt2, cat_tgt3=cat_tgt3, cat_sa=cat_sa, tgt=tgt, tgt_id=tgt_id))
        return tgt, tgt_id, [[curr_K, curr_V, None, curr_ID_V],
                             [global_K, global_V, None, global_ID_V],
                             [local_K, local_V, None, local_ID_V]]

    def fuse_key_value_id(self, i, value, id_emb):
        # transformer.py:1238-1244
        p = "LSTT.layers.%d.linear_ID_V" % i
        if value is not None:
            return silu(linear(torch.cat([value, id_emb], dim=2), self.sd, p))
        return silu(linear(id_emb, self.sd, p))

    def _aot_block(self, i, tgt, long_mem, short_mem, curr_id_emb, self_pos, size_2d, tpe, save):
        # SimplifiedTransformerBlock.forward, transformer.py:553-692 (MODEL_LINEAR_Q = False)
        sd = self.sd
        p = "LSTT.layers.%d" % i
        _tgt = layer_norm(tgt, sd, p + ".norm1")
        qk = _tgt if self_pos is None else _tgt + self_pos
        tgt2, _ = multihead_attention(qk, qk, _tgt, sd, p + ".self_attn", 8, True)
        tgt = tgt + tgt2
        _tgt = layer_norm(tgt, sd, p + ".norm2")
        curr_Q = linear(_tgt, sd, p + ".linear_Q")
        curr_K = curr_Q
        curr_V = _tgt
        if curr_id_emb is not None:
            global_K = curr_K
            global_V = linear(curr_V + curr_id_emb, sd, p + ".linear_V")
            local_K, local_V = global_K, global_V
            global_K, global_V = global_K[None], global_V[None]
        else:
            global_K, global_V = long_mem
            local_K, local_V = short_mem
        Kf, Qt = self._tpe(global_K, curr_Q, tpe)
        tgt2, attn = multihead_attention(Qt, Kf, global_V.flatten(0, 1), sd, p + ".long_term_attn", 8, False)
        if save:
            self.record[i] = record_attn_weight(attn, tgt.shape[0])
        tgt3, _ = multihead_attention(curr_Q, layer_norm(local_K + curr_K, sd, p + ".norm4"),
                                      layer_norm(local_V + curr_V, sd, p + ".norm4"), sd,
                                      p + ".short_term_attn", 8, False)
        new_local_K = linear(tgt3, sd, p + ".linear_QMem")
        new_local_V = tgt3
        if curr_id_emb is not None:
            new_local_V = linear(new_local_V + curr_id_emb, sd, p + ".linear_VMem")
        tgt = tgt + tgt2 + tgt3
        _tgt = layer_norm(tgt, sd, p + ".norm3")
        y = linear(_tgt, sd, p + ".linear1")
        # GNActDWConv2d (basic.py:15-35): GroupNorm(32) -> GELU -> 5x5 depthwise
        h, w = size_2d
        _, bs, c = y.shape
        y2 = y.view(h, w, bs, c).permute(2, 3, 0, 1)
        y2 = F.group_norm(y2, 32, sd[p + ".activation.gn.weight"], sd[p + ".activation.gn.bias"], 1e-5)
        y2 = F.conv2d(F.gelu(y2), sd[p + ".activation.conv.weight"], None, 1, 2, 1, c)
        y = y2.reshape(bs, c, h * w).permute(2, 0, 1)
        tgt = tgt + linear(y, sd, p + ".linear2")
        if self.trace is not None:
            self.trace.setdefault("layers", []).append(dict(curr_Q=curr_Q, tgt2=tgt2, tgt3=tgt3, tgt=tgt))
        return tgt, [[curr_K, curr_V], [global_K, global_V], [new_local_K, new_local_V]]

    # -- forward ---------------------------------------------------------------------------
    def forward(self, tgt, curr_id_emb, self_pos, size_2d, tpe, save_atten_weights):
        sd = self.sd
        mems = []
        outs = []
        out_id = None
        output = tgt
        for i in range(self.num_layers):
            lm = self.long_term_memories[i] if self.long_term_memories is not None else None
            sm = self.short_term_memories[i] if self.short_term_memories is not None else None
            if self.deaot:
                output, out_id, m = self._deaot_block(i, output, out_id, lm, sm, curr_id_emb, size_2d, tpe,
                                                      save_atten_weights)
                outs.append(torch.cat([output, out_id], dim=2))
            else:
                output, m = self._aot_block(i, output, lm, sm, curr_id_emb, self_pos, size_2d, tpe,
                                            save_atten_weights)
                outs.append(output)
            mems.append(m)
        if self.deaot:
            # one GroupNorm1D(512, groups=2) on the last output (transformer.py:806-812;
            # intermediate_norm False => num_norms = 1)
            outs[-1] = group_norm_1d(outs[-1], sd, "LSTT.decoder_norms.0", 2)
        else:
            # 3 LayerNorms (transformer.py:248-259)
            outs[-1] = layer_norm(outs[-1], sd, "LSTT.decoder_norms.2")
            for i in range(self.num_layers - 1):
                outs[i] = layer_norm(outs[i], sd, "LSTT.decoder_norms.%d" % i)
        self.curr = [list(m[0]) for m in mems]
        self.long = [list(m[1]) for m in mems]
        self.short = [list(m[2]) for m in mems]
        return outs

    # -- write path ------------------------------------------------------------------------
    def update_short_memories(self, curr_id_emb, size_2d, is_update_long_memory):
        sd = self.sd
        if self.deaot:
            # transformer.py:826-857
            new_short = []
            for i in range(self.num_layers):
                curr_k, curr_v, _, curr_id_v = self.curr[i]
                id_v = self.fuse_key_value_id(i, curr_id_v, curr_id_emb)
                self.curr[i][3] = id_v
                new_short.append([seq_to_2d(curr_k, size_2d), seq_to_2d(curr_v, size_2d), None,
                                  seq_to_2d(id_v, size_2d)])
            self.short_term_memories = new_short
        else:
            # transformer.py:269-304
            new_short = []
            for i in range(self.num_layers):
                p = "LSTT.layers.%d" % i
                self.curr[i][1] = linear(self.curr[i][1] + curr_id_emb, sd, p + ".linear_V")
                self.short[i][1] = linear(self.short[i][1] + curr_id_emb, sd, p + ".linear_VMem")
                new_short.append([self.short[i][0], self.short[i][1]])
            self.short_term_memories = new_short
        if is_update_long_memory:
            # transformer.py:306-322 / 859-878
            upd = []
            for new_m, last_m in zip(self.curr, self.long_term_memories):
                row = []
                for new_e, last_e in zip(new_m, last_m):
                    if new_e is None or last_e is None:
                        row.append(None)
                    else:
                        row.append(torch.cat([last_e, new_e[None]], dim=0))
                upd.append(row)
            self.long_term_memories = upd

    def restrict_long_memories(self, former, latter, long_memories_indexes, fg):
        """transformer.py:324-436 (AOT) / 880-991 (DeAOT), use_atten_weight=True.
        Returns a dict describing the decision (None if the AOT variant returned early)."""
        cap = former + latter
        if not self.deaot and self.long_term_memories[0][0].shape[0] <= cap:
            return None
        fgv = fg.squeeze().flatten() if fg is not None else None
        drop, detail = eviction_scores(self.record[0], fgv, long_memories_indexes, self.evict)
        detail["drop"] = drop
        detail["dropped"] = False
        is_drop = False
        for i in range(self.num_layers):
            for j, mem in enumerate(self.long_term_memories[i]):
                if mem is None:
                    continue
                if mem.shape[0] > cap:
                    is_drop = True
                    self.long_term_memories[i][j] = torch.cat([mem[0:drop], mem[drop + 1:]], dim=0)
        if is_drop:
            detail["dropped"] = True
            long_memories_indexes.remove(long_memories_indexes[drop])
        return detail


# ------------------------------------------------------------------------------------------------
# engine (AOTEngine / DeAOTEngine state machine, aot_engine.py:18-568, single engine, B = 1)


class Engine:
    def __init__(self, sd, deaot, former=1, latter=4, long_term_mem_gap=9999, max_obj_num=10):
        self.sd = sd
        self.deaot = deaot
        self.former, self.latter = former, latter
        self.long_term_mem_gap = long_term_mem_gap
        self.max_obj_num = max_obj_num
        self.lstt = LSTT(sd, deaot)
        self.restart_engine()

    def restart_engine(self):
        # aot_engine.py:533-563
        self.frame_step = 0
        self.last_mem_step = -1
        self.obj_nums = None
        self.pos_emb = None
        self.enc_size_2d = None
        self.enc_hw = None
        self.input_size_2d = None
        self.lstt.clear_memory()
        self.long_memories_indexes = []
        self.pred_id_logits = None
        self.evict_log = []

    def _tpe(self):
        return torch.cat((self.sd["cur_pos_emb"], self.sd["mem_pos_emb"]), dim=0)

    def assign_identity(self, one_hot, ignore_mask=None):
        # aot_engine.py:208-232 with MODEL_IGNORE_TOKEN = True, no id shuffle
        if ignore_mask is None:
            ignore_mask = torch.zeros(one_hot.shape[0], 1, one_hot.shape[2], one_hot.shape[3], device=one_hot.device)
        one_hot = one_hot.clone()
        one_hot[:, 0] = one_hot[:, 0] * (ignore_mask == 0).float().squeeze()
        x = torch.cat((one_hot, ignore_mask), 1)
        return get_id_emb(x, self.sd, self.deaot).view(1, -1, self.enc_hw).permute(2, 0, 1)

    def _lstt_forward(self, embs, id_emb, save):
        x = embs[-1]
        n, c, h, w = x.shape
        tgt = x.view(n, c, h * w).permute(2, 0, 1).contiguous()   # bchw_2_lbc (utils/tensor.py:3-6)
        return self.lstt.forward(tgt, id_emb, self.pos_emb, self.enc_size_2d, self._tpe(), save)

    def add_reference_frame(self, img, mask, obj_nums, frame_step=-1, img_embs=None):
        # aot_engine.py:241-325
        self.obj_nums = obj_nums
        if frame_step == -1:
            frame_step = self.frame_step
        embs = encode_image(img, self.sd) if img_embs is None else img_embs
        one_hot, _ = one_hot_mask(mask, self.max_obj_num)
        if self.input_size_2d is None:
            self.input_size_2d = tuple(img.shape[2:]) if img is not None else None
            self.enc_size_2d = tuple(embs[-1].shape[2:])
            self.enc_hw = self.enc_size_2d[0] * self.enc_size_2d[1]
        if self.pos_emb is None:
            h, w = self.enc_size_2d
            self.pos_emb = position_embedding_sine(h, w, device=embs[-1].device).view(
                1, -1, self.enc_hw).permute(2, 0, 1)
        id_emb = self.assign_identity(one_hot)
        out = self._lstt_forward(embs, id_emb, False)
        self.last_mem_step = frame_step
        self.lstt.init_memory()
        self.long_memories_indexes.append(self.frame_step)
        self.decode_current_logits(embs, out)

    def match_propogate_one_frame(self, img=None, img_embs=None, output_size=None):
        # aot_engine.py:398-436
        self.frame_step += 1
        embs = encode_image(img, self.sd) if img_embs is None else img_embs
        out = self._lstt_forward(embs, None, True)
        return self.decode_current_logits(embs, out, output_size)

    def decode_current_logits(self, embs, lstt_out, output_size=None):
        # aot_engine.py:438-465, models/aot.py:136-142, deaot.py:57-63
        n, c, h, w = embs[-1].shape
        inputs = [embs[-1]] + [e.view(h, w, n, -1).permute(2, 3, 0, 1) for e in lstt_out]
        logits = fpn_decoder(inputs, embs, self.sd, intermediate=not self.deaot)
        for b, on in enumerate(self.obj_nums):
            logits[b, on + 1:] = -1e10
        self.pred_id_logits = logits
        if output_size is not None:
            logits = F.interpolate(logits, size=output_size, mode="bilinear", align_corners=True)
        return logits

    def update_short_term_memory(self, curr_mask):
        # aot_engine.py:327-369
        if curr_mask.dim() == 3 or curr_mask.shape[0] == 1:
            one_hot, ignore = one_hot_mask(curr_mask, self.max_obj_num)
        else:
            one_hot, ignore = curr_mask, None
        id_emb = self.assign_identity(one_hot, ignore)
        upd = False
        if self.frame_step - self.last_mem_step >= self.long_term_mem_gap:
            upd = True
            self.last_mem_step = self.frame_step
        self.lstt.update_short_memories(id_emb, self.enc_size_2d, upd)
        if upd:
            self.long_memories_indexes.append(self.frame_step)
            pl = F.interpolate(self.pred_id_logits, size=self.enc_size_2d, mode="bilinear", align_corners=True)
            fg = 1 - torch.softmax(pl, dim=1)[:, 0:1]
            d = self.lstt.restrict_long_memories(self.former, self.latter, self.long_memories_indexes, fg)
            if d is not None:
                d["frame_step"] = self.frame_step
                self.evict_log.append(d)


class InferEngine:
    """AOTInferEngine / DeAOTInferEngine (aot_engine.py:571-725, deaot_engine.py:20-56): more than
    ``max_aot_obj_num`` objects are split over several engines whose logits are merged every frame.

    One deliberate difference from the reference: there every AOTEngine of the list drives the SAME model object,
    and the memory banks hang off ``model.LSTT`` (transformer.py:306-322, 438-443), so with two object groups the
    second engine re-initialises and overwrites the first one's memories ("ownership").  That
    is an artefact of where the reference keeps its state, not part of the algorithm; here (and in the product)
    every object group owns its memories -- the single-group case, which is what every fixture pins, is identical."""

    def __init__(self, sd, deaot, former=1, latter=4, long_term_mem_gap=9999, max_aot_obj_num=10):
        self.sd, self.deaot, self.former, self.latter = sd, deaot, former, latter
        self.long_term_mem_gap = long_term_mem_gap
        self.max_aot_obj_num = max_aot_obj_num
        self.aot_engines = []
        self.obj_nums = None

    def separate_mask(self, mask):
        # aot_engine.py:604-628 (label-map branch; the probability branch is reachable with batch > 1 only)
        if mask is None:
            return [None] * len(self.aot_engines)
        if len(self.aot_engines) == 1:
            return [mask]
        out = []
        for idx in range(len(self.aot_engines)):
            start_id = idx * self.max_aot_obj_num + 1
            end_id = (idx + 1) * self.max_aot_obj_num
            fg_mask = ((mask >= start_id) & (mask <= end_id)).float()
            out.append((fg_mask * mask - start_id + 1) * fg_mask)
        return out

    def soft_logit_aggregation(self, all_logits):
        # aot_engine.py:650-673
        if len(all_logits) == 1:
            return all_logits[0]
        fg_probs, bg_probs = [], []
        for logit in all_logits:
            prob = torch.softmax(logit, dim=1)
            bg_probs.append(prob[:, 0:1])
            fg_probs.append(prob[:, 1:1 + self.max_aot_obj_num])
        bg_prob = torch.prod(torch.cat(bg_probs, dim=1), dim=1, keepdim=True)
        merged_prob = torch.cat([bg_prob] + fg_probs, dim=1).clamp(1e-5, 1 - 1e-5)
        return torch.logit(merged_prob)

    def add_reference_frame(self, img, mask, obj_nums, frame_step=-1):
        # aot_engine.py:675-702
        if isinstance(obj_nums, list):
            obj_nums = obj_nums[0]
        self.obj_nums = obj_nums
        aot_num = max(int(math.ceil(obj_nums / self.max_aot_obj_num)), 1)
        while aot_num > len(self.aot_engines):
            self.aot_engines.append(Engine(self.sd, self.deaot, self.former, self.latter, self.long_term_mem_gap,
                                           self.max_aot_obj_num))
        embs = encode_image(img, self.sd)                     # the same image for every group
        for eng, m in zip(self.aot_engines, self.separate_mask(mask)):
            eng.long_term_mem_gap = self.long_term_mem_gap
            if eng.input_size_2d is None:
                eng.input_size_2d = tuple(img.shape[2:])
                eng.enc_size_2d = tuple(embs[-1].shape[2:])
                eng.enc_hw = eng.enc_size_2d[0] * eng.enc_size_2d[1]
            eng.add_reference_frame(img, m, obj_nums=[self.max_aot_obj_num], frame_step=frame_step, img_embs=embs)
        self.input_size_2d = self.aot_engines[0].input_size_2d

    def match_propogate_one_frame(self, img=None, output_size=None):
        # aot_engine.py:704-712
        embs = encode_image(img, self.sd)
        return self.soft_logit_aggregation([eng.match_propogate_one_frame(img, img_embs=embs, output_size=output_size)
                                            for eng in self.aot_engines])

    def update_memory(self, curr_mask):
        # aot_engine.py:714-720
        for eng, m in zip(self.aot_engines, self.separate_mask(curr_mask)):
            eng.update_short_term_memory(m)


def run_clip(engine, images, label0, n_obj, output_size=None):
    """The per-sequence frame loop of Evaluator.evaluating (networks/managers/evaluator.py:391-441,
    518-523), single scale, no flip: returns the list of argmax label maps for frames 1..N-1."""
    H, W = images[0].shape[-2:]
    osz = output_size or (H, W)
    masks = []
    with torch.no_grad():
        engine.add_reference_frame(images[0], label0, obj_nums=[engine.max_obj_num], frame_step=0)
        for f in range(1, len(images)):
            logit = engine.match_propogate_one_frame(images[f], output_size=osz)
            prob = torch.softmax(logit, dim=1)
            label = torch.argmax(prob, dim=1, keepdim=True).float()
            masks.append(label[0, 0].to(torch.uint8))
            label_in = F.interpolate(label, size=engine.input_size_2d, mode="nearest")
            engine.update_short_term_memory(label_in)
    return masks
