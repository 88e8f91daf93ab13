"""Thin Python wrappers over the C ABI: torch tensors in, raw device pointers out.

PyTorch is used for device memory and streams only; every function here launches hand-written
sm_100a kernels from librmem_b200.so on the current CUDA stream.  No fallbacks.
"""
import ctypes

import torch

from . import _lib
from ._lib import (ACT_NONE, ACT_SILU, OUT_BF16, OUT_F32, OUT_F32_ACCUM, GemmSegment)  # noqa: F401


# number of librmem_b200 kernels launched through this module (bench.py reports it as gpu_launches)
LAUNCHES = 0
_KERNELS_PER_CALL = {"rmem_groupnorm_tokens": 2, "rmem_groupnorm_nhwc": 2, "rmem_groupnorm_gelu_tokens": 2}


def _count(name):
    global LAUNCHES
    LAUNCHES += _KERNELS_PER_CALL.get(name, 1)


def _p(t):
    return None if t is None else t.data_ptr()


def _stream():
    return torch.cuda.current_stream().cuda_stream


def _chk_cuda(*ts):
    for t in ts:
        if t is not None and not t.is_cuda:
            raise _lib.RmemError("rmem ops need CUDA tensors (there is no CPU path)")


# ---------------------------------------------------------------------------------------------
def linear(a, w, bias, segments, M=None):
    """C = a @ w.T + bias with a segmented epilogue.

    a: bf16 [M, K] (row stride may exceed K); w: bf16 [N, K]; bias: f32 [N] or None;
    segments: list of (col_begin, col_end, act, mode, dst_tensor, dst_col0)."""
    _chk_cuda(a, w)
    M = a.shape[0] if M is None else M
    K = a.shape[1]
    N = w.shape[0]
    if a.dtype != w.dtype or a.dtype not in (torch.bfloat16, torch.float16):
        raise _lib.RmemError("linear: a and w must both be bf16 or both fp16")
    segs = (GemmSegment * len(segments))()
    for i, (c0, c1, act, mode, dst, dcol) in enumerate(segments):
        segs[i] = GemmSegment(c0, c1, act, mode, dst.data_ptr(), dst.stride(0), dcol)
    if a.dtype == torch.float16:
        _lib.check(_lib.lib().rmem_linear_ex(_p(a), a.stride(0), _p(w), w.stride(0), _p(bias), M, N, K,
                                             ctypes.cast(segs, ctypes.c_void_p), len(segments), 1, _stream()),
                   "rmem_linear_ex")
    else:
        _lib.check(_lib.lib().rmem_linear(_p(a), a.stride(0), _p(w), w.stride(0), _p(bias), M, N, K,
                                          ctypes.cast(segs, ctypes.c_void_p), len(segments), _stream()),
                   "rmem_linear")
    _count("rmem_linear")


def layernorm(x0, g0, b0, out0, col0=0, x1=None, g1=None, b1=None, out1=None, col1=0):
    """LayerNorm(256) of one or two [L,256] streams (f32 or bf16 in), bf16 out at a column offset."""
    _chk_cuda(x0, out0)
    L, C = x0.shape
    in_bf16 = 1 if x0.dtype == torch.bfloat16 else 0
    _lib.check(_lib.lib().rmem_layernorm(_p(x0), _p(g0), _p(b0), _p(out0), out0.stride(0), col0,
                                         _p(x1), _p(g1), _p(b1), _p(out1),
                                         out1.stride(0) if out1 is not None else 0, col1, L, C, in_bf16,
                                         _stream()), "rmem_layernorm")
    _count("rmem_layernorm")


def layernorm_copy(x, gamma, beta, out, col0, copy_f32):
    """LN(x) -> bf16 out[:, col0:col0+256] and x itself as f32 into copy_f32 [L, 256]; x [L, 256] f32 / bf16 / fp16."""
    kind = {torch.float32: 0, torch.bfloat16: 1, torch.float16: 2}[x.dtype]
    if not x.is_contiguous():
        raise _lib.RmemError("layernorm_copy: x must be contiguous [L, 256]")
    L, C = x.shape
    _lib.check(_lib.lib().rmem_layernorm_copy(_p(x), kind, _p(gamma), _p(beta), _p(out), out.stride(0), col0,
                                              _p(copy_f32), L, C, _stream()), "rmem_layernorm_copy")
    _count("rmem_layernorm_copy")


def dwconv5(x, xcol0, wt, y, ycol0, h, w, C, S=1):
    """x, y: [S * h * w, ld] token rows of S stacked (h, w) grids."""
    _lib.check(_lib.lib().rmem_dwconv5_batched(_p(x), x.stride(0), xcol0, _p(wt), _p(y), y.stride(0), ycol0, h, w, C,
                                               S, _stream()), "rmem_dwconv5_batched")
    _count("rmem_dwconv5")


def groupnorm_tokens(xa, xb, gamma, beta, out, groups, workspace, S=1):
    """xa / xb / out: [S * L, C] rows of S stacked sequences (statistics per sequence; workspace: S blocks)."""
    L, Ca = xa.shape[0] // S, xa.shape[1]
    C = Ca + (xb.shape[1] if xb is not None else 0)
    kinds = {torch.bfloat16: 0, torch.float32: 1, torch.float16: 2}
    if out.dtype not in kinds:
        raise _lib.RmemError("groupnorm_tokens: out must be bf16, f32 or fp16")
    kind = kinds[out.dtype]
    if S == 1:
        _lib.check(_lib.lib().rmem_groupnorm_tokens(_p(xa), Ca, _p(xb), _p(gamma), _p(beta), _p(out), kind, L, C, groups,
                                                    _p(workspace), _stream()), "rmem_groupnorm_tokens")
    else:
        _lib.check(_lib.lib().rmem_groupnorm_tokens_batched(_p(xa), Ca, _p(xb), _p(gamma), _p(beta), _p(out), kind, L, C,
                                                            groups, S, _p(workspace), _stream()),
                   "rmem_groupnorm_tokens_batched")
    _count("rmem_groupnorm_tokens")


def groupnorm_nhwc(x, gamma, beta, y, groups, relu, workspace, pre_bias=None, S=1):
    """GroupNorm(+ReLU) of (x + pre_bias), a channels-last [S, C, H, W] f32 tensor viewed as [S*H*W, C]; statistics
    per image, workspace: S consecutive (2 * groups + 1)-double blocks."""
    P, C = x.shape[0] // S, x.shape[1]
    if x.dtype != y.dtype or x.dtype not in (torch.float32, torch.float16):
        raise _lib.RmemError("groupnorm_nhwc: x and y must both be fp32 or both fp16")
    _lib.check(_lib.lib().rmem_groupnorm_nhwc_ex(_p(x), _p(pre_bias), _p(gamma), _p(beta), _p(y), P, C, groups,
                                                 1 if relu else 0, S, 1 if x.dtype == torch.float16 else 0,
                                                 _p(workspace), _stream()), "rmem_groupnorm_nhwc_ex")
    _count("rmem_groupnorm_nhwc")


def upsample_add_nhwc(x, hi, wi, a, out, ho, wo, a_bias=None, S=1):
    """out = a + a_bias + bilinear_align_corners(x); x [S*hi*wi, C] and out [S*ho*wo, C] fp32 or fp16, a f32."""
    C = x.shape[1]
    if x.dtype != out.dtype or x.dtype not in (torch.float32, torch.float16) or (a is not None and a.dtype != torch.float32):
        raise _lib.RmemError("upsample_add_nhwc: x and out both fp32 or both fp16, a fp32")
    _lib.check(_lib.lib().rmem_upsample_add_nhwc_ex(_p(x), hi, wi, _p(a), _p(a_bias), _p(out), ho, wo, C, S,
                                                    1 if x.dtype == torch.float16 else 0, _stream()),
               "rmem_upsample_add_nhwc_ex")
    _count("rmem_upsample_add_nhwc")


def gn_upsample_add_nhwc(x, pre_bias, gamma, beta, hi, wi, a, out, ho, wo, workspace, a_bias=None, groups=8, S=1):
    """out = a + a_bias + bilinear_align_corners(relu(groupnorm(x + pre_bias))): x [S*hi*wi, C] raw conv output and out
    [S*ho*wo, C] fp32 or fp16, a f32; statistics per image."""
    C = x.shape[1]
    if x.dtype != out.dtype or x.dtype not in (torch.float32, torch.float16) or (a is not None and a.dtype != torch.float32):
        raise _lib.RmemError("gn_upsample_add_nhwc: x and out both fp32 or both fp16, a fp32")
    _lib.check(_lib.lib().rmem_gn_upsample_add_nhwc(_p(x), _p(pre_bias), _p(gamma), _p(beta), hi, wi, C, groups, _p(a),
                                                    _p(a_bias), _p(out), ho, wo, S, 1 if x.dtype == torch.float16 else 0,
                                                    _p(workspace), _stream()), "rmem_gn_upsample_add_nhwc")
    global LAUNCHES
    LAUNCHES += 2


def finish_logits(conv, bias, pred, n_valid, out=None):
    """conv: f32 [S, C, Hs, Ws] (any strides) -> pred f32 [S, C, Hs, Ws] contiguous (+ bias, channels >= n_valid[s]
    set to -1e10) and, if given, out f32 [S, C, Ho, Wo] = bilinear align_corners resize of pred.  n_valid: an int
    (S = 1) or a list of S ints."""
    S, C, Hs, Ws = conv.shape
    Ho, Wo = (out.shape[-2], out.shape[-1]) if out is not None else (0, 0)
    nv = [int(n_valid)] * S if not isinstance(n_valid, (list, tuple)) else [int(v) for v in n_valid]
    for s0 in range(0, S, 64):                    # the kernel takes up to 64 images per call
        s1 = min(S, s0 + 64)
        arr = (ctypes.c_int32 * (s1 - s0))(*nv[s0:s1])
        _lib.check(_lib.lib().rmem_finish_logits_batched(_p(conv[s0:]), conv.stride(0), conv.stride(1), conv.stride(2),
                                                         conv.stride(3), _p(bias), _p(pred[s0:]), C, Hs, Ws,
                                                         ctypes.cast(arr, ctypes.c_void_p), s1 - s0,
                                                         _p(out[s0:]) if out is not None else None, Ho, Wo, _stream()),
                   "rmem_finish_logits_batched")
        global LAUNCHES
        LAUNCHES += 2 if out is not None else 1


def logits_head(x, pre_bias, gamma, beta, w_out, b_out, n_valid, pred, out, label, workspace, groups=8):
    """x: f32 or fp16 channels-last [S, C, Hs, Ws] raw conv output; w_out f32 [Co, C]; pred f32 [S, Co, Hs, Ws]; out f32
    [S, Co, Ho, Wo] or None; label f32 [S, 1, Ho, Wo] or None (with out).  GroupNorm + ReLU + 1x1 conv + bias + class
    mask (+ resize + argmax): the decoder's last stage in three launches."""
    S, C, Hs, Ws = x.shape
    Co = w_out.shape[0]
    Ho, Wo = (out.shape[-2], out.shape[-1]) if out is not None else (0, 0)
    nv = [int(n_valid)] * S if not isinstance(n_valid, (list, tuple)) else [int(v) for v in n_valid]
    arr = (ctypes.c_int32 * S)(*nv)
    xt = x.permute(0, 2, 3, 1)
    if not xt.is_contiguous():
        raise _lib.RmemError("logits_head: x must be channels-last")
    if x.dtype not in (torch.float32, torch.float16):
        raise _lib.RmemError("logits_head: x must be fp32 or fp16")
    _lib.check(_lib.lib().rmem_logits_head_batched(_p(x), 1 if x.dtype == torch.float16 else 0, _p(pre_bias), _p(gamma),
                                                   _p(beta), Hs, Ws, C, groups, _p(w_out),
                                                   _p(b_out), Co, ctypes.cast(arr, ctypes.c_void_p), S, _p(pred),
                                                   _p(out), Ho, Wo, _p(label), _p(workspace), _stream()),
               "rmem_logits_head_batched")
    global LAUNCHES
    LAUNCHES += 3 if out is not None else 2


def pack_image(img, out):
    """img f32 [1, 3, H, W] contiguous -> out bf16 [1, 8, H, W] channels-last (channels 3..7 zero)."""
    _chk_cuda(img, out)
    if not img.is_contiguous():
        img = img.contiguous()
    H, W = img.shape[-2:]
    _lib.check(_lib.lib().rmem_pack_image(_p(img), H, W, _p(out), _stream()), "rmem_pack_image")
    _count("rmem_pack_image")


def pack_image_s2d(img, out, mean=None, std=None):
    """img f32 [1, 3, H, W] contiguous -- or u8 [1, H, W, 3] / [H, W, 3] interleaved pixels, normalised here as
    (u8 / 255 - mean) / std -- -> out bf16 or fp16 [1, 16, (H+1)//2+3, (W+1)//2+3] channels-last: the 2x2
    space-to-depth image with the stem convolution's zero padding built in (see include/rmem_b200.h)."""
    _chk_cuda(img, out)
    if not img.is_contiguous():
        img = img.contiguous()
    u8 = img.dtype == torch.uint8
    S = out.shape[0]
    if u8:
        if img.shape[-1] != 3:
            raise _lib.RmemError("pack_image_s2d: u8 frames are interleaved [H, W, 3]")
        H, W = img.shape[-3:-1]
        mean = (0.0, 0.0, 0.0) if mean is None else mean
        std = (1.0, 1.0, 1.0) if std is None else std
        m3 = (ctypes.c_float * 3)(*[float(v) for v in mean])
        s3 = (ctypes.c_float * 3)(*[1.0 / float(v) for v in std])
    else:
        if img.dtype != torch.float32:
            raise _lib.RmemError("pack_image_s2d: img must be f32 planes or u8 pixels")
        H, W = img.shape[-2:]
        m3 = s3 = None
    if out.dtype not in (torch.bfloat16, torch.float16):
        raise _lib.RmemError("pack_image_s2d: out must be bf16 or fp16")
    if img.numel() != S * 3 * H * W:
        raise _lib.RmemError("pack_image_s2d: image / output batch mismatch")
    _lib.check(_lib.lib().rmem_pack_image_s2d_batched(_p(img), int(u8), m3, s3, H, W, _p(out),
                                                      int(out.dtype == torch.float16), S, _stream()),
               "rmem_pack_image_s2d_batched")
    _count("rmem_pack_image_s2d")


def maxpool3s2_nhwc(x, out, bias=None, relu=False):
    """x bf16 [1, C, H, W] channels-last -> out bf16 [1, C, (H-1)//2+1, (W-1)//2+1] channels-last:
    MaxPool2d(3, 2, 1) of the encoder stem; with bias (f32 [C]) / relu: relu(maxpool(x) + bias)."""
    _chk_cuda(x, out)
    n, C, H, W = x.shape
    if (x.dtype not in (torch.bfloat16, torch.float16) or x.stride()[1:] != (1, W * C, C)
            or (n > 1 and x.stride(0) != H * W * C)):
        raise ValueError("maxpool3s2_nhwc: expects bf16 / fp16 channels-last images")
    if out.dtype != x.dtype:
        raise ValueError("maxpool3s2_nhwc: out must have the input's dtype")
    _lib.check(_lib.lib().rmem_maxpool3s2_nhwc_batched(_p(x), H, W, C, _p(bias), 1 if relu else 0, _p(out),
                                                       int(x.dtype == torch.float16), n, _stream()),
               "rmem_maxpool3s2_nhwc_batched")
    _count("rmem_maxpool3s2_nhwc")


def argmax_label(logits, label=None):
    """logits f32 [1, C, H, W] contiguous -> label f32 [1, 1, H, W] (argmax over the object channel)."""
    _chk_cuda(logits)
    n, C, H, W = logits.shape
    if label is None:
        label = torch.empty(n, 1, H, W, device=logits.device, dtype=torch.float32)
    _lib.check(_lib.lib().rmem_argmax_label_batched(_p(logits), C, H * W, _p(label), n, _stream()),
               "rmem_argmax_label_batched")
    _count("rmem_argmax_label")
    return label


def tpe_prepare(q, cur_pos_emb, mem_pos_emb, q_time, tpe_bias):
    L, C = q.shape
    _lib.check(_lib.lib().rmem_tpe_prepare(_p(q), _p(cur_pos_emb), _p(mem_pos_emb), _p(q_time), _p(tpe_bias),
                                           L, C, _stream()), "rmem_tpe_prepare")
    _count("rmem_tpe_prepare")


# ---------------------------------------------------------------------------------------------
def lt_attn_partial(q, kbank, vbank, order, T, cpf, part_o, part_m, part_l):
    L = q.shape[0]
    _lib.check(_lib.lib().rmem_lt_attn_deaot_partial(_p(q), _p(kbank), _p(vbank), _p(order), T, L,
                                                     kbank.shape[0], cpf, _p(part_o), _p(part_m), _p(part_l),
                                                     _stream()), "rmem_lt_attn_deaot_partial")
    _count("rmem_lt_attn_deaot_partial")


def lt_attn_combine(part_o, part_m, part_l, tpe_bias, tpe_rows, T, L, cpf, gate, out, attn_mass):
    rows = (ctypes.c_int32 * max(T, 1))(*(tpe_rows if tpe_rows is not None else [0] * T))
    _lib.check(_lib.lib().rmem_lt_attn_deaot_combine(_p(part_o), _p(part_m), _p(part_l), _p(tpe_bias),
                                                     ctypes.cast(rows, ctypes.c_void_p), T, L, cpf, _p(gate),
                                                     _p(out), _p(attn_mass), _stream()),
               "rmem_lt_attn_deaot_combine")
    _count("rmem_lt_attn_deaot_combine")


def lt_attn_single(q, k, v, gate, out):
    """One frame of keys, one launch: out = gate * softmax(q k^T / sqrt(128)) v.  q, k [L,128]; v, gate, out [L,1024]."""
    L = q.shape[0]
    _lib.check(_lib.lib().rmem_lt_attn_deaot_single(_p(q), _p(k), _p(v), L, _p(gate), _p(out), _stream()),
               "rmem_lt_attn_deaot_single")
    _count("rmem_lt_attn_deaot_single")


def lt_attn_sk_workspace(L, T_max, device, n_clusters=0, S=1):
    """Workspace of the stream-K long-term attention for up to T_max frames in the bank (uint8 tensor)."""
    n = max(_lib.lib().rmem_lt_attn_deaot_sk_workspace_bytes(L, T, S, n_clusters) for T in range(1, T_max + 1))
    return torch.empty(n, device=device, dtype=torch.uint8)


def lt_attn_sk_partial(q, kbank, vbank, order, T, tpe_bias, tpe_rows, record, workspace, n_clusters=0, S=1,
                       order_stride=0):
    """S = 1: q [L,128], kbank [n_slots,L,128], vbank [n_slots,L,1024].  S > 1: the same with a leading [S] dimension
    (q may be given as [S*L,128]); `order` of sequence s starts order_stride int32 words after that of s - 1."""
    L = q.shape[-2] if q.dim() == 3 else q.shape[0] // S
    n_slots = kbank.shape[-3]
    rows = (ctypes.c_int32 * max(T, 1))(*(tpe_rows if tpe_rows is not None else [0] * T))
    _lib.check(_lib.lib().rmem_lt_attn_deaot_sk_partial(_p(q), _p(kbank), _p(vbank), _p(order), order_stride, T, L,
                                                        n_slots, S, _p(tpe_bias), ctypes.cast(rows, ctypes.c_void_p),
                                                        1 if record else 0, n_clusters, _p(workspace),
                                                        workspace.numel(), _stream()),
               "rmem_lt_attn_deaot_sk_partial")
    _count("rmem_lt_attn_deaot_sk_partial")


def lt_attn_sk_combine(T, L, gate, out, attn_mass, workspace, n_clusters=0, S=1):
    _lib.check(_lib.lib().rmem_lt_attn_deaot_sk_combine(T, L, S, _p(gate), _p(out), _p(attn_mass), n_clusters,
                                                        _p(workspace), workspace.numel(), _stream()),
               "rmem_lt_attn_deaot_sk_combine")
    _count("rmem_lt_attn_deaot_sk_combine")


def local_banded_buffer(h, w, device, S=1):
    """Zero-filled bf16 buffer for the banded probabilities of S (h, w) token grids (zero-fill once: the
    scores kernel only ever writes the in-window slots)."""
    n = _lib.lib().rmem_local_banded_elems(h, w)
    return torch.zeros(S * n, device=device, dtype=torch.bfloat16)


def local_attn_scores(q, k, rel_bias, probs, banded, h, w, S=1):
    _lib.check(_lib.lib().rmem_local_attn_scores_batched(_p(q), _p(k), _p(rel_bias), rel_bias.stride(0), _p(probs),
                                                         _p(banded), h, w, S, _stream()),
               "rmem_local_attn_scores_batched")
    _count("rmem_local_attn_scores")


def local_attn(q, k, v, rel_bias, gate, out, h, w):
    """Fused 15x15 local gated attention: q, k bf16 [L,128]; v, gate, out bf16 [L,1024]; rel_bias f32 [L, ld]."""
    _chk_cuda(q, k, v, rel_bias, out)
    _lib.check(_lib.lib().rmem_local_attn(_p(q), _p(k), _p(v), _p(rel_bias), rel_bias.stride(0), _p(gate), _p(out),
                                          h, w, _stream()), "rmem_local_attn")
    _count("rmem_local_attn")


def local_attn_aggregate(banded, v, gate, out, h, w, S=1):
    _lib.check(_lib.lib().rmem_local_attn_aggregate_batched(_p(banded), _p(v), _p(gate), _p(out), h, w, S, _stream()),
               "rmem_local_attn_aggregate_batched")
    _count("rmem_local_attn_aggregate")


# ---------------------------------------------------------------------------------------------
def bank_init(state, frame_step):
    _lib.check(_lib.lib().rmem_bank_init(_p(state), frame_step, _stream()), "rmem_bank_init")
    _count("rmem_bank_init")


def bank_tick(state):
    """state: int32 [BANK_STATE_WORDS] or a stack [S, BANK_STATE_WORDS] (one launch)."""
    S = state.shape[0] if state.dim() == 2 else 1
    _lib.check(_lib.lib().rmem_bank_tick_batched(_p(state), S, _stream()), "rmem_bank_tick_batched")
    _count("rmem_bank_tick")


def id_prefix_table(weight):
    """patch_wise_id_bank.weight [256, n_cls, 17, 17] -> the lookup table rmem_id_embed reads (host f32, flat):
    row prefixes [n_cls][17][18][256] followed by 2-D prefixes [n_cls][18][18][256] (include/rmem_b200.h)."""
    wk = weight.detach().double().cpu().permute(1, 2, 3, 0)                 # [n_cls, ky, kx, c]
    n_cls = wk.shape[0]
    p1 = torch.zeros(n_cls, 17, 18, 256, dtype=torch.float64)
    p1[:, :, 1:, :] = torch.cumsum(wk, dim=2)
    p2 = torch.zeros(n_cls, 18, 18, 256, dtype=torch.float64)
    p2[:, 1:, :, :] = torch.cumsum(p1, dim=1)
    return torch.cat([p1.reshape(-1), p2.reshape(-1)]).float().contiguous()


def id_embed(label, prefix, bias, gamma, beta, n_cls, dsts, h, w):
    """label: f32 [H, W] or a stack [S, H, W]; dsts: list of (bf16 tensor [S*h*w, ld], col0)."""
    H, W = label.shape[-2:]
    S = label.shape[0] if label.dim() == 3 else 1
    n = len(dsts)
    dp = (ctypes.c_void_p * n)(*[d.data_ptr() for d, _ in dsts])
    ld = (ctypes.c_int32 * n)(*[d.stride(0) for d, _ in dsts])
    c0 = (ctypes.c_int32 * n)(*[c for _, c in dsts])
    _lib.check(_lib.lib().rmem_id_embed_batched(_p(label), H, W, _p(prefix), _p(bias), _p(gamma), _p(beta), n_cls,
                                                ctypes.cast(dp, ctypes.c_void_p), ctypes.cast(ld, ctypes.c_void_p),
                                                ctypes.cast(c0, ctypes.c_void_p), n, h, w, S, _stream()),
               "rmem_id_embed_batched")
    _count("rmem_id_embed")


def id_embed_soft(prob, ignore, taps, bias, gamma, beta, n_cls, dsts, h, w):
    """prob: f32 [S, n_ch, H, W] one-hot or probability masks; ignore: f32 [S, H, W] or None; taps: id_tap_table()."""
    S, n_ch, H, W = prob.shape
    n = len(dsts)
    dp = (ctypes.c_void_p * n)(*[d.data_ptr() for d, _ in dsts])
    ld = (ctypes.c_int32 * n)(*[d.stride(0) for d, _ in dsts])
    c0 = (ctypes.c_int32 * n)(*[c for _, c in dsts])
    _lib.check(_lib.lib().rmem_id_embed_soft(_p(prob), n_ch, _p(ignore), H, W, _p(taps), _p(bias), _p(gamma), _p(beta),
                                             n_cls, ctypes.cast(dp, ctypes.c_void_p), ctypes.cast(ld, ctypes.c_void_p),
                                             ctypes.cast(c0, ctypes.c_void_p), n, h, w, S, _stream()),
               "rmem_id_embed_soft")
    _count("rmem_id_embed_soft")


def id_tap_table(weight):
    """conv.weight [256, n_cls, 17, 17] -> [n_cls, 17, 17, 256] fp32 (output channel innermost)."""
    return weight.detach().float().permute(1, 2, 3, 0).contiguous()


def bank_append(state, T, frame_step, srcs, dsts):
    """Copy each src tensor (one slot's worth) into slot order[T] of the matching bank tensor."""
    n = len(srcs)
    sp = (ctypes.c_void_p * n)(*[s.data_ptr() for s in srcs])
    dp = (ctypes.c_void_p * n)(*[d.data_ptr() for d in dsts])
    rb = (ctypes.c_int64 * n)(*[s.numel() * s.element_size() for s in srcs])
    _lib.check(_lib.lib().rmem_bank_append(_p(state), T, frame_step, ctypes.cast(sp, ctypes.c_void_p),
                                           ctypes.cast(dp, ctypes.c_void_p), ctypes.cast(rb, ctypes.c_void_p),
                                           n, _stream()), "rmem_bank_append")
    _count("rmem_bank_append")


def bank_score_evict(state, attn_mass, fg, L, T_old, capacity):
    _lib.check(_lib.lib().rmem_bank_score_evict(_p(state), _p(attn_mass), _p(fg), L, T_old, capacity,
                                                _stream()), "rmem_bank_score_evict")
    _count("rmem_bank_score_evict")


def foreground_prob(logits, fg, h, w):
    C, Hs, Ws = logits.shape[-3:]
    _lib.check(_lib.lib().rmem_foreground_prob(_p(logits), C, Hs, Ws, _p(fg), h, w, _stream()),
               "rmem_foreground_prob")
    _count("rmem_foreground_prob")


# --------------------------------------------------------------------------------------------- AOT (8 heads x 32)
def mha_partial(q, k, v, order, T, cpf, part_o, part_m, part_l):
    """q bf16 [L, >=256] (row stride = q.stride(0)); k, v bf16 [n_slots, L, >=256] or [L, >=256] (one slot)."""
    L = q.shape[0]
    if k.dim() == 2:
        n_slots, ldk, ks, ldv, vs = 1, k.stride(0), 0, v.stride(0), 0
    else:
        n_slots, ldk, ks, ldv, vs = k.shape[0], k.stride(1), k.stride(0), v.stride(1), v.stride(0)
    _lib.check(_lib.lib().rmem_mha_partial(_p(q), q.stride(0), _p(k), ldk, ks, _p(v), ldv, vs, _p(order), T, L,
                                           n_slots, cpf, _p(part_o), _p(part_m), _p(part_l), _stream()),
               "rmem_mha_partial")
    _count("rmem_mha_partial")


def mha_combine(part_o, part_m, part_l, tpe_bias, tpe_rows, T, L, cpf, out, out_col0, attn_mass):
    rows = (ctypes.c_int32 * max(T, 1))(*(tpe_rows if tpe_rows is not None else [0] * T))
    _lib.check(_lib.lib().rmem_mha_combine(_p(part_o), _p(part_m), _p(part_l), _p(tpe_bias),
                                           ctypes.cast(rows, ctypes.c_void_p), T, L, cpf, _p(out), out.stride(0),
                                           out_col0, _p(attn_mass), _stream()), "rmem_mha_combine")
    _count("rmem_mha_combine")


def tpe_prepare_mha(q, cur_pos_emb, mem_pos_emb, q_time, tpe_bias):
    _lib.check(_lib.lib().rmem_tpe_prepare_mha(_p(q), q.stride(0), _p(cur_pos_emb), _p(mem_pos_emb), _p(q_time),
                                               _p(tpe_bias), q.shape[0], _stream()), "rmem_tpe_prepare_mha")
    _count("rmem_tpe_prepare_mha")


def layernorm_ex(xa, gamma, beta, xb=None, out=None, col_out=0, post_add=None, out_plus=None, col_plus=0,
                 out_f32=None, col_f32=0):
    """LN(xa [+ xb]) -> out (bf16) / out_plus = LN + post_add (bf16) / out_f32; all [L, 256] column windows."""
    _chk_cuda(xa)
    L = xa.shape[0]
    in_bf16 = 1 if xa.dtype == torch.bfloat16 else 0
    st = lambda t: t.stride(0) if t is not None else 0
    _lib.check(_lib.lib().rmem_layernorm_ex(_p(xa), xa.stride(0), _p(xb), st(xb), in_bf16, _p(gamma), _p(beta),
                                            _p(out), st(out), col_out, _p(post_add), st(post_add),
                                            _p(out_plus), st(out_plus), col_plus, _p(out_f32), st(out_f32),
                                            col_f32, L, 256, _stream()), "rmem_layernorm_ex")
    _count("rmem_layernorm_ex")


def groupnorm_gelu_tokens(x, gamma, beta, out, groups, workspace):
    L, C = x.shape
    _lib.check(_lib.lib().rmem_groupnorm_gelu_tokens(_p(x), _p(gamma), _p(beta), _p(out), L, C, groups,
                                                     _p(workspace), _stream()), "rmem_groupnorm_gelu_tokens")
    _count("rmem_groupnorm_gelu_tokens")
