"""Model containers for the RMem hot path: parameters under the reference's state_dict key names
 so a reference checkpoint / state_dict loads unchanged.

The LSTT blocks here hold parameters only -- their arithmetic runs in the CUDA kernels driven by
``engine.py``.  The ResNet-50 encoder (stride 16, no layer4; reference
networks/encoders/resnet.py:71-194), the FPN head (networks/decoders/fpn.py:36-68) and the 17x17
ID-bank convolution (networks/models/aot.py:64-83) are the black boxes of this tier: they stay
library (cuDNN) calls.
"""
import zlib
from dataclasses import dataclass

import numpy as np
import torch
import torch.nn as nn
import torch.nn.functional as F


@dataclass
class RMemConfig:
    """The subset of configs/default.py + configs/models/r50_{aotl,deaotl}.py the path reads."""
    model_vos: str = "deaot"            # 'deaot' | 'aot'
    embed_dim: int = 256                # MODEL_ENCODER_EMBEDDING_DIM
    lstt_num: int = 3                   # MODEL_LSTT_NUM
    max_obj_num: int = 10               # MODEL_MAX_OBJ_NUM
    ignore_token: bool = True           # MODEL_IGNORE_TOKEN (configs/pre_vost.py)
    former_mem_len: int = 1             # FORMER_MEM_LEN
    latter_mem_len: int = 4             # LATTER_MEM_LEN
    tpe_rows: int = 4                   # TEMPORAL_POSITIONAL_EMBEDDING_SLOT_4
    test_long_term_mem_gap: int = 5     # TEST_LONG_TERM_MEM_GAP
    encoder_dims: tuple = (256, 512, 1024, 1024)

    # attribute names the reference's cfg object exposes (read by callers of the engine)
    @property
    def MODEL_ALIGN_CORNERS(self):
        return True

    @property
    def FORMER_MEM_LEN(self):
        return self.former_mem_len

    @property
    def LATTER_MEM_LEN(self):
        return self.latter_mem_len

    @property
    def MODEL_MAX_OBJ_NUM(self):
        return self.max_obj_num


# ------------------------------------------------------------------------------------------------
class FrozenBN(nn.Module):
    def __init__(self, n):
        super().__init__()
        self.register_buffer("weight", torch.ones(n))
        self.register_buffer("bias", torch.zeros(n))
        self.register_buffer("running_mean", torch.zeros(n))
        self.register_buffer("running_var", torch.ones(n) - 1e-5)

    def scale_shift(self):
        scale = self.weight * (self.running_var + 1e-5).rsqrt()
        return scale, self.bias - self.running_mean * scale

    def forward(self, x):
        s, b = self.scale_shift()
        return x * s.view(1, -1, 1, 1).to(x.dtype) + b.view(1, -1, 1, 1).to(x.dtype)


class _Bottleneck(nn.Module):
    def __init__(self, inplanes, planes, stride, downsample):
        super().__init__()
        self.conv1 = nn.Conv2d(inplanes, planes, 1, bias=False)
        self.bn1 = FrozenBN(planes)
        self.conv2 = nn.Conv2d(planes, planes, 3, stride, 1, bias=False)
        self.bn2 = FrozenBN(planes)
        self.conv3 = nn.Conv2d(planes, planes * 4, 1, bias=False)
        self.bn3 = FrozenBN(planes * 4)
        self.downsample = downsample

    def forward(self, x):
        out = F.relu(self.bn1(self.conv1(x)))
        out = F.relu(self.bn2(self.conv2(out)))
        out = self.bn3(self.conv3(out))
        if self.downsample is not None:
            x = self.downsample(x)
        return F.relu(out + x)


class ResNet50S16(nn.Module):
    """ResNet-50 trunk to stride 16 (layer1..layer3), returns [x4, x8, x16, x16]."""

    def __init__(self):
        super().__init__()
        self.conv1 = nn.Conv2d(3, 64, 7, 2, 3, bias=False)
        self.bn1 = FrozenBN(64)
        inplanes = 64
        for li, (planes, blocks, stride) in enumerate([(64, 3, 1), (128, 4, 2), (256, 6, 2)], start=1):
            layers = []
            for b in range(blocks):
                down = None
                if b == 0:
                    down = nn.Sequential(nn.Conv2d(inplanes, planes * 4, 1, stride, bias=False),
                                         FrozenBN(planes * 4))
                layers.append(_Bottleneck(inplanes, planes, stride if b == 0 else 1, down))
                inplanes = planes * 4
            setattr(self, "layer%d" % li, nn.Sequential(*layers))

    def forward(self, x):
        x = F.relu(self.bn1(self.conv1(x)))
        x = F.max_pool2d(x, 3, 2, 1)
        xs = []
        for li in (1, 2, 3):
            x = getattr(self, "layer%d" % li)(x)
            xs.append(x)
        xs.append(x)
        return xs

    @torch.no_grad()
    def fold_bn_(self):
        """Inference-time folding of the frozen BN affine into the preceding conv (exact algebra)."""
        def fold(conv, bn):
            s, b = bn.scale_shift()
            conv.weight.mul_(s.view(-1, 1, 1, 1).to(conv.weight.dtype))
            conv.bias = nn.Parameter(b.to(conv.weight.dtype), requires_grad=False)
            bn.weight.fill_(1.0); bn.bias.zero_(); bn.running_mean.zero_(); bn.running_var.fill_(1.0 - 1e-5)
        fold(self.conv1, self.bn1)
        for li in (1, 2, 3):
            for blk in getattr(self, "layer%d" % li):
                fold(blk.conv1, blk.bn1); fold(blk.conv2, blk.bn2); fold(blk.conv3, blk.bn3)
                if blk.downsample is not None:
                    fold(blk.downsample[0], blk.downsample[1])
        self._folded = True


class _ConvGN(nn.Module):
    def __init__(self, cin, cout, k):
        super().__init__()
        self.conv = nn.Conv2d(cin, cout, k, padding=k // 2)
        self.gn = nn.GroupNorm(8, cout)

    def forward(self, x):
        return self.gn(self.conv(x))


class FPNHead(nn.Module):
    def __init__(self, in_dim, out_dim, intermediate, hidden=256, shortcut_dims=(256, 512, 1024, 1024)):
        super().__init__()
        self.intermediate = intermediate
        self.conv_in = _ConvGN(in_dim, hidden, 1)
        self.conv_16x = _ConvGN(hidden, hidden, 3)
        self.conv_8x = _ConvGN(hidden, hidden // 2, 3)
        self.conv_4x = _ConvGN(hidden // 2, hidden // 2, 3)
        self.adapter_16x = nn.Conv2d(shortcut_dims[-2], hidden, 1)
        self.adapter_8x = nn.Conv2d(shortcut_dims[-3], hidden, 1)
        self.adapter_4x = nn.Conv2d(shortcut_dims[-4], hidden // 2, 1)
        self.conv_out = nn.Conv2d(hidden // 2, out_dim, 1)

    def forward(self, inputs, shortcuts):
        x = torch.cat(inputs, dim=1) if self.intermediate else inputs[-1]
        x = F.relu(self.conv_in(x))
        x = F.relu(self.conv_16x(self.adapter_16x(shortcuts[-2]) + x))
        x = F.interpolate(x, size=shortcuts[-3].shape[-2:], mode="bilinear", align_corners=True)
        x = F.relu(self.conv_8x(self.adapter_8x(shortcuts[-3]) + x))
        x = F.interpolate(x, size=shortcuts[-4].shape[-2:], mode="bilinear", align_corners=True)
        x = F.relu(self.conv_4x(self.adapter_4x(shortcuts[-4]) + x))
        return self.conv_out(x)


# ------------------------------------------------------------------------------------------------
# LSTT parameter containers (names = reference module attribute names; no forward here)
class _GatedPropagationParams(nn.Module):
    """Parameters of GatedPropagation (networks/layers/attention.py:93-137)."""

    def __init__(self, d_qk, d_vu, d_att, use_linear):
        super().__init__()
        expand = d_vu * 2
        if use_linear:
            self.linear_QK = nn.Linear(d_qk, d_att)
            self.linear_V1 = nn.Linear(d_vu // 2, expand // 2)
            self.linear_V2 = nn.Linear(d_vu // 2, expand // 2)
            self.linear_U1 = nn.Linear(d_vu // 2, expand // 2)
            self.linear_U2 = nn.Linear(d_vu // 2, expand // 2)
        self.dw_conv = nn.Module()
        self.dw_conv.conv = nn.Conv2d(expand, expand, 5, padding=2, groups=expand, bias=False)
        self.projection = nn.Linear(expand, d_vu)


class _LocalGatedPropagationParams(nn.Module):
    """Parameters of LocalGatedPropagation (attention.py:220-287), use_linear=False, 1 head."""

    def __init__(self, d_vu, d_att, max_dis=7):
        super().__init__()
        expand = d_vu * 2
        win = 2 * max_dis + 1
        self.relative_emb_k = nn.Conv2d(d_att, win * win, 1)
        self.dw_conv = nn.Module()
        self.dw_conv.conv = nn.Conv2d(expand, expand, 5, padding=2, groups=expand, bias=False)
        self.projection = nn.Linear(expand, d_vu)


class GPMLayerParams(nn.Module):
    """GatedPropagationModule (networks/layers/transformer.py:1010-1082)."""

    def __init__(self, d_model, layer_idx):
        super().__init__()
        d_att, expand = d_model // 2, d_model * 2
        self.norm1 = nn.LayerNorm(d_model)
        self.linear_QV = nn.Linear(d_model, d_att + expand)
        self.linear_U = nn.Linear(d_model, expand)
        if layer_idx == 0:
            self.linear_ID_V = nn.Linear(d_model, expand)
        else:
            self.id_norm1 = nn.LayerNorm(d_model)
            self.linear_ID_V = nn.Linear(d_model * 2, expand)
            self.linear_ID_U = nn.Linear(d_model, expand)
        self.long_term_attn = _GatedPropagationParams(d_model, d_model * 2, d_att, use_linear=False)
        self.short_term_attn = _LocalGatedPropagationParams(d_model * 2, d_att)
        self.norm2 = nn.LayerNorm(d_model)
        self.id_norm2 = nn.LayerNorm(d_model)
        self.self_attn = _GatedPropagationParams(d_model * 2, d_model * 2, d_att, use_linear=True)


class _GroupNorm1DParams(nn.Module):
    def __init__(self, dim, groups):
        super().__init__()
        self.gn = nn.GroupNorm(groups, dim)


class DualBranchGPMParams(nn.Module):
    """DualBranchGPM (transformer.py:700-763): 3 layers + one GroupNorm1D(512, 2)."""

    def __init__(self, num_layers, d_model):
        super().__init__()
        self.layers = nn.ModuleList([GPMLayerParams(d_model, i) for i in range(num_layers)])
        self.decoder_norms = nn.ModuleList([_GroupNorm1DParams(d_model * 2, 2)])


class DeAOTModel(nn.Module):
    """Parameter-compatible with reference networks/models/deaot.py:DeAOT (R50-DeAOT-L)."""

    def __init__(self, cfg: RMemConfig = None):
        super().__init__()
        self.cfg = cfg or RMemConfig()
        c = self.cfg
        self.max_obj_num = c.max_obj_num
        self.encoder = ResNet50S16()
        self.encoder_projector = nn.Conv2d(c.encoder_dims[-1], c.embed_dim, 1)
        self.LSTT = DualBranchGPMParams(c.lstt_num, c.embed_dim)
        self.decoder = FPNHead(c.embed_dim * 2, c.max_obj_num + 1, intermediate=False, hidden=c.embed_dim,
                               shortcut_dims=c.encoder_dims)
        id_dim = c.max_obj_num + (2 if c.ignore_token else 1)
        self.patch_wise_id_bank = nn.Conv2d(id_dim, c.embed_dim, 17, 16, 8)
        self.id_norm = nn.LayerNorm(c.embed_dim)
        self.cur_pos_emb = nn.Parameter(torch.zeros(1, c.embed_dim // 2))
        self.mem_pos_emb = nn.Parameter(torch.zeros(c.tpe_rows, c.embed_dim // 2))
        self.use_temporal_pe = True
        self.eval()
        for p in self.parameters():
            p.requires_grad_(False)


class _MHAParams(nn.Module):
    """Parameters of MultiheadAttention (networks/layers/attention.py:8-26)."""

    def __init__(self, d_model, use_linear):
        super().__init__()
        if use_linear:
            self.linear_Q = nn.Linear(d_model, d_model)
            self.linear_K = nn.Linear(d_model, d_model)
            self.linear_V = nn.Linear(d_model, d_model)
        self.projection = nn.Linear(d_model, d_model)


class _GNActDWConvParams(nn.Module):
    """GNActDWConv2d (networks/layers/basic.py:15-35): GroupNorm(32) -> GELU -> 5x5 depth-wise."""

    def __init__(self, dim):
        super().__init__()
        self.gn = nn.GroupNorm(32, dim)
        self.conv = nn.Conv2d(dim, dim, 5, padding=2, groups=dim, bias=False)


class AOTLayerParams(nn.Module):
    """SimplifiedTransformerBlock (networks/layers/transformer.py:466-551), MODEL_LINEAR_Q = False."""

    def __init__(self, d_model, dim_feedforward=1024):
        super().__init__()
        self.norm1 = nn.LayerNorm(d_model)
        self.self_attn = _MHAParams(d_model, True)
        self.norm2 = nn.LayerNorm(d_model)
        self.linear_Q = nn.Linear(d_model, d_model)
        self.linear_V = nn.Linear(d_model, d_model)
        self.linear_QMem = nn.Linear(d_model, d_model)
        self.linear_VMem = nn.Linear(d_model, d_model)
        self.norm4 = nn.LayerNorm(d_model)
        self.linear_KMem = nn.Linear(d_model, d_model)        # present in checkpoints, unused in forward
        self.long_term_attn = _MHAParams(d_model, False)
        self.short_term_attn = _MHAParams(d_model, False)
        self.norm3 = nn.LayerNorm(d_model)
        self.linear1 = nn.Linear(d_model, dim_feedforward)
        self.activation = _GNActDWConvParams(dim_feedforward)
        self.linear2 = nn.Linear(dim_feedforward, d_model)


class LSTTParams(nn.Module):
    """LongShortTermTransformer (transformer.py:133-197): 3 blocks + 3 output LayerNorms."""

    def __init__(self, num_layers, d_model):
        super().__init__()
        self.layers = nn.ModuleList([AOTLayerParams(d_model) for _ in range(num_layers)])
        self.decoder_norms = nn.ModuleList([nn.LayerNorm(d_model) for _ in range(num_layers)])


class AOTModel(nn.Module):
    """Parameter-compatible with reference networks/models/aot.py:AOT (R50-AOT-L)."""

    def __init__(self, cfg: RMemConfig = None):
        super().__init__()
        self.cfg = cfg or RMemConfig(model_vos="aot")
        c = self.cfg
        self.max_obj_num = c.max_obj_num
        self.encoder = ResNet50S16()
        self.encoder_projector = nn.Conv2d(c.encoder_dims[-1], c.embed_dim, 1)
        self.LSTT = LSTTParams(c.lstt_num, c.embed_dim)
        self.decoder = FPNHead(c.embed_dim * (c.lstt_num + 1), c.max_obj_num + 1, intermediate=True,
                               hidden=c.embed_dim, shortcut_dims=c.encoder_dims)
        id_dim = c.max_obj_num + (2 if c.ignore_token else 1)
        self.patch_wise_id_bank = nn.Conv2d(id_dim, c.embed_dim, 17, 16, 8)
        self.cur_pos_emb = nn.Parameter(torch.zeros(1, c.embed_dim))
        self.mem_pos_emb = nn.Parameter(torch.zeros(c.tpe_rows, c.embed_dim))
        self.use_temporal_pe = True
        self.eval()
        for p in self.parameters():
            p.requires_grad_(False)


def deterministic_init_(state_dict, seed=0, attn_sharpness=2.5):
    """Fill every tensor of a state_dict in place with values that depend only on (seed, key name,
    shape): numpy's legacy RandomState is stable across platforms, so the build container (reference
    model) and the GPU box (this model) get bit-identical weights without shipping a checkpoint."""
    for name in sorted(state_dict.keys()):
        t = state_dict[name]
        rs = np.random.RandomState((seed * 1000003 + zlib.crc32(name.encode())) % (2 ** 31))
        v = rs.standard_normal(size=tuple(t.shape)).astype(np.float32)
        leaf = name.rsplit(".", 1)[-1]
        if t.dim() > 1:                                   # conv / linear weights: xavier-like scale
            fan_out = t.shape[0] * int(np.prod(t.shape[2:]))
            fan_in = t.shape[1] * int(np.prod(t.shape[2:]))
            if "encoder." in name:
                std = np.sqrt(2.0 / fan_in) * 0.7         # keeps the residual trunk's activations O(1)
            elif "patch_wise_id_bank" in name:
                std = 0.02
            elif "dw_conv" in name or "activation.conv" in name:
                std = 0.2                                 # depth-wise 5x5: 25 taps -> O(1) gain
            elif "pos_emb" in name:
                std = 0.05
            else:
                std = np.sqrt(2.0 / (fan_in + fan_out))
            v *= std
            # sharper attention than xavier gives (q = k scale), so relevance is not uniform
            if name.endswith("linear_QV.weight"):
                v[: t.shape[1] // 2] *= attn_sharpness
            elif name.endswith("linear_QK.weight") or name.endswith(".linear_Q.weight"):
                v *= attn_sharpness
        elif leaf == "running_var":
            v = 1.0 + 0.1 * np.abs(v)
        elif leaf == "running_mean":
            v *= 0.1
        elif leaf == "weight":                            # norm scales
            v = 1.0 + 0.1 * v
        elif leaf == "bias":
            v *= 0.02
        else:
            v *= 0.05
        with torch.no_grad():
            t.copy_(torch.from_numpy(v).to(t.dtype))
    return state_dict
