"""Drop-in engines for the RMem per-frame path: AOTEngine / DeAOTEngine / *InferEngine with the
reference's method names and argument lists (reference networks/engines/aot_engine.py:18-725,
deaot_engine.py:9-56), driving the sm_100a kernels of librmem_b200.so through ``ops``.

State layout (per engine = per video sequence, one object group of <= 10 objects):
  * long-term bank : per layer K [n_slots, L, 128] and V|ID_V [n_slots, L, 1024] in bf16, fixed
    capacity (former + latter + 1 slots), plus a device-resident slot table (``bank_state``) --
    appends are one strided copy, evictions an index shuffle done by the device kernel;
  * short-term memory: previous frame's K and V|ID_V per layer;
  * residual streams in fp32, everything the tensor cores read in bf16.
The frame path issues no host synchronisation; ``long_memories_indexes`` is mirrored lazily.
There is no CPU / PyTorch fallback: without CUDA or the shared library the engines raise.
"""

import functools

import collections
import os

import torch
import torch.nn.functional as F

from . import _lib, ops
from .tpe import tpe_row_indices

BF16 = torch.bfloat16


def _bf(t, dev):
    return t.detach().to(device=dev, dtype=BF16).contiguous()


def _f32(t, dev):
    return t.detach().to(device=dev, dtype=torch.float32).contiguous()


def _on_device(fn):
    """Run a public engine method with the engine's GPU current: the kernels launch on
    ``torch.cuda.current_stream()`` and set per-device kernel attributes, so an engine built for ``gpu_id`` must
    not depend on which device the calling thread happens to have selected."""
    @functools.wraps(fn)
    def wrapped(self, *args, **kwargs):
        with torch.cuda.device(self.device):
            return fn(self, *args, **kwargs)
    return wrapped


class PreparedDeAOT:
    """Device-side repack of a DeAOT model's weights (done once per model per device)."""

    deaot = True

    def __init__(self, model, device):
        sd = model.state_dict()
        self.device = device
        self.cfg = model.cfg
        self.num_layers = sum(1 for k in sd if k.startswith("LSTT.layers.") and k.endswith(".norm1.weight"))
        self.layers = []
        d = 256
        for i in range(self.num_layers):
            p = "LSTT.layers.%d." % i
            g = lambda n: sd[p + n]
            L = {}
            L["norm1"] = (_f32(g("norm1.weight"), device), _f32(g("norm1.bias"), device))
            L["norm2"] = (_f32(g("norm2.weight"), device), _f32(g("norm2.bias"), device))
            L["id_norm2"] = (_f32(g("id_norm2.weight"), device), _f32(g("id_norm2.bias"), device))
            # one GEMM for K | V | U  (linear_QV 256->128+512, linear_U 256->512)
            L["w_qvu"] = _bf(torch.cat([g("linear_QV.weight"), g("linear_U.weight")], 0), device)
            L["b_qvu"] = _f32(torch.cat([g("linear_QV.bias"), g("linear_U.bias")], 0), device)
            if i > 0:
                L["id_norm1"] = (_f32(g("id_norm1.weight"), device), _f32(g("id_norm1.bias"), device))
                L["w_idu"] = _bf(g("linear_ID_U.weight"), device)
                L["b_idu"] = _f32(g("linear_ID_U.bias"), device)
            L["w_idv"] = _bf(g("linear_ID_V.weight"), device)          # [512, 256] or [512, 512]
            L["b_idv"] = _f32(g("linear_ID_V.bias"), device)
            L["w_rel"] = _bf(g("short_term_attn.relative_emb_k.weight").reshape(225, 128), device)
            L["b_rel"] = _f32(g("short_term_attn.relative_emb_k.bias"), device)
            tapmajor = lambda w: _f32(w.reshape(1024, 25).t(), device)
            L["dw_lt"] = tapmajor(g("long_term_attn.dw_conv.conv.weight"))
            L["dw_st"] = tapmajor(g("short_term_attn.dw_conv.conv.weight"))
            L["dw_sa"] = tapmajor(g("self_attn.dw_conv.conv.weight"))
            # long-term and short-term projections share the residual add: one K = 2048 GEMM
            L["w_proj_ls"] = _bf(torch.cat([g("long_term_attn.projection.weight"),
                                            g("short_term_attn.projection.weight")], 1), device)
            L["b_proj_ls"] = _f32(g("long_term_attn.projection.bias") + g("short_term_attn.projection.bias"),
                                  device)
            # self-attention input GEMM: QK (512->128) | V1,V2 block-diagonal | U1,U2 block-diagonal
            z = torch.zeros(512, 256)
            w_sa = torch.cat([g("self_attn.linear_QK.weight").cpu(),
                              torch.cat([g("self_attn.linear_V1.weight").cpu(), z], 1),
                              torch.cat([z, g("self_attn.linear_V2.weight").cpu()], 1),
                              torch.cat([g("self_attn.linear_U1.weight").cpu(), z], 1),
                              torch.cat([z, g("self_attn.linear_U2.weight").cpu()], 1)], 0)
            L["w_sa"] = _bf(w_sa, device)
            L["b_sa"] = _f32(torch.cat([g("self_attn.linear_QK.bias"), g("self_attn.linear_V1.bias"),
                                        g("self_attn.linear_V2.bias"), g("self_attn.linear_U1.bias"),
                                        g("self_attn.linear_U2.bias")], 0), device)
            L["w_proj_sa"] = _bf(g("self_attn.projection.weight"), device)
            L["b_proj_sa"] = _f32(g("self_attn.projection.bias"), device)
            self.layers.append(L)
        self.gn_out = (_f32(sd["LSTT.decoder_norms.0.gn.weight"], device),
                       _f32(sd["LSTT.decoder_norms.0.gn.bias"], device))
        self.id_norm = (_f32(sd["id_norm.weight"], device), _f32(sd["id_norm.bias"], device))
        self.cur_pos_emb = _f32(sd["cur_pos_emb"].reshape(-1), device)
        self.mem_pos_emb = _f32(sd["mem_pos_emb"], device)
        # ID bank as a lookup table: row-prefix sums of the 17x17 kernel per class (csrc/id_embed.cu)
        wid = sd["patch_wise_id_bank.weight"].detach().float().cpu()            # [256, n_cls, 17, 17]
        self.n_id_cls = wid.shape[1]
        self.id_prefix = ops.id_prefix_table(wid).to(device)
        self.id_bias = _f32(sd["patch_wise_id_bank.bias"], device)
        self._id_w, self._id_taps = wid, None
        # black boxes (library calls): encoder / projector in bf16 channels-last with the frozen BN folded
        # into the convs, FPN head in fp32 (it decides the argmax mask)
        import os
        # RMEM_ENCODER_DTYPE=fp32 runs the cuDNN trunk in fp32 (TF32 off): used by the parity tests so mask
        # agreement measures the hot-path kernels, not the black-box backbone's bf16 rounding
        self.encoder = InferenceResNet(model.encoder, model.encoder_projector, device, encoder_dtype())
        self.decoder = InferenceFPN(sd, device, intermediate=False, shortcut_dtype=self.encoder.dtype)


    def id_taps(self):
        """patch_wise_id_bank.weight as [n_cls, 17, 17, 256] fp32 on the device (csrc/id_embed.cu, soft masks); built on
        first use -- label maps, the evaluated form, never need it."""
        if self._id_taps is None:
            self._id_taps = ops.id_tap_table(self._id_w).to(self.device)
        return self._id_taps

class InferenceFPN:
    """FPNSegmentationHead (reference decoders/fpn.py:36-68) for inference, channels-last fp32 (it decides
    the argmax mask): the convolutions are cuDNN calls (a black box of this tier, TF32 as PyTorch's default),
    GroupNorm(8)+ReLU and the bilinear upsample + shortcut add run in csrc/decoder.cu."""

    def __init__(self, sd, device, intermediate, prefix="decoder.", shortcut_dtype=BF16):
        self.intermediate = intermediate
        self.shortcut_dtype = shortcut_dtype if shortcut_dtype in (BF16, torch.float16) else BF16
        # activations and 3x3 convolutions in fp16 (fp32 accumulation; GroupNorm statistics, normalisation, interpolation
        # and the logits head in fp32) or, with RMEM_DECODER_DTYPE=fp32, everything in fp32 / TF32
        self.dtype = torch.float16 if os.environ.get("RMEM_DECODER_DTYPE", "fp16") == "fp16" else torch.float32
        cl = lambda w: w.detach().to(device=device, dtype=self.dtype).contiguous(memory_format=torch.channels_last)
        f = lambda t: _f32(t, device)
        self.w = {}
        for name in ("conv_in", "conv_16x", "conv_8x", "conv_4x"):
            self.w[name] = (cl(sd[prefix + name + ".conv.weight"]), f(sd[prefix + name + ".conv.bias"]),
                            f(sd[prefix + name + ".gn.weight"]), f(sd[prefix + name + ".gn.bias"]))
        for name in ("adapter_16x", "adapter_8x", "adapter_4x", "conv_out"):
            self.w[name] = (cl(sd[prefix + name + ".weight"]), f(sd[prefix + name + ".bias"]))
        self.w_head = f(sd[prefix + "conv_out.weight"].reshape(sd[prefix + "conv_out.weight"].shape[0], -1))   # [Co, C]
        # the 1x1 adapters read the encoder's bf16 feature maps: as [Cout, Cin] bf16 matrices they run on the
        # tcgen05 GEMM straight off those maps (no f32 copies of the shortcuts, no f32 convolution)
        # (in the backbone's 16-bit type: fp16 by default)
        self.wb = {name: sd[prefix + name + ".weight"].detach().reshape(sd[prefix + name + ".weight"].shape[0], -1)
                   .to(device=device, dtype=self.shortcut_dtype).contiguous()
                   for name in ("adapter_16x", "adapter_8x", "adapter_4x")}
        self._emulate_f16 = os_env_flag("RMEM_EXP_DECODER_F16", False)
        self.fused_stage = os_env_flag("RMEM_FUSED_GN_UP", True)
        self.gn_ws = torch.zeros(1, 17, device=device, dtype=torch.float64)  # per image: 2 x 8 group sums + counter
        self._ws = self.gn_ws       # the workspace of the call in flight (engines pass their own: they may run
                                    # concurrently on different streams and share this object through `prepare`)

    @staticmethod
    def _tok(x):
        """channels-last [n, C, H, W] -> [n*H*W, C] view of the same memory."""
        n, c, h, w = x.shape
        return x.permute(0, 2, 3, 1).reshape(n * h * w, c)

    def _conv_gn_relu(self, x, name):
        """ConvGN + ReLU: the conv bias is applied inside the GroupNorm kernels (no separate bias pass)."""
        w, b, g, be = self.w[name]
        if self._emulate_f16:      # experiment: the numerics an fp16 decoder would have (inputs, weights, outputs rounded)
            y = F.conv2d(x.half().float(), w.half().float(), None, 1, w.shape[-1] // 2).half().float()
            y = y.contiguous(memory_format=torch.channels_last)
        else:
            y = F.conv2d(x, w, None, 1, w.shape[-1] // 2).contiguous(memory_format=torch.channels_last)
        t = self._tok(y)
        # statistics per image (a batch = independent sequences), in place
        ops.groupnorm_nhwc(t, g, be, t, 8, True, self._ws, pre_bias=b, S=y.shape[0])
        return y

    ADAPTERS = (("adapter_16x", -2), ("adapter_8x", -3), ("adapter_4x", -4))

    def _adapter(self, shortcut, name):
        """1x1 adapter convolution of a bf16 shortcut as a GEMM (+ bias in the epilogue) -> f32 channels-last."""
        w, b = self.w[name]
        shortcut = shortcut.contiguous(memory_format=torch.channels_last)
        ho, wo = shortcut.shape[-2:]
        a = torch.empty(shortcut.shape[0], ho, wo, w.shape[0], device=shortcut.device,
                        dtype=torch.float32).permute(0, 3, 1, 2)
        ops.linear(self._tok(shortcut), self.wb[name], b, [(0, w.shape[0], ops.ACT_NONE, ops.OUT_F32, self._tok(a), 0)])
        return a

    def adapters(self, shortcuts):
        """The three adapter outputs, which depend on the encoder alone: the engine computes them at the end of
        the encoder pass (off the LSTT -> decoder chain when the encoder runs ahead) and hands them to __call__."""
        if any(shortcuts[i].dtype != self.shortcut_dtype for _, i in self.ADAPTERS):
            return None
        return [self._adapter(shortcuts[i], name) for name, i in self.ADAPTERS]

    def _conv(self, x, name):
        w = self.w[name][0]
        return F.conv2d(x, w, None, 1, w.shape[-1] // 2).contiguous(memory_format=torch.channels_last)

    def _gn_up_add(self, x, gn_name, shortcut, name, pre=None):
        """relu(GroupNorm(x + conv bias)) resized to the shortcut's size + adapter(shortcut): statistics + one kernel."""
        _, cb, g, be = self.w[gn_name]
        w, b = self.w[name]
        if pre is not None:
            a, b = pre, None
        elif shortcut.dtype == self.shortcut_dtype:
            a, b = self._adapter(shortcut, name), None                              # bias added by the GEMM epilogue
        else:
            a = F.conv2d(shortcut.float(), w.float(), None).contiguous(memory_format=torch.channels_last)
        hi, wi = x.shape[-2:]
        ho, wo = a.shape[-2:]
        out = torch.empty(a.shape[0], ho, wo, a.shape[1], device=a.device, dtype=self.dtype).permute(0, 3, 1, 2)
        ops.gn_upsample_add_nhwc(self._tok(x), cb, g, be, hi, wi, self._tok(a), self._tok(out), ho, wo, self._ws,
                                 a_bias=b, S=x.shape[0])
        return out

    def _up_add(self, x, shortcut, name, pre=None):
        """adapter(shortcut) + bias + bilinear(x) (identity resample when the sizes agree)."""
        w, b = self.w[name]
        if pre is not None:
            a, ta, b = pre, self._tok(pre), None
        elif shortcut.dtype == self.shortcut_dtype:
            a = self._adapter(shortcut, name)
            ta = self._tok(a)
            b = None                                                                # added by the GEMM epilogue
        else:
            a = F.conv2d(shortcut.float(), w.float(), None).contiguous(memory_format=torch.channels_last)
            ta = self._tok(a)
        hi, wi = x.shape[-2:]
        ho, wo = a.shape[-2:]
        if self.dtype == torch.float32:
            ops.upsample_add_nhwc(self._tok(x), hi, wi, ta, ta, ho, wo, a_bias=b, S=x.shape[0])   # in place
            return a
        out = torch.empty(a.shape[0], ho, wo, a.shape[1], device=a.device, dtype=self.dtype).permute(0, 3, 1, 2)
        ops.upsample_add_nhwc(self._tok(x), hi, wi, ta, self._tok(out), ho, wo, a_bias=b, S=x.shape[0])
        return out

    def __call__(self, inputs, shortcuts, gn_ws=None, adapters=None, head=False):
        self._ws = self.gn_ws if gn_ws is None else gn_ws
        pre = adapters if adapters is not None else (None, None, None)
        x = torch.cat(inputs, dim=1) if (self.intermediate and len(inputs) > 1) else inputs[-1]
        x = x.to(self.dtype).contiguous(memory_format=torch.channels_last)
        sc = shortcuts
        if self.fused_stage and not self._emulate_f16:
            # a stage = convolution (cuDNN) + statistics + one kernel for GroupNorm, ReLU, resize and adapter add
            x = self._gn_up_add(self._conv(x, "conv_in"), "conv_in", sc[-2], "adapter_16x", pre[0])
            x = self._gn_up_add(self._conv(x, "conv_16x"), "conv_16x", sc[-3], "adapter_8x", pre[1])
            x = self._gn_up_add(self._conv(x, "conv_8x"), "conv_8x", sc[-4], "adapter_4x", pre[2])
        else:
            x = self._conv_gn_relu(x, "conv_in")
            x = self._conv_gn_relu(self._up_add(x, sc[-2], "adapter_16x", pre[0]), "conv_16x")
            x = self._conv_gn_relu(self._up_add(x, sc[-3], "adapter_8x", pre[1]), "conv_8x")
            x = self._up_add(x, sc[-4], "adapter_4x", pre[2])
        if not head:
            x = self._conv_gn_relu(x, "conv_4x")
            w, b = self.w["conv_out"]
            return F.conv2d(x, w, None).float(), b      # the caller adds the bias while finishing the logits
        # the caller finishes the stage with ops.logits_head: GroupNorm + ReLU + conv_out + bias + class mask in one pass
        w4 = self.w["conv_4x"][0]
        if self._emulate_f16:
            return F.conv2d(x.half().float(), w4.half().float(), None, 1, w4.shape[-1] // 2).half().float().contiguous(
                memory_format=torch.channels_last), None
        return F.conv2d(x, w4, None, 1, w4.shape[-1] // 2).contiguous(memory_format=torch.channels_last), None


class InferenceResNet:
    """ResNet-50 trunk (stride 16) + encoder_projector for inference: frozen-BN affine folded into the
    conv weights (exact algebra), bf16 channels-last, conv + bias + ReLU (+ residual) issued as single
    fused cuDNN calls where the installed PyTorch exposes them.  A black box of this tier (cuDNN)."""

    def __init__(self, enc, projector, device, dtype=BF16):
        # `enc` may be the repo's ResNet50S16 or the reference's own ResNet (networks/encoders/resnet.py:71-194,
        # BatchNorm = FrozenBatchNorm2d, networks/layers/normalization.py:6-45, or nn.BatchNorm2d in eval mode): the
        # frozen BN affine is folded into this object's own copies of the conv weights -- the caller's module is
        # neither mutated nor deep-copied
        self.dtype = dtype

        def folded(conv, bn):
            """(w * s, shift + b * s) in f32 for conv -> bn with frozen statistics (exact algebra)."""
            w = conv.weight.detach().float()
            if bn is None:
                b = conv.bias.detach().float() if conv.bias is not None else torch.zeros(w.shape[0])
                return w, b
            eps = float(getattr(bn, "epsilon", getattr(bn, "eps", 1e-5)))
            s = bn.weight.detach().float() * (bn.running_var.detach().float() + eps).rsqrt()
            b = bn.bias.detach().float() - bn.running_mean.detach().float() * s
            if conv.bias is not None:
                b = b + conv.bias.detach().float() * s
            return w * s.view(-1, 1, 1, 1), b

        def cv(conv, bn=None):
            w, b = folded(conv, bn)
            return (w.to(device=device, dtype=dtype).contiguous(memory_format=torch.channels_last),
                    b.to(device=device, dtype=dtype), tuple(conv.stride), tuple(conv.padding))

        w1, b1 = folded(enc.conv1, enc.bn1)
        # 3 input channels starve the tensor-core conv kernels: pad to 8 (zeros) -> same result
        w1p = torch.zeros(w1.shape[0], 8, *w1.shape[2:])
        w1p[:, :3] = w1
        self.conv1 = (w1p.to(device=device, dtype=dtype).contiguous(memory_format=torch.channels_last),
                      b1.to(device=device, dtype=dtype), tuple(enc.conv1.stride), tuple(enc.conv1.padding))
        # space-to-depth form of the stem (exact algebra): out[y,x] = sum_{a,b in -2..1} sum_{dy,dx,c}
        # W[c][2a+dy+3][2b+dx+3] X[c][2(y+a)+dy][2(x+b)+dx] -- a 4x4 stride-1 conv over 12 (padded to 16) channels
        # of the 2x2 space-to-depth image that ops.pack_image_s2d writes (with the zero border built in)
        self.conv1_s2d = None
        if tuple(w1.shape[1:]) == (3, 7, 7) and tuple(enc.conv1.stride) == (2, 2) and tuple(enc.conv1.padding) == (3, 3):
            ws = torch.zeros(w1.shape[0], 16, 4, 4)
            for dy in range(2):
                for dx in range(2):
                    ch = (dy * 2 + dx) * 3
                    for a in range(-2, 2):
                        for b in range(-2, 2):
                            i, j = 2 * a + dy + 3, 2 * b + dx + 3
                            if 0 <= i < 7 and 0 <= j < 7:
                                ws[:, ch:ch + 3, a + 2, b + 2] = w1[:, :, i, j]
            self.conv1_s2d = (ws.to(device=device, dtype=dtype).contiguous(memory_format=torch.channels_last),
                              self.conv1[1], (1, 1), (0, 0))
            self.conv1_bias_f32 = b1.to(device=device, dtype=torch.float32)
        self.blocks = []
        for li in (1, 2, 3):
            stage = []
            for blk in getattr(enc, "layer%d" % li):
                if tuple(getattr(blk.conv2, "dilation", (1, 1))) != (1, 1):
                    raise _lib.RmemError("InferenceResNet: dilated bottlenecks (output stride 8) are not on the path")
                down = None
                c3 = cv(blk.conv3, blk.bn3)
                if blk.downsample is not None:
                    # relu(conv3(y) + b3 + down(x) + bd): both biases ride on conv3's fused epilogue
                    wd, bd = folded(blk.downsample[0], blk.downsample[1])
                    _, b3 = folded(blk.conv3, blk.bn3)
                    c3 = (c3[0], (b3 + bd).to(device=device, dtype=dtype), c3[2], c3[3])
                    down = (wd.to(device=device, dtype=dtype).contiguous(memory_format=torch.channels_last), None,
                            tuple(blk.downsample[0].stride), tuple(blk.downsample[0].padding))
                stage.append((cv(blk.conv1, blk.bn1), cv(blk.conv2, blk.bn2), c3, down))
            self.blocks.append(stage)
        self.proj = cv(projector)
        self.fused = (hasattr(torch, "cudnn_convolution_relu") and os_env_flag("RMEM_FUSED_CONV", True)
                      and torch.device(device).type == "cuda")
        # measured: forking the three projection shortcuts costs ~2 % (980 vs 1000 frames/s) -- the convolutions
        # already fill the GPU and the extra graph edges add latency -- so it stays off
        self.fork = self.fused and os_env_flag("RMEM_FORK_ENCODER", False)
        self._side_streams = {}                    # one side stream per calling stream (engines may run concurrently)

    @staticmethod
    def _conv(x, c, relu):
        w, b, stride, pad = c
        if relu:
            return torch.cudnn_convolution_relu(x, w, b, stride, pad, (1, 1), 1)
        return F.conv2d(x, w, b, stride, pad)

    def __call__(self, img):
        if self.dtype == torch.float32:
            with torch.backends.cudnn.flags(enabled=True, allow_tf32=False):
                return self._forward(img)
        return self._forward(img)

    # normalisation of the reference's dataloader (dataloaders/video_transforms.py:537-539, 676-678), applied on
    # the device when a caller hands the engine a decoded u8 frame [1, H, W, 3] instead of normalised f32 planes
    U8_MEAN, U8_STD = (0.485, 0.456, 0.406), (0.229, 0.224, 0.225)

    def _forward(self, img):
        half = self.dtype in (BF16, torch.float16)
        if img.dtype == torch.uint8:
            if img.dim() == 3:
                img = img[None]
            n, h, w = img.shape[0], img.shape[1], img.shape[2]
        else:
            n, c, h, w = img.shape
        # n > 1: a batch of independent sequences stepping in lock-step (the convolutions take the batch; the small
        # per-image kernels around them are issued per image)
        s2d = (half and img.dtype in (torch.float32, torch.uint8) and self.fused and self.conv1_s2d is not None)
        if img.dtype == torch.uint8 and not s2d:
            img = img.permute(0, 3, 1, 2).float().div(255.0)
            img = (img - img.new_tensor(self.U8_MEAN).view(1, 3, 1, 1)) / img.new_tensor(self.U8_STD).view(1, 3, 1, 1)
        if s2d:
            x = torch.empty(n, (h + 1) // 2 + 3, (w + 1) // 2 + 3, 16, device=img.device,
                            dtype=self.dtype).permute(0, 3, 1, 2)
            # f32 planes / u8 pixels -> 16-bit space-to-depth image, conv padding built in
            ops.pack_image_s2d(img, x, self.U8_MEAN, self.U8_STD)
        elif self.dtype == BF16 and n == 1 and img.dtype == torch.float32:
            x = torch.empty(1, 8, h, w, device=img.device, dtype=BF16, memory_format=torch.channels_last)
            ops.pack_image(img, x)                  # f32 planes -> bf16 channels-last, channels 3..7 = 0
        else:
            x = torch.zeros(n, 8, h, w, device=img.device, dtype=self.dtype).contiguous(
                memory_format=torch.channels_last)
            x[:, :3] = img
        if s2d:
            # bias + ReLU commute with the max-pool (monotone per channel): applied by the pool kernel on a
            # quarter of the pixels instead of by a separate pass over the full-resolution map
            x = F.conv2d(x, self.conv1_s2d[0], None)
        elif self.fused:
            x = self._conv(x, self.conv1, True)
        else:
            x = F.relu(F.conv2d(x, self.conv1[0], self.conv1[1], self.conv1[2], self.conv1[3]))
        if half and (n == 1 or s2d):
            pooled = torch.empty(n, x.shape[1], (x.shape[2] - 1) // 2 + 1, (x.shape[3] - 1) // 2 + 1,
                                 device=x.device, dtype=self.dtype, memory_format=torch.channels_last)
            c_, h_, w_ = x.shape[1:]
            if x.stride()[1:] != (1, w_ * c_, c_):                   # cuDNN returns NHWC strides; never copy then
                x = x.contiguous(memory_format=torch.channels_last)
            ops.maxpool3s2_nhwc(x, pooled, self.conv1_bias_f32 if s2d else None, s2d)
            x = pooled
        else:
            x = F.max_pool2d(x, 3, 2, 1)
        xs = []
        for stage in self.blocks:
            for (c1, c2, c3, down) in stage:
                forked = down is not None and self.fork
                if forked:
                    # the projection shortcut only shares its input with conv1 -> conv2: a parallel branch (side
                    # stream; inside a captured graph a parallel path) that meets the trunk again in conv3's add
                    cur = torch.cuda.current_stream()
                    side = self._side_streams.get(cur.cuda_stream)
                    if side is None:
                        side = self._side_streams[cur.cuda_stream] = torch.cuda.Stream(device=x.device)
                    ev = torch.cuda.Event()
                    ev.record()
                    side.wait_event(ev)
                    with torch.cuda.stream(side):
                        idt = F.conv2d(x, down[0], down[1], down[2], down[3])
                        done = torch.cuda.Event()
                        done.record()
                    if not torch.cuda.is_current_stream_capturing():      # graph pools never recycle mid-capture
                        x.record_stream(side)
                        idt.record_stream(cur)
                else:
                    idt = x if down is None else F.conv2d(x, down[0], down[1], down[2], down[3])
                if self.fused:
                    y = self._conv(self._conv(x, c1, True), c2, True)
                    if forked:
                        torch.cuda.current_stream().wait_event(done)
                    x = torch.cudnn_convolution_add_relu(y, c3[0], idt, 1.0, c3[1], c3[2], c3[3], (1, 1), 1)
                else:
                    y = F.relu(F.conv2d(x, c1[0], c1[1], c1[2], c1[3]))
                    y = F.relu(F.conv2d(y, c2[0], c2[1], c2[2], c2[3]))
                    x = F.relu(F.conv2d(y, c3[0], c3[1], c3[2], c3[3]) + idt)
            xs.append(x)
        xs.append(F.conv2d(x, self.proj[0], self.proj[1]))
        return xs


# fp16, not bf16: same storage and tensor-core rate, 3 more mantissa bits -- measured on the reference fixtures,
# masks agree >= 99.92 % with the fp32 reference on every frame with an fp16 backbone and only 99.78-99.85 % with a
# bf16 one (profiles/r02_exp_encoder_dtype.log); the reference's own --amp path is fp16 autocast as well
ENCODER_DTYPE_DEFAULT = "fp16"


def encoder_dtype():
    """Arithmetic type of the cuDNN backbone (a black box of this tier): RMEM_ENCODER_DTYPE = bf16 | fp16 | fp32."""
    import os
    name = os.environ.get("RMEM_ENCODER_DTYPE", ENCODER_DTYPE_DEFAULT)
    try:
        return {"bf16": BF16, "fp16": torch.float16, "fp32": torch.float32}[name]
    except KeyError:
        raise _lib.RmemError("RMEM_ENCODER_DTYPE must be bf16, fp16 or fp32, not %r" % name)


class _EmbList(list):
    """A list of embedding tensors that may carry the `_rmem_static` mark (see DeAOTEngine.encode_static)."""


def os_env_flag(name, default):
    import os
    v = os.environ.get(name)
    return default if v is None else v not in ("0", "false", "False")


def prepare(model, device):
    cache = model.__dict__.setdefault("_rmem_prepared", {})
    import os
    key = (str(device), os.environ.get("RMEM_ENCODER_DTYPE", ENCODER_DTYPE_DEFAULT),
           os.environ.get("RMEM_DECODER_DTYPE", "fp32"))
    if key not in cache:
        if is_deaot_model(model):
            cache[key] = PreparedDeAOT(model, device)
        else:
            from .engine_aot import PreparedAOT
            cache[key] = PreparedAOT(model, device, InferenceResNet, InferenceFPN, encoder_dtype())
    return cache[key]


def is_deaot_model(model):
    """DeAOT (gated propagation, two branches) or AOT (multi-head attention), told apart by the parameters."""
    return any(k.endswith("linear_QV.weight") for k in model.state_dict().keys())


class _Workspace:
    """Per-sequence device buffers for one token grid (h, w); allocated once, reused every frame."""

    def __init__(self, prep, h, w, n_slots, device, S=1):
        """S = number of independent sequences stepping in lock-step: every token matrix is the stack [S * L, C] of
        the sequences' matrices (row-wise kernels take it as it is, spatial kernels get S as a grid dimension)."""
        L, nl = h * w, prep.num_layers
        f32 = torch.float32
        R = S * L                                             # rows of a stacked token matrix

        def z(*s, dt=BF16):
            return torch.zeros(*s, device=device, dtype=dt)

        self.h, self.w, self.L, self.n_slots, self.S = h, w, L, n_slots, S
        self.tgt, self.tgt_id = z(R, 256, dt=f32), z(R, 256, dt=f32)
        self.xn = z(R, 512)                                   # LayerNorm outputs (norm | id_norm)
        self.idcat = [z(R, 512) for _ in range(nl)]           # [curr_ID_V | id_emb] per layer
        # K and V | ID_V of a frame, two sets: the valid short-term memory (previous frame) and the set the
        # current frame is written into; the write path completes ID_V in place and the roles swap (no copies)
        self.kst = [[z(R, 128) for _ in range(nl)] for _ in range(2)]
        self.vst = [[z(R, 1024) for _ in range(nl)] for _ in range(2)]
        self.bind(0)
        self.kbank = [z(S, n_slots, L, 128) for _ in range(nl)]   # long-term bank, one per sequence
        self.vbank = [z(S, n_slots, L, 1024) for _ in range(nl)]
        self.u = [z(R, 1024) for _ in range(nl)]              # cat_curr_U
        self.u[0][:, 512:] = 1.0                              # layer 0: U = [silu(U), 1]
        self.qt, self.tpe_bias = z(R, 128), z(R, 4, dt=f32)
        self.gated_a, self.gated_b = z(R, 1024), z(R, 1024)
        self.dw = z(R, 2048)
        self.rel = z(R, 256, dt=f32)
        self.banded = ops.local_banded_buffer(h, w, device, S)    # local-attention probabilities, banded per tile
        self.sa_qk, self.sa_v, self.sa_u = z(R, 128), z(R, 1024), z(R, 1024)
        if S == 1:                                            # partials of the first-generation attention (T = 1)
            n_chunks = max(T * _chunks_per_frame(T, L) for T in range(1, n_slots + 1))
            self.part_o = z(n_chunks, L, 1024, dt=f32)
            self.part_m, self.part_l = z(n_chunks, L, dt=f32), z(n_chunks, L, dt=f32)
        self.mass = z(R * _lib.MAX_BANK, dt=f32)              # layer-0 record_attn_weight, [S][L][T] packed
        self.gn_out = z(R, 512, dt=prep.decoder.dtype)        # in the decoder's activation type: its first convolution reads it as it is
        self.gn_ws = z(5 * S, dt=torch.float64)               # per sequence: 2 x 2 group sums + arrival counter
        self.fg = z(S, L, dt=f32)
        self.id_emb = z(R, 256)
        self.bank_state = z(S, _lib.BANK_STATE_WORDS, dt=torch.int32)
        self.zero_order = z(S, dt=torch.int32)
        self.sk_ws = None                                     # stream-K attention workspace, allocated on first use
        self.pred_logits = None                               # [1, 11, H/4, W/4] f32, allocated on first decode


def default_sk_clusters(device):
    """2-CTA clusters the stream-K long-term attention may occupy inside the frame graph: ~3/8 of the SM pairs
    (56 on a 148-SM B200; measured 48 / 52 / 56 / 60 / 64 / 68 / all: 56 is the maximum of frames/s) -- the
    short-term branch and the next frame's encoder run beside it and need SMs of their own."""
    n_sms = torch.cuda.get_device_properties(device).multi_processor_count
    return max(1, int(0.38 * n_sms + 0.5))


def _chunks_per_frame(T, L, n_sms=148):
    """Split-KV granularity of the long-term attention grid (q tiles x 4 value splits x T x cpf CTAs, one
    CTA per SM): the cpf in 1..4 that minimises waves x key tiles per chunk."""
    q_tiles = -(-L // 128)
    best, best_cost = 1, None
    for cpf in range(1, _lib.MAX_CHUNKS_PER_FRAME + 1):
        ctas = q_tiles * 4 * T * cpf
        tiles = -(-(-(-L // cpf)) // 128)
        cost = -(-ctas // n_sms) * (tiles + 1.5)          # + fixed prologue / epilogue per CTA
        if best_cost is None or cost < best_cost:
            best, best_cost = cpf, cost
    return best


def _ws_bind(ws, st):
    """Point kprev / vprev at short-term set `st` and kcur / vcur at the other one."""
    ws.kprev, ws.vprev = ws.kst[st], ws.vst[st]
    ws.kcur, ws.vcur = ws.kst[1 - st], ws.vst[1 - st]


_Workspace.bind = _ws_bind


class DeAOTEngine:
    """Per-sequence state machine with the reference AOTEngine / DeAOTEngine surface
    (aot_engine.py:18-568, deaot_engine.py:9-18): add_reference_frame, match_propogate_one_frame,
    update_short_term_memory, decode_current_logits, restart_engine, update_size."""

    def __init__(self, aot_model, gpu_id=0, long_term_mem_gap=9999, short_term_mem_skip=1,
                 layer_loss_scaling_ratio=2., n_sequences=1):
        """``n_sequences`` (not in the reference, whose recorded-attention path asserts batch size 1,
        transformer.py:641): S > 1 steps S independent sequences of one frame size in lock-step through the same
        launches -- images [S, 3, H, W], label maps [S, 1, H, W], obj_nums a list of S counts, logits [S, C, H, W],
        ``long_memories_indexes`` a list of S lists.  Every sequence keeps its own bank and its own eviction history;
        the sequence is a grid dimension of the kernels (one sequence alone is a fraction of a wave on 148 SMs)."""
        if not torch.cuda.is_available():
            raise _lib.RmemError("rmem_ocu_b200 engines need a CUDA device: there is no CPU fallback")
        _lib.lib()  # fail loudly now if the shared library is missing
        if short_term_mem_skip != 1:
            raise NotImplementedError("short_term_mem_skip != 1 is not on the evaluated path")
        self.AOT = aot_model
        self.cfg = aot_model.cfg
        self.align_corners = True
        self.max_obj_num = aot_model.max_obj_num
        self.gpu_id = gpu_id
        self.device = torch.device("cuda", gpu_id)
        self.long_term_mem_gap = long_term_mem_gap
        self.short_term_mem_skip = short_term_mem_skip
        self.S = int(n_sequences)
        if self.S < 1:
            raise ValueError("n_sequences must be >= 1")
        self.prep = prepare(aot_model, self.device)
        self.capacity = int(self.cfg.FORMER_MEM_LEN) + int(self.cfg.LATTER_MEM_LEN)
        if self.capacity + 1 > _lib.MAX_BANK:
            raise _lib.RmemError("FORMER_MEM_LEN + LATTER_MEM_LEN + 1 exceeds RMEM_MAX_BANK")
        self.deaot = self.prep.deaot
        if self.S > 1 and not self.deaot:
            raise NotImplementedError("n_sequences > 1 is implemented for the DeAOT family")
        if self.deaot != self._family_deaot:
            raise _lib.RmemError("%s was given a%s model" % (type(self).__name__, "n AOT" if self._family_deaot
                                                             else " DeAOT"))
        self.ws = None
        self.trace = None          # optional dict: per-layer intermediates for the parity tests
        self.use_cuda_graphs = os_env_flag("RMEM_CUDA_GRAPHS", True)
        self.fork_branches = os_env_flag("RMEM_FORK_BRANCHES", True)
        # single-kernel local attention (csrc/local_attn_fused.cu): parity-green but 17.5 us against 16.5 us for
        # the scores + aggregation pair at 480p, so it is opt-in until it wins
        self.local_fused = os_env_flag("RMEM_LOCAL_FUSED", False) and self.S == 1
        import os
        self.lt_streamk = self.deaot and os_env_flag("RMEM_LT_STREAMK", True)
        self.fused_head = os_env_flag("RMEM_FUSED_HEAD", True)    # decoder tail: GN + ReLU + conv_out + mask in one pass
        self.pred_label = None       # argmax of the last frame's logits at the output size (f32 [S, 1, Ho, Wo]) or None
        self.ln_copy = os_env_flag("RMEM_LN_COPY", True)          # layer 0's LayerNorm starts the fp32 residual stream
        self.label_alias = os_env_flag("RMEM_LABEL_ALIAS", True)  # pred_label lives in the write path's label buffer
        self.check_prefetch = os_env_flag("RMEM_CHECK_PREFETCH", False)   # compare each frame with the one encoded ahead
        self.lt_single = os_env_flag("RMEM_LT_SINGLE", True)      # one-launch self-attention (no partials / combine)
        # 0 = one cluster per SM pair; k > 0 caps the clusters so that the short-term branch running beside the
        # long-term attention finds free SMs
        self.lt_sk_clusters = int(os.environ.get("RMEM_LT_SK_CLUSTERS", "-1"))
        if self.lt_sk_clusters < 0:
            self.lt_sk_clusters = default_sk_clusters(self.device) if self.fork_branches else 0
        self._side = None
        self._cap_stream = None
        self._graphs_baked = None
        # the write path (update_short_term_memory) runs on its own stream: the next frame's encoder does not
        # depend on it, so the two overlap; the next frame's LSTT waits for it (self._w_done)
        self.async_write = os_env_flag("RMEM_ASYNC_WRITE", True)
        self._wstream = None
        self._w_done = None
        self._graph_pool_w = None
        # deferred write path (DeAOT): update_short_term_memory only stores the label map; the next
        # match_propogate_one_frame replays ONE graph holding that write path as a parallel branch of the LSTT
        # (joined where the new short-term memory is first read), which removes two graph boundaries from the
        # frame-to-frame dependency chain and hides the write path behind layer 0's long-term attention.
        # Anything else that reads the memory (reference frames, bank indexes) runs the pending write path first.
        self.lazy_write = self.deaot and os_env_flag("RMEM_LAZY_WRITE", True)
        self._wside = None
        self._label_buf = None
        self._w_events = None
        # encoder prefetch (match_propogate_one_frame(..., next_img=)): the next frame's encoder runs on its own
        # stream, into one of two alternating sets of embeddings, while this frame's LSTT + decoder run
        self._estream = None
        self._enc_done = None        # event: the prefetched embeddings are complete
        self._pq = collections.deque()   # frames encoded ahead: (embeddings, (set, batch, index)) in call order
        self._enc_set = 0            # the embedding set (encoder graph 0 / 1) the next prefetch writes
        self._hp_stream = None
        self._graphs = {}
        self._graph_pool = None
        self.restart_engine()

    _family_deaot = True

    def _new_workspace(self, h, w, n_slots):
        return _Workspace(self.prep, h, w, n_slots, self.device, self.S)

    @staticmethod
    def _tpe_rows(T):
        return tpe_row_indices(T)

    def _decoder_inputs(self, embs, lstt_out):
        n, c, h, w = embs[-1].shape
        return [embs[-1], lstt_out.view(n, h, w, -1).permute(0, 3, 1, 2)]   # channels-last view, no copy

    # ------------------------------------------------------------------ bookkeeping
    def restart_engine(self, batch_size=1, enable_id_shuffle=False):
        if batch_size != 1 or enable_id_shuffle:
            raise NotImplementedError("evaluation path: batch_size 1, no id shuffle")
        self.batch_size = 1
        self.frame_step = 0
        self.last_mem_step = -1
        self.obj_nums = None
        self.enc_size_2d = None
        self.enc_hw = None
        self.input_size_2d = None
        self.pred_id_logits = None if getattr(self, "ws", None) is None else self.ws.pred_logits
        self.st = 0                # which K / V set holds the valid short-term memory (host mirror)
        self.T = 0                 # frames in the long-term bank (host mirror; deterministic)
        self.T_recorded = 0        # bank length when record_attn_weight was last written
        self._pending_w = None     # (T, update_long, do_evict, st) of a deferred write path
        self._prefetching = False  # the last propagated frame started the next frame's encoder
        self._img_sig = None       # (dtype, shape) of the sequence's frames, fixed by the reference frame
        self.n_long_updates = 0

    @_on_device
    def update_size(self, input_size, enc_size):
        self.input_size_2d = tuple(input_size)
        self.enc_size_2d = tuple(enc_size)
        self.enc_hw = self.enc_size_2d[0] * self.enc_size_2d[1]
        h, w = self.enc_size_2d
        n_slots = self.capacity + 1
        if self.ws is None or (self.ws.h, self.ws.w, self.ws.n_slots) != (h, w, n_slots):
            self.ws = self._new_workspace(h, w, n_slots)
            if self.lt_streamk:          # allocated here, never inside a graph capture
                self.ws.sk_ws = ops.lt_attn_sk_workspace(self.ws.L, n_slots, self.device, self.lt_sk_clusters, self.S)
            self._wait_write()
            self._graphs, self._graph_pool, self._graph_pool_w = {}, None, None   # graphs bake the workspace pointers
            self._pq.clear()
            self._enc_set = 0

    @property
    def long_memories_indexes(self):
        """Frame ids in the bank, oldest first (host mirror; reading it synchronises)."""
        if self.ws is None or self.T == 0:
            return []
        with torch.cuda.device(self.device):
            self._wait_write()
            st = self.ws.bank_state.cpu()
        per_seq = [[int(st[s, _lib.MAX_BANK + int(st[s, t])]) for t in range(self.T)] for s in range(self.S)]
        return per_seq[0] if self.S == 1 else per_seq

    @_on_device
    def eviction_debug(self, sequence=0):
        self._wait_write()
        st = self.ws.bank_state[sequence].cpu()
        base = 5 * _lib.MAX_BANK
        return dict(last_drop=int(st[base]), last_drop_frame=int(st[base + 1]),
                    margin=float(st[base + 2:base + 3].view(torch.float32)), n_evictions=int(st[base + 3]),
                    last_choice=int(st[base + 4]))

    # ------------------------------------------------------------------ black boxes
    @_on_device
    def encode_image(self, img):
        """ResNet-50 trunk + encoder_projector (cuDNN, bf16 channels-last) -> [x4, x8, x16, emb16]."""
        return self.prep.encoder(img.to(device=self.device, non_blocking=True))

    @_on_device
    def encode_static(self, img):
        """encode_image + adapters through this engine's encoder graph: the outputs live at stable addresses, so other
        engines (the object groups of one sequence) can replay their read graphs on them."""
        if self._img_sig is not None and (img.dtype, tuple(img.shape)) != self._img_sig:
            raise _lib.RmemError("frames of one sequence must share the reference frame's dtype and shape")
        embs = _EmbList(self._run(("enc", 0, 1), self._encode_body, [img.to(self.device, non_blocking=True)]))
        embs._rmem_static = self.use_cuda_graphs and ("enc", 0, 1) in self._graphs
        return embs

    def _encode_body(self, img):
        """The encoder graph: encode_image + the decoder's adapter convolutions of its shortcuts (they depend on
        the encoder alone) -> [x4, x8, x16, emb16, a16, a8, a4]."""
        embs = list(self.encode_image(img))
        pre = self.prep.decoder.adapters(embs)
        return embs + pre if pre is not None else embs

    def _tokens(self, x_nchw):
        """[S, C, h, w] channels-last tensor -> [S * L, C] view of the same memory."""
        n, c, h, w = x_nchw.shape
        return x_nchw.permute(0, 2, 3, 1).reshape(n * h * w, c)

    def _id_embed(self, label_hw):
        """assign_identity + get_id_emb (aot_engine.py:208-232, models/deaot.py:65-69) from a hard label
        map [H, W] (0..10 object ids, 255 = ignore): one gather kernel writes id_emb next to curr_ID_V
        for every layer's ID_V GEMM."""
        ws, prep = self.ws, self.prep
        g, b = prep.id_norm
        dsts = [(ws.id_emb, 0)] + [(ws.idcat[i], 256) for i in range(prep.num_layers)]
        if label_hw.dim() == 4:                 # one-hot / probability masks [S, n_ch, H, W]: the conv as a weighted sum
            ops.id_embed_soft(label_hw, None, prep.id_taps(), prep.id_bias, g, b, prep.n_id_cls, dsts, ws.h, ws.w)
            return
        ops.id_embed(label_hw, prep.id_prefix, prep.id_bias, g, b, prep.n_id_cls, dsts, ws.h, ws.w)   # [S, H, W] labels

    def _label_map(self, mask, allow_soft=False):
        """Accepts [H,W], [S,H,W] or [S,1,H,W] label maps -> [S, H, W].  With ``allow_soft`` a 4-D tensor with more than
        one channel is a one-hot / probability mask [S, max_obj + 1, H, W] (update_short_term_memory's second branch,
        aot_engine.py:330-336; the reference reaches it with batch > 1 only, because it tells the two forms apart by
        ``size()[0] == 1`` -- here the channel count decides, so one sequence can use it too) and is returned as
        [S, n_ch, H, W] fp32."""
        if mask.dim() == 4 and mask.shape[1] != 1:
            if not allow_soft:
                raise ValueError("the reference frame takes a label map (aot_engine.py:162-164), got a %s tensor" %
                                 (tuple(mask.shape),))
            if mask.shape[0] != self.S or mask.shape[1] > self.prep.n_id_cls:
                raise ValueError("expected one-hot / probability masks [%d, <= %d, H, W], got %s" %
                                 (self.S, self.prep.n_id_cls, tuple(mask.shape)))
            return mask.float().contiguous()
        if mask.numel() != self.S * mask.shape[-2] * mask.shape[-1]:
            raise ValueError("expected label maps for %d sequence(s), got a tensor of shape %s" %
                             (self.S, tuple(mask.shape)))
        return mask.reshape(self.S, mask.shape[-2], mask.shape[-1])

    # ------------------------------------------------------------------ LSTT (kernels)
    def _id_value(self, i, dst):
        """fuse_key_value_id (transformer.py:1238-1244): ID_V = silu(linear_ID_V([curr_ID_V | id_emb]))
        written into columns 512.. of ``dst`` (a V | ID_V buffer)."""
        ws, Lw = self.ws, self.prep.layers[i]
        a = ws.idcat[i][:, 256:] if i == 0 else ws.idcat[i]
        ops.linear(a, Lw["w_idv"], Lw["b_idv"], [(0, 512, ops.ACT_SILU, ops.OUT_BF16, dst, 512)])

    def _gated_global_attention(self, q, kbank, vbank, order, T, tpe, gate, out, mass, gate_ready=None,
                                order_stride=0):
        """kbank / vbank: [S, n_slots, L, C]; order: int32, sequence s's slot table order_stride words after s - 1's."""
        ws = self.ws
        if (self.lt_streamk and T >= 2) or ws.S > 1:
            # second-generation kernel (csrc/lt_attn_sk.cu): persistent stream-K on 2-CTA clusters; wins from two
            # frames in the bank on (one sequence at T = 1, i.e. its self-attention and a fresh bank, stays on the
            # first generation); the sequence-batched form exists in this kernel only
            ops.lt_attn_sk_partial(q, kbank, vbank, order, T, ws.tpe_bias if tpe else None,
                                   tpe_row_indices(T) if tpe else None, mass is not None, ws.sk_ws,
                                   self.lt_sk_clusters, ws.S, order_stride)
            if gate_ready is not None:
                torch.cuda.current_stream().wait_event(gate_ready)
            ops.lt_attn_sk_combine(T, ws.L, gate, out, mass, ws.sk_ws, self.lt_sk_clusters, ws.S)
            return
        kbank, vbank = kbank.view(-1, ws.L, kbank.shape[-1]), vbank.view(-1, ws.L, vbank.shape[-1])
        if T == 1 and mass is None and order is ws.zero_order and self.lt_single:
            # one frame of keys in slot 0 and nothing to record (a layer's self-attention): the softmax completes
            # inside one CTA per (query tile, value split), so the kernel writes the gated bf16 output itself
            if gate_ready is not None:
                torch.cuda.current_stream().wait_event(gate_ready)
            ops.lt_attn_single(q, kbank[0], vbank[0], gate, out)
            return
        cpf = _chunks_per_frame(T, ws.L)
        ops.lt_attn_partial(q, kbank, vbank, order, T, cpf, ws.part_o, ws.part_m, ws.part_l)
        if gate_ready is not None:
            torch.cuda.current_stream().wait_event(gate_ready)    # the gate is computed on another branch
        ops.lt_attn_combine(ws.part_o, ws.part_m, ws.part_l, ws.tpe_bias if tpe else None,
                            tpe_row_indices(T) if tpe else None, T, ws.L, cpf, gate, out, mass)

    def _fork(self):
        """Make the side stream wait for everything issued so far on the current stream."""
        if not self.fork_branches:
            return False
        if self._side is None:
            self._side = torch.cuda.Stream(device=self.device)
        ev = torch.cuda.Event()
        ev.record()
        self._side.wait_event(ev)
        return True

    def _join(self):
        """Make the current stream wait for everything issued on the side stream."""
        ev = torch.cuda.Event()
        ev.record(self._side)
        torch.cuda.current_stream().wait_event(ev)

    def _lstt_forward(self, curr_emb16, is_reference, save_atten_weights, T):
        """DualBranchGPM.forward (transformer.py:765-824): three GatedPropagationModule blocks + output norm."""
        ws, prep = self.ws, self.prep
        tok = self._tokens(curr_emb16)
        self._tgt_src = tok if (self.ln_copy and self.deaot and tok.is_contiguous() and
                                tok.dtype in (BF16, torch.float16)) else None
        if self._tgt_src is None:
            ws.tgt.copy_(tok)                 # (the DeAOT block's first LayerNorm makes this copy itself)
        we = self._w_events
        if we is not None and prep.num_layers == 1:
            we["l0"] = we["all"]
        for i in range(prep.num_layers):
            self._block_forward(i, is_reference, save_atten_weights and i == 0, T)
        if we is not None and not we["waited"]:
            torch.cuda.current_stream().wait_event(we["all"])
            we["waited"] = True
        # the single output GroupNorm1D(512, 2) (transformer.py:806-808)
        ops.groupnorm_tokens(ws.tgt, ws.tgt_id, *prep.gn_out, ws.gn_out, 2, ws.gn_ws, S=ws.S)
        return ws.gn_out

    def _block_forward(self, i, is_reference, record, T):
        """GatedPropagationModule.forward (transformer.py:1091-1236) of layer ``i`` on the workspace's residual
        streams; ``record``: write this layer's record_attn_weight into ws.mass (the engines record layer 0)."""
        ws, prep = self.ws, self.prep
        h, w, L, S = ws.h, ws.w, ws.L, ws.S
        Lw = prep.layers[i]
        tr = self.trace
        we = self._w_events
        if we is not None and i == 1 and not we["waited"]:
            # a deferred write path is running beside layer 0: its ID_V GEMMs read curr_ID_V of the previous
            # frame, which the next LayerNorm overwrites, and write the short-term values layers 1.. read
            torch.cuda.current_stream().wait_event(we["all"])
            we["waited"] = True
        # norm1 (+ id_norm1 -> curr_ID_V, kept next to id_emb for the write path)
        if i == 0 and getattr(self, "_tgt_src", None) is not None:
            # norm1 of layer 0 straight off the encoder's 16-bit embedding; the kernel also starts the fp32 residual stream
            ops.layernorm_copy(self._tgt_src, *Lw["norm1"], ws.xn, 0, ws.tgt)
        elif i == 0:
            ops.layernorm(ws.tgt, *Lw["norm1"], ws.xn, 0)
        else:
            ops.layernorm(ws.tgt, *Lw["norm1"], ws.xn, 0, ws.tgt_id, *Lw["id_norm1"], ws.idcat[i], 0)
        # K (= Q) | silu(V) | silu(U) in one GEMM
        ops.linear(ws.xn[:, :256], Lw["w_qvu"], Lw["b_qvu"],
                   [(0, 128, ops.ACT_NONE, ops.OUT_BF16, ws.kcur[i], 0),
                    (128, 640, ops.ACT_SILU, ops.OUT_BF16, ws.vcur[i], 0),
                    (640, 1152, ops.ACT_SILU, ops.OUT_BF16, ws.u[i], 0)])
        idu = (lambda: ops.linear(ws.idcat[i][:, :256], Lw["w_idu"], Lw["b_idu"],
                                  [(0, 512, ops.ACT_SILU, ops.OUT_BF16, ws.u[i], 512)])) if i > 0 else None
        if idu is not None and not self.fork_branches:
            idu()
        if is_reference:
            # the reference frame is its own long- and short-term memory (transformer.py:1125-1135)
            self._id_value(i, ws.vcur[i])
            ws.kprev[i].copy_(ws.kcur[i]); ws.vprev[i].copy_(ws.vcur[i])
            ws.kbank[i][:, 0].copy_(ws.kcur[i].view(S, L, 128)); ws.vbank[i][:, 0].copy_(ws.vcur[i].view(S, L, 1024))
        # The long-term and the short-term branch only share their inputs (K = Q, U) and meet again in the
        # projection GEMM: the short-term branch is forked onto a side stream (a parallel branch of the
        # captured graph), so its sub-wave kernels fill the SMs the long-term attention's ragged second wave
        # leaves idle.
        forked = self._fork()
        idu_done = None
        with torch.cuda.stream(self._side if forked else torch.cuda.current_stream()):
            if idu is not None and forked:
                # the ID half of the gate U is first read by the attention epilogues (combine / aggregate):
                # its GEMM leaves the long-term chain for the short-term branch, which has slack
                idu()
                idu_done = torch.cuda.Event()
                idu_done.record()
            # short-term 15x15 local attention over the previous frame, gate, dw5x5
            ops.linear(ws.kcur[i], Lw["w_rel"], Lw["b_rel"], [(0, 225, ops.ACT_NONE, ops.OUT_F32, ws.rel, 0)])
            if we is not None and i == 0 and not we["waited"]:
                torch.cuda.current_stream().wait_event(we["l0"])     # ID_V of layer 0 (deferred write path)
            if self.local_fused:
                ops.local_attn(ws.kcur[i], ws.kprev[i], ws.vprev[i], ws.rel, ws.u[i], ws.gated_b, h, w)
            else:
                ops.local_attn_scores(ws.kcur[i], ws.kprev[i], ws.rel, None, ws.banded, h, w, S=S)
                ops.local_attn_aggregate(ws.banded, ws.vprev[i], ws.u[i], ws.gated_b, h, w, S=S)
            ops.dwconv5(ws.gated_b, 0, Lw["dw_st"], ws.dw, 1024, h, w, 1024, S=S)
        # long-term attention over the bank (TPE folded into a per-frame score bias), gate, dw5x5
        ops.tpe_prepare(ws.kcur[i], prep.cur_pos_emb, prep.mem_pos_emb, ws.qt, ws.tpe_bias)
        mass = None
        if record:
            mass = ws.mass
        self._gated_global_attention(ws.qt, ws.kbank[i], ws.vbank[i], ws.bank_state, T, True, ws.u[i],
                                     ws.gated_a, mass, gate_ready=idu_done, order_stride=_lib.BANK_STATE_WORDS)
        ops.dwconv5(ws.gated_a, 0, Lw["dw_lt"], ws.dw, 0, h, w, 1024, S=S)
        if forked:
            self._join()
        # (projection_lt + projection_st) -> residual adds, one GEMM
        if tr is not None:
            tr.setdefault("layers", []).append(dict(
                curr_Q=ws.kcur[i].float().cpu(), curr_V=ws.vcur[i][:, :512].float().cpu(),
                cat_curr_U=ws.u[i].float().cpu(), lt_gated=ws.gated_a.float().cpu(),
                st_gated=ws.gated_b.float().cpu()))
        id_mode = ops.OUT_F32 if i == 0 else ops.OUT_F32_ACCUM
        ops.linear(ws.dw, Lw["w_proj_ls"], Lw["b_proj_ls"],
                   [(0, 256, ops.ACT_NONE, ops.OUT_F32_ACCUM, ws.tgt, 0),
                    (256, 512, ops.ACT_NONE, id_mode, ws.tgt_id, 0)])
        # self-attention (gated, d_att 128, value width 1024)
        ops.layernorm(ws.tgt, *Lw["norm2"], ws.xn, 0, ws.tgt_id, *Lw["id_norm2"], ws.xn, 256)
        ops.linear(ws.xn, Lw["w_sa"], Lw["b_sa"],
                   [(0, 128, ops.ACT_NONE, ops.OUT_BF16, ws.sa_qk, 0),
                    (128, 1152, ops.ACT_SILU, ops.OUT_BF16, ws.sa_v, 0),
                    (1152, 2176, ops.ACT_SILU, ops.OUT_BF16, ws.sa_u, 0)])
        self._gated_global_attention(ws.sa_qk, ws.sa_qk.view(S, 1, L, 128), ws.sa_v.view(S, 1, L, 1024),
                                     ws.zero_order, 1, False, ws.sa_u, ws.gated_a, None, order_stride=1)
        ops.dwconv5(ws.gated_a, 0, Lw["dw_sa"], ws.dw, 0, h, w, 1024, S=S)
        ops.linear(ws.dw[:, :1024], Lw["w_proj_sa"], Lw["b_proj_sa"],
                   [(0, 256, ops.ACT_NONE, ops.OUT_F32_ACCUM, ws.tgt, 0),
                    (256, 512, ops.ACT_NONE, ops.OUT_F32_ACCUM, ws.tgt_id, 0)])
        if tr is not None:
            tr["layers"][-1].update(tgt=ws.tgt.cpu().clone(), tgt_id=ws.tgt_id.cpu().clone())

    # ------------------------------------------------------------------ device work (graph-capturable)
    def _read_body(self, img, img_embs, T, output_size, st):
        """Everything match_propogate_one_frame puts on the stream; a pure function of (T, output_size, st)."""
        self.ws.bind(st)
        we = self._w_events
        if we is not None and we["update_long"]:
            # the deferred write path appends to / evicts from the bank this frame reads, and stamps the slot with
            # the device frame counter: it has to be complete before the counter ticks
            torch.cuda.current_stream().wait_event(we["all"])
            we["waited"] = True
        ops.bank_tick(self.ws.bank_state)                               # frame_step += 1 on the device (every sequence)
        embs = self._encode_body(img) if img_embs is None else img_embs
        embs, pre = (embs[:4], embs[4:]) if len(embs) == 7 else (embs, None)
        lstt_out = self._lstt_forward(embs[-1], False, True, T)
        return self._decode(embs, lstt_out, output_size, adapters=pre)

    def _fused_body(self, label_hw, embs, Tw, update_long, do_evict, st_w, T, output_size, st):
        """The previous frame's write path as a parallel branch + this frame's read path."""
        if self._wside is None:
            self._wside = torch.cuda.Stream(device=self.device)
        start = torch.cuda.Event()
        start.record()
        self._wside.wait_event(start)
        we = dict(l0=torch.cuda.Event(), all=torch.cuda.Event(), update_long=update_long, waited=False)
        with torch.cuda.stream(self._wside):
            self._write_body(label_hw, Tw, update_long, do_evict, st_w, events=we)
            we["all"].record()
        self._w_events = we
        try:
            return self._read_body(None, embs, T, output_size, st)
        finally:
            self._w_events = None

    def _write_body(self, label_hw, T, update_long, do_evict, st, events=None):
        """Everything update_short_term_memory puts on the stream (transformer.py:826-991).  The frame's K and
        V already sit in the "current" set; fusing ID_V into it completes the new short-term memory, and the
        caller swaps the roles of the two sets afterwards."""
        ws = self.ws
        ws.bind(st)
        self._id_embed(label_hw)
        # the layers' ID_V GEMMs are independent and sub-wave: every second one goes to the side stream (a parallel
        # branch of the captured graph); the next frame's LSTT waits for this whole path
        forked = events is None and self.prep.num_layers > 1 and self._fork()
        for i in range(self.prep.num_layers):
            if events is not None and i == 1:
                events["l0"].record()                             # layer 0's short-term memory is complete
            with torch.cuda.stream(self._side if forked and i % 2 == 1 else torch.cuda.current_stream()):
                self._id_value(i, ws.vcur[i])
        if forked:
            self._join()
        if update_long:
            L = ws.L
            for s in range(ws.S):                # every sequence owns its bank, slot table and eviction history
                srcs, dsts = [], []
                for i in range(self.prep.num_layers):
                    srcs += [ws.kcur[i][s * L:(s + 1) * L], ws.vcur[i][s * L:(s + 1) * L]]
                    dsts += [ws.kbank[i][s], ws.vbank[i][s]]
                ops.bank_append(ws.bank_state[s], T, -1, srcs, dsts)    # frame id from the device counter
                if do_evict:
                    ops.foreground_prob(self.pred_id_logits[s], ws.fg[s], ws.h, ws.w)
                    ops.bank_score_evict(ws.bank_state[s], ws.mass[s * L * T:(s + 1) * L * T], ws.fg[s], L, T,
                                         self.capacity)

    def _decode(self, embs, lstt_out, output_size, adapters=None):
        """decode_current_logits body (aot_engine.py:438-465, models/deaot.py:57-63)."""
        if getattr(self, "_dec_gn_ws", None) is None:
            self._dec_gn_ws = torch.zeros(self.S, 17, device=self.device, dtype=torch.float64)
        dec = self.prep.decoder
        fused = self.fused_head and dec.w["conv_4x"][0].shape[0] in (64, 128) and dec.w_head.shape[0] <= 16
        conv, bias = dec(self._decoder_inputs(embs, lstt_out), embs, gn_ws=self._dec_gn_ws, adapters=adapters,
                         head=fused)
        ws = self.ws
        n, Co = conv.shape[0], dec.w_head.shape[0]
        shape = (n, Co) + tuple(conv.shape[-2:])
        if ws.pred_logits is None or tuple(ws.pred_logits.shape) != shape:
            ws.pred_logits = torch.empty(shape, device=conv.device, dtype=torch.float32)
        out = label = None
        if output_size is not None:
            out = torch.empty(n, Co, output_size[0], output_size[1], device=conv.device, dtype=torch.float32)
        n_valid = [int(k) + 1 for k in self.obj_nums]
        # + conv bias, channels beyond the objects -> -1e10 (aot_engine.py:455-457), persistent pred_id_logits
        # (read by the write path), bilinear resize to the output size (:459-463)
        if fused:
            # ... and, in the same passes, conv_4x's GroupNorm + ReLU and the 1x1 conv_out (the normalised map is never
            # stored) and the evaluator's argmax (pred_label)
            if out is not None:
                shape = (n,) + tuple(output_size)
                same_size = self.input_size_2d is not None and tuple(output_size) == tuple(self.input_size_2d)
                have = self._label_buf is not None and tuple(self._label_buf.shape) == shape
                if self.label_alias and same_size and (have or not torch.cuda.is_current_stream_capturing()):
                    # at the input resolution the label map is what update_short_term_memory is handed next: written
                    # straight into the buffer the deferred write path reads, so that call copies nothing (the buffer
                    # is never allocated inside a capture: the first propagated frame always runs eagerly)
                    label = self._label_buffer_shape(shape).view(n, 1, *output_size)
                else:
                    label = torch.empty(n, 1, output_size[0], output_size[1], device=conv.device, dtype=torch.float32)
            _, b4, g4, be4 = dec.w["conv_4x"]
            ops.logits_head(conv, b4, g4, be4, dec.w_head, dec.w["conv_out"][1], n_valid, ws.pred_logits, out, label,
                            self._dec_gn_ws)
        else:
            ops.finish_logits(conv, bias, ws.pred_logits, n_valid, out)
        self.pred_id_logits = ws.pred_logits
        res = ws.pred_logits if out is None else out
        res._rmem_label = label        # travels with the graph's output tensor: a replay does not re-run this method
        return res

    # ------------------------------------------------------------------ CUDA graphs
    def _wait_write(self):
        """Order the current stream after the last write path (it runs on its own stream, or is still deferred)."""
        if self._pending_w is not None:
            (T, update_long, do_evict, st), self._pending_w = self._pending_w, None
            self._run(("write", T, update_long, do_evict, st),
                      lambda m: self._write_body(m, T, update_long, do_evict, st), [self._label_buf])
        if self._w_done is not None:
            torch.cuda.current_stream().wait_event(self._w_done)

    def _capture(self, key, body, inputs, alias=False):
        """Capture ``body`` into a CUDA graph.  ``alias``: the inputs are outputs of another graph of this
        engine (stable addresses), so they are baked in as they are and never copied.  Write-path graphs use
        their own memory pool: they replay concurrently with the next frame's encoder graph."""
        static_in = list(inputs) if alias else [torch.empty(t.shape, dtype=t.dtype, device=self.device)
                                                 for t in inputs]
        is_write = key[0] == "write"
        if key[0] == "enc":                                       # may replay concurrently with the read graphs
            return self._capture_enc(key, body, static_in)
        pool = self._graph_pool_w if is_write else self._graph_pool
        launches0 = ops.LAUNCHES
        g = torch.cuda.CUDAGraph()
        torch.cuda.synchronize()
        # RMEM_HP_CAPTURE=1 records on a high-priority stream (kernel nodes inherit the capture stream's priority,
        # so the frame's own kernels win SMs over the concurrently running encoder of the next frame).  Measured
        # +1.5% device-resident but -1% host-to-host (graph launches cost 0.34 instead of 0.22 ms of host time): off.
        # (the capture stream is always this engine's own: torch's default capture stream is created once per process
        # on whichever device was current then, and would move an engine on another device onto that device)
        if self._hp_stream is None:
            self._hp_stream = torch.cuda.Stream(device=self.device,
                                                priority=-1 if os_env_flag("RMEM_HP_CAPTURE", False) else 0)
        with torch.cuda.graph(g, pool=pool, stream=self._hp_stream):
            out = body(*static_in)
        ent = (g, static_in, out, ops.LAUNCHES - launches0)
        ops.LAUNCHES = launches0
        self._graphs[key] = ent
        if is_write and self._graph_pool_w is None:
            self._graph_pool_w = g.pool()
        elif not is_write and self._graph_pool is None:
            self._graph_pool = g.pool()
        return ent

    def _capture_enc(self, key, body, static_in):
        launches0 = ops.LAUNCHES
        g = torch.cuda.CUDAGraph()
        torch.cuda.synchronize()
        # one memory pool per encoder graph: graphs sharing a pool reuse each other's scratch, and encoder graph
        # 0 replays while the outputs of encoder graph 1 are being read (and the other way round)
        if self._cap_stream is None:
            self._cap_stream = torch.cuda.Stream(device=self.device)      # see _capture: never the process-wide default
        with torch.cuda.graph(g, stream=self._cap_stream):
            out = body(*static_in)
        ent = (g, static_in, out, ops.LAUNCHES - launches0)
        ops.LAUNCHES = launches0
        self._graphs[key] = ent
        return ent

    def _run(self, key, body, inputs, alias=False):
        """Run ``body(*static_inputs)`` eagerly, or through a CUDA graph captured once per key."""
        if not self.use_cuda_graphs or self.frame_step < 2 or self.trace is not None:
            return body(*inputs)                                  # frame 1 is eager: cuDNN picks its algorithms
        ent = self._graphs.get(key)
        if ent is not None and alias and any(s.data_ptr() != t.data_ptr() for s, t in zip(ent[1], inputs)):
            return body(*inputs)                                  # not the tensors the graph was recorded on
        if ent is None:
            ent = self._capture(key, body, inputs, alias)
        g, static_in, out, n_kernels = ent
        if not alias:
            for s, t in zip(static_in, inputs):
                if t.is_cuda and t.is_floating_point() and t.is_contiguous():
                    # an SM kernel, not a copy-engine memcpy: the copy engines may be busy with a caller's
                    # upload of the next frame (pipeline.FrameStreamer) and this copy is on the critical path
                    torch.mul(t, 1, out=s)
                else:
                    s.copy_(t, non_blocking=True)
        g.replay()
        ops.LAUNCHES += n_kernels
        return out

    def _launch_encoder(self, img):
        """Encode ``img`` -- one frame, or a stack of the next B frames of this sequence in ONE batched pass (the cuDNN
        trunk is latency-bound at one 480p frame: 293 us for one frame, 198 / 179 / 170 us per frame at B = 2 / 3 / 4)
        -- on the encoder stream, ordered after the point the current stream has reached.  Two encoder graphs with
        their own output sets alternate: the set written now was last read by the frame before the current one, which
        the current stream has already passed, so nothing is ever copied.  Frame j of the stack is the batch slice
        [j:j+1] of every output: B read-graph variants per set, recorded once."""
        if self._estream is None:
            self._estream = torch.cuda.Stream(device=self.device)
        es = self._estream
        ready = torch.cuda.Event()
        ready.record()
        es.wait_event(ready)
        if img.is_cuda:
            img.record_stream(es)
        B = img.shape[0] // self.S            # lock-step engines: frame-major stack [B * S, ...], frame j = rows [j S, (j + 1) S)
        p = self._enc_set
        self._enc_set = 1 - p
        with torch.cuda.stream(es):
            embs = list(self._run(("enc", p, B), self._encode_body, [img]))
            self._enc_done = torch.cuda.Event()
            self._enc_done.record()
        for j in range(B):
            self._pq.append(([e[j * self.S:(j + 1) * self.S] for e in embs] if B > 1 else embs, (p, B, j),
                             img[j * self.S:(j + 1) * self.S] if self.check_prefetch else None))

    def _take_prefetched(self, img=None):
        """The oldest frame encoded ahead and the (set, batch, index) of its embeddings; the current stream waits for
        the encoder pass that produced them."""
        cur = torch.cuda.current_stream()
        cur.wait_event(self._enc_done)
        embs, p, frame = self._pq.popleft()
        if frame is not None and img is not None and not torch.equal(frame, img.to(frame.device)):
            # RMEM_CHECK_PREFETCH=1 (debugging aid; it synchronises): the frame handed over now is not the one that was
            # encoded ahead for this call
            raise _lib.RmemError("match_propogate_one_frame: this frame differs from the one passed as next_img")
        for e in embs:
            e.record_stream(cur)
        return embs, p

    def frames_ahead(self):
        """How many frames have been encoded ahead (``next_img``) and not yet consumed."""
        return len(self._pq)

    @_on_device
    def precapture(self, output_size=None, prefetch=False, lookahead=1):
        """Capture the read / write graphs of every bank length now (capturing records, it does not
        execute, so the sequence state is untouched). Call after the first propagated frame.  ``prefetch``: record
        the read graphs on the embedding sets used by ``match_propogate_one_frame(..., next_img=)``; ``lookahead``: the
        number of frames a ``next_img`` stack will hold."""
        if not self.use_cuda_graphs or self.frame_step < 1:
            return 0
        H, W = self.input_size_2d
        # the encoder graphs are recorded for the frame format this sequence uses (normalised f32 planes
        # [1, 3, H, W] or decoded u8 pixels [1, H, W, 3]); one format per sequence
        if self._img_sig is not None and self._img_sig[0] == torch.uint8:
            img = torch.zeros(self.S, H, W, 3, device=self.device, dtype=torch.uint8)
        else:
            img = torch.zeros(self.S, 3, H, W, device=self.device)
        label = torch.zeros(self.S, H, W, device=self.device)
        saved = dict(self._graphs)
        self._wait_write()
        sets = {}
        for p in ((0, 1) if prefetch else (0,)):                  # prefetching alternates between two sets
            for B in sorted({1, lookahead} if prefetch else {1}):
                if ("enc", p, B) not in self._graphs:
                    self._capture(("enc", p, B), self._encode_body, [img if B == 1 else torch.cat([img] * B)])
                outs = list(self._graphs[("enc", p, B)][2])       # the encoder graphs' outputs feed the read graphs
                for j in range(B):
                    sets[(p, B, j)] = [e[j * self.S:(j + 1) * self.S] for e in outs] if B > 1 else outs
        for T in range(1, self.capacity + 1):
            for key, body, inputs, alias in self._variants(T, output_size, sets, label):
                if key not in self._graphs:
                    self._capture(key, body, inputs, alias)
        return len(self._graphs) - len(saved)

    def _label_buffer(self, like):
        return self._label_buffer_shape(tuple(like.shape))

    def _label_buffer_shape(self, shape):
        if self._label_buf is None or tuple(self._label_buf.shape) != tuple(shape):
            self._label_buf = torch.empty(shape, dtype=torch.float32, device=self.device)
        return self._label_buf

    def _variants(self, T, output_size, sets, label):
        ev = self.deaot or T + 1 > self.capacity
        if self.lazy_write and len(sets) > 1:
            # steady state of the deferred schedule (prefetching loops): write path of the previous frame (bank length T, with or
            # without a long-term update) + read path of this frame; the stand-alone graphs are recorded on demand
            buf = self._label_buffer(label)
            for st in (0, 1):
                for p, embs in sets.items():
                    for ul, de, Tr in ((False, False, T), (True, ev, min(T + 1, self.capacity))):
                        yield (("wr", T, ul, de, st, Tr, output_size, 1 - st, p),
                               lambda m, *e, T=T, ul=ul, de=de, st=st, Tr=Tr: self._fused_body(
                                   m, list(e), T, ul, de, st, Tr, output_size, 1 - st),
                               [buf] + embs, True)
            return
        for st in ((0, 1) if self.deaot else (0,)):
            for p, embs in sets.items():
                yield (("read", T, output_size, st, p),
                       lambda *e, T=T, st=st: self._read_body(None, list(e), T, output_size, st), embs, True)
            yield (("write", T, False, False, st), lambda m, T=T, st=st: self._write_body(m, T, False, False, st),
                   [label], False)
            yield (("write", T, True, ev, st), lambda m, T=T, ev=ev, st=st: self._write_body(m, T, True, ev, st),
                   [label], False)

    # ------------------------------------------------------------------ public frame API
    @_on_device
    def add_reference_frame(self, img=None, mask=None, frame_step=-1, obj_nums=None, img_embs=None):
        """aot_engine.py:241-325 (always eager: once per sequence)."""
        if self.obj_nums is None and obj_nums is None:
            raise ValueError("No objects for reference frame!")
        self._wait_write()
        if self._enc_done is not None:                           # an encoder pass left over from another sequence
            torch.cuda.current_stream().wait_event(self._enc_done)
            self._pq.clear()
        if obj_nums is not None:
            self.obj_nums = obj_nums
        if frame_step == -1:
            frame_step = self.frame_step
        if img_embs is None and img is None:
            raise ValueError("No image for reference frame!")
        if mask is None:
            raise ValueError("No mask for reference frame!")
        if img is not None:
            self._img_sig = (img.dtype, tuple(img.shape))
        # an engine that is reused for another sequence (restart_engine keeps the workspace and the captured graphs)
        # drops graphs that bake what has changed: the valid-class counts of the logits kernels, the frame format
        baked = (tuple(int(n) for n in self.obj_nums), self._img_sig)
        if self._graphs and self._graphs_baked != baked:
            self._graphs, self._graph_pool, self._graph_pool_w = {}, None, None
            self._enc_set = 0
        self._graphs_baked = baked
        curr_enc_embs = self.encode_image(img) if img_embs is None else img_embs
        label = self._label_map(mask.to(self.device)).float()
        if self.input_size_2d is None:
            # (a decoded u8 frame is [1, H, W, 3]; normalised f32 planes are [1, 3, H, W])
            in_size = (img.shape[-3:-1] if img.dtype == torch.uint8 else img.shape[2:]) if img is not None \
                else mask.shape[-2:]
            self.update_size(in_size, curr_enc_embs[-1].shape[2:])
        if len(self.obj_nums) != self.S:
            raise ValueError("obj_nums must list one object count per sequence (%d)" % self.S)
        self._id_embed(label)
        for s in range(self.S):
            ops.bank_init(self.ws.bank_state[s], self.frame_step)
        self.T = 1
        self.st = 0
        self.ws.bind(0)
        lstt_out = self._lstt_forward(curr_enc_embs[-1], True, False, 1)
        self.last_mem_step = frame_step
        self._decode(curr_enc_embs, lstt_out, None)

    @_on_device
    def match_propogate_one_frame(self, img=None, img_embs=None, mask=None, output_size=None, next_img=None):
        """aot_engine.py:398-436 (save_atten_weights=True in eval). With CUDA graphs the returned logits
        live in a buffer that the next call overwrites -- consume them (softmax / argmax) first, as the
        reference's evaluator does.

        ``next_img`` (optional, not in the reference; its closest relative there is ``offline_encoder``,
        aot_engine.py:174-196, which encodes a whole clip in one batch): the frame that will be passed to the NEXT call,
        or a stack of the frames of the next B calls.  Their encoder pass -- one batched pass -- starts now on a second
        stream and overlaps this frame's LSTT + decoder (the encoder depends on nothing but the images); the next B
        calls must then be given those same frames, and the next stack comes with the call that consumes the last."""
        if img is not None and self._img_sig is not None and (img.dtype, tuple(img.shape)) != self._img_sig:
            raise _lib.RmemError("frames of one sequence must share the reference frame's dtype and shape: got %s %s, "
                                 "the sequence started with %s %s" % (img.dtype, tuple(img.shape), *self._img_sig))
        self.frame_step += 1
        T = self.T
        self.T_recorded = T
        osz = tuple(output_size) if output_size is not None else None
        st = self.st
        if img_embs is not None:
            self._wait_write()
            if getattr(img_embs, "_rmem_static", False):
                # embeddings that live at stable addresses (another engine's encoder graph: the object groups of one
                # DeAOTInferEngine share the backbone pass): the read path replays as a graph on them
                return self._logits_out(self._run(("read", T, osz, st, "ext"),
                                                  lambda *e: self._read_body(None, _EmbList(e), T, osz, st),
                                                  list(img_embs), alias=True))
            return self._logits_out(self._read_body(None, img_embs, T, osz, st))
        # two graphs per frame: the encoder (independent of the previous frame's write path, which may still be
        # running on its own stream) and, after waiting for that write path, LSTT + decoder
        if next_img is not None or self._pq:
            if not self._pq:
                self._launch_encoder(img)                 # first frame of a prefetching loop: nothing encoded ahead
            embs, p = self._take_prefetched(img)
            if next_img is not None:
                if self._pq:
                    raise _lib.RmemError("next_img given while %d frame(s) encoded ahead are still waiting: pass the "
                                         "next stack with the call that consumes the last one" % len(self._pq))
                self._launch_encoder(next_img)
        else:
            embs, p = self._run(("enc", 0, 1), self._encode_body, [img]), (0, 1, 0)
        self._prefetching = bool(self._pq)                # the next call finds its embeddings ready
        if self._pending_w is not None:
            pend, self._pending_w = self._pending_w, None
            return self._logits_out(self._run(("wr",) + pend + (T, osz, st, p),
                                              lambda m, *e: self._fused_body(m, list(e), *pend, T, osz, st),
                                              [self._label_buf] + list(embs), alias=True))
        self._wait_write()
        return self._logits_out(self._run(("read", T, osz, st, p),
                                          lambda *e: self._read_body(None, list(e), T, osz, st), list(embs), alias=True))

    def _logits_out(self, logits):
        """``pred_label``: the argmax of these logits (f32 [S, 1, Ho, Wo]) when the decoder tail computed it."""
        self.pred_label = getattr(logits, "_rmem_label", None)
        return logits

    @_on_device
    def decode_current_logits(self, curr_enc_embs, curr_lstt_embs, output_size=None):
        emb = curr_lstt_embs[-1] if isinstance(curr_lstt_embs, (list, tuple)) else curr_lstt_embs
        return self._decode(curr_enc_embs, emb, output_size)

    @_on_device
    def update_short_term_memory(self, curr_mask, curr_id_emb=None, step=0):
        """aot_engine.py:327-369."""
        if curr_id_emb is not None:
            raise NotImplementedError("pre-computed id embeddings are a training-only path")
        label = self._label_map(curr_mask.to(self.device), allow_soft=True).float()
        soft = label.dim() == 4
        update_long = False
        if self.frame_step - self.last_mem_step >= self.long_term_mem_gap:
            update_long = True
            self.last_mem_step = self.frame_step
        T = self.T
        do_evict = update_long and (self.deaot or T + 1 > self.capacity)
        if do_evict and self.T_recorded != T:
            raise _lib.RmemError("record_attn_weight is stale: call match_propogate_one_frame first")
        st = self.st
        body = lambda m: self._write_body(m, T, update_long, do_evict, st)
        if soft:
            # rare form: its own graph key (the static input has another shape), on the calling stream
            self._wait_write()
            self._run(("write", T, update_long, do_evict, st, tuple(label.shape)), body, [label])
        elif self.lazy_write and self._prefetching:
            # (a loop that does not prefetch keeps the write path on its own stream: there it overlaps the next
            # frame's encoder, which is worth more than the two graph boundaries -- 1055 vs 1017 frames/s)
            self._wait_write()                                    # an earlier deferred write path, if any
            buf = self._label_buffer(label)
            if label.data_ptr() != buf.data_ptr():                # pred_label at the input size already lives there
                torch.mul(label, 1, out=buf)                      # SM copy (see _run)
            self._pending_w = (T, update_long, do_evict, st)
        elif self.async_write:
            if self._wstream is None:
                self._wstream = torch.cuda.Stream(device=self.device)
            ready = torch.cuda.Event()
            ready.record()                                        # the label (and this frame's read path) is done
            self._wstream.wait_event(ready)
            label.record_stream(self._wstream)
            with torch.cuda.stream(self._wstream):
                self._run(("write", T, update_long, do_evict, st), body, [label])
                self._w_done = torch.cuda.Event()
                self._w_done.record()
        else:
            self._run(("write", T, update_long, do_evict, st), body, [label])
        if self.deaot:
            self.st = 1 - st                                      # the set just completed is the short-term memory now
        if update_long:
            self.T = min(T + 1, self.capacity)
            self.n_long_updates += 1

    def eval(self):
        return self


from .engine_aot import AOTBodies, WorkspaceAOT, mha_chunks_per_frame  # noqa: E402


class AOTEngine(AOTBodies, DeAOTEngine):
    """AOTEngine (aot_engine.py:18-568) on the multi-head kernels: the state machine of DeAOTEngine with the
    AOT block (engine_aot.AOTBodies).  Eviction statistics only run once the bank is full (transformer.py:332-334)."""

    _family_deaot = False

    def _new_workspace(self, h, w, n_slots):
        return WorkspaceAOT(self.prep, h, w, n_slots, self.device, mha_chunks_per_frame)


class DeAOTInferEngine:
    """AOTInferEngine / DeAOTInferEngine (aot_engine.py:571-725, deaot_engine.py:20-56): splits more
    than ``max_aot_obj_num`` objects over several engines and merges their logits every frame."""

    def __init__(self, aot_model, gpu_id=0, long_term_mem_gap=9999, short_term_mem_skip=1,
                 max_aot_obj_num=None):
        self.cfg = aot_model.cfg
        self.AOT = aot_model
        if max_aot_obj_num is None or max_aot_obj_num > aot_model.max_obj_num:
            self.max_aot_obj_num = aot_model.max_obj_num
        else:
            self.max_aot_obj_num = max_aot_obj_num
        self.gpu_id = gpu_id
        self.long_term_mem_gap = long_term_mem_gap
        self.short_term_mem_skip = short_term_mem_skip
        self.aot_engines = []
        self._engine_cls = DeAOTEngine if is_deaot_model(aot_model) else AOTEngine
        self.restart_engine()

    def eval(self):
        return self

    def restart_engine(self):
        """aot_engine.py:688-693 drops the engines; here they go back to a pool: an engine keeps its workspace (~70 MB
        at 480p) and its captured CUDA graphs, and the next sequence of the same frame size reuses both (a sequence of
        another size re-allocates in update_size, other object counts / frame formats re-capture)."""
        for engine in self.aot_engines:
            engine.restart_engine()
        self._pool = self.aot_engines + getattr(self, "_pool", [])
        self.aot_engines = []
        self.obj_nums = None

    def separate_mask(self, mask):
        # aot_engine.py:604-628
        if mask is None:
            return [None] * len(self.aot_engines)
        if len(self.aot_engines) == 1:
            return [mask]
        if mask.dim() == 3 or mask.shape[0] == 1:
            out = []
            for idx in range(len(self.aot_engines)):
                start_id = idx * self.max_aot_obj_num + 1
                end_id = (idx + 1) * self.max_aot_obj_num
                fg = ((mask >= start_id) & (mask <= end_id)).float()
                out.append((fg * mask - start_id + 1) * fg)
            return out
        out = []
        for idx in range(len(self.aot_engines)):
            start_id = idx * self.max_aot_obj_num + 1
            end_id = (idx + 1) * self.max_aot_obj_num
            fg_prob = mask[start_id:(end_id + 1)]
            bg_prob = 1. - torch.sum(fg_prob, dim=1, keepdim=True)
            out.append(torch.cat([bg_prob, fg_prob], dim=1))
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
        merged = torch.cat([bg_prob] + fg_probs, dim=1).clamp(1e-5, 1 - 1e-5)
        return torch.logit(merged)

    def add_reference_frame(self, img, mask, obj_nums, frame_step=-1):
        if isinstance(obj_nums, list):
            obj_nums = obj_nums[0]
        self.obj_nums = obj_nums
        aot_num = max(-(-int(obj_nums) // self.max_aot_obj_num), 1)
        while aot_num > len(self.aot_engines):
            self.aot_engines.append(self._pool.pop(0) if self._pool else
                                    self._engine_cls(self.AOT, self.gpu_id, self.long_term_mem_gap,
                                                     self.short_term_mem_skip))
        separated = self.separate_mask(mask)
        img_embs = None
        for engine, m in zip(self.aot_engines, separated):
            engine.long_term_mem_gap = self.long_term_mem_gap
            if img_embs is None:
                img_embs = engine.encode_image(img)        # the backbone runs once per frame
            engine.add_reference_frame(img, m, obj_nums=[self.max_aot_obj_num], frame_step=frame_step,
                                       img_embs=img_embs)
        self.update_size()

    def match_propogate_one_frame(self, img=None, mask=None, output_size=None, next_img=None):
        if len(self.aot_engines) == 1:                      # <= 10 objects: one engine, CUDA graphs end to end
            eng = self.aot_engines[0]
            eng.long_term_mem_gap = self.long_term_mem_gap
            return eng.match_propogate_one_frame(img, mask=mask, output_size=output_size, next_img=next_img)
        # more than 10 objects: the backbone runs once (first engine's encoder graph: outputs at stable addresses),
        # every object group replays its own read graph on those embeddings, the logits are merged (aot_engine.py:695-712)
        img_embs = self.aot_engines[0].encode_static(img)
        all_logits = []
        for engine in self.aot_engines:
            engine.long_term_mem_gap = self.long_term_mem_gap
            all_logits.append(engine.match_propogate_one_frame(img, img_embs=img_embs, mask=mask,
                                                               output_size=output_size))
        return self.soft_logit_aggregation(all_logits)

    def update_memory(self, curr_mask):
        for engine, m in zip(self.aot_engines, self.separate_mask(curr_mask)):
            engine.update_short_term_memory(m)

    def precapture(self, output_size=None, prefetch=False, lookahead=1):
        return sum(e.precapture(output_size, prefetch, lookahead) or 0 for e in self.aot_engines[:1])

    def frames_ahead(self):
        return min(e.frames_ahead() for e in self.aot_engines) if self.aot_engines else 0

    @property
    def pred_label(self):
        """argmax of the logits the last match_propogate_one_frame returned, computed inside the decoder tail (one
        engine, output_size given); None when several engines' logits were merged (more than 10 objects)."""
        return self.aot_engines[0].pred_label if len(self.aot_engines) == 1 else None

    def update_size(self):
        self.input_size_2d = self.aot_engines[0].input_size_2d
        self.enc_size_2d = self.aot_engines[0].enc_size_2d
        self.enc_hw = self.aot_engines[0].enc_hw


AOTInferEngine = DeAOTInferEngine


def build_engine(name, phase="train", **kwargs):
    """networks/engines/__init__.py:5-21."""
    if name in ("aotengine", "deaotengine"):
        if phase == "train":
            cls = DeAOTEngine if is_deaot_model(kwargs["aot_model"]) else AOTEngine
            return cls(**kwargs)
        if phase == "eval":
            return DeAOTInferEngine(**kwargs)
    raise NotImplementedError(name)
