"""AOT (multi-head, 8 x 32) half of the drop-in engines: weight repack, per-sequence workspace and the
LSTT forward / memory-write bodies of ``AOTEngine`` on the sm_100a kernels (csrc/mha.cu, csrc/aot_glue.cu,
csrc/gemm.cu, csrc/fused.cu).  Reference: SimplifiedTransformerBlock.forward
(aot_plus/networks/layers/transformer.py:553-692), LongShortTermTransformer.forward / update_short_memories
(:199-304), AOTEngine (networks/engines/aot_engine.py:18-568).  The state machine itself is shared with
DeAOTEngine (engine.py); this module only supplies the AOT-specific device work.
"""
import math

import torch

from . import _lib, ops

BF16 = torch.bfloat16


def _bf(t, dev):
    return t.detach().to(device=dev, dtype=BF16).contiguous()


def _f32(t, dev):
    return t.detach().to(device=dev, dtype=torch.float32).contiguous()


def position_embedding_sine(h, w, num_pos_feats=128, temperature=10000.0):
    """PositionEmbeddingSine(normalize=True, scale=2*pi) of the reference (networks/layers/position.py:35-79,
    constructed models/aot.py:87-90) as an [h*w, 2*num_pos_feats] f32 token matrix (host, once per size)."""
    gy, gx = torch.meshgrid(torch.arange(h), torch.arange(w), indexing="ij")
    y_embed, x_embed = gy.unsqueeze(0).float(), gx.unsqueeze(0).float()
    eps, scale = 1e-6, 2 * math.pi
    y_embed = y_embed / (y_embed[:, -1:, :] + eps) * scale
    x_embed = x_embed / (x_embed[:, :, -1:] + eps) * scale
    dim_t = torch.arange(num_pos_feats, dtype=torch.float32)
    dim_t = temperature ** (2 * torch.div(dim_t, 2, rounding_mode="trunc") / num_pos_feats)
    pos_x = x_embed[:, :, :, None] / dim_t
    pos_y = y_embed[:, :, :, None] / dim_t
    pos_x = torch.stack((pos_x[:, :, :, 0::2].sin(), pos_x[:, :, :, 1::2].cos()), dim=4).flatten(3)
    pos_y = torch.stack((pos_y[:, :, :, 0::2].sin(), pos_y[:, :, :, 1::2].cos()), dim=4).flatten(3)
    return torch.cat((pos_y, pos_x), dim=3).reshape(h * w, 2 * num_pos_feats).contiguous()


class PreparedAOT:
    """Device-side repack of an AOT model's weights (once per model per device)."""

    deaot = False

    def __init__(self, model, device, encoder_cls, decoder_cls, enc_dtype):
        sd = model.state_dict()
        self.device = device
        self.cfg = model.cfg
        self.num_layers = sum(1 for k in sd if k.startswith("LSTT.layers.") and k.endswith(".norm1.weight"))
        self.layers = []
        z = torch.zeros(256, 256)
        for i in range(self.num_layers):
            p = "LSTT.layers.%d." % i
            g = lambda n: sd[p + n].detach().float().cpu()
            nb = lambda n: (_f32(g(n + ".weight"), device), _f32(g(n + ".bias"), device))
            L = {n: nb(n) for n in ("norm1", "norm2", "norm3", "norm4")}
            # self-attention input GEMM over A = [norm1(x)+pos | norm1(x)]: Q, K from the first half, V from the second
            w_qkv = torch.cat([torch.cat([g("self_attn.linear_Q.weight"), z], 1),
                               torch.cat([g("self_attn.linear_K.weight"), z], 1),
                               torch.cat([z, g("self_attn.linear_V.weight")], 1)], 0)
            L["w_sa_qkv"] = _bf(w_qkv, device)
            L["b_sa_qkv"] = _f32(torch.cat([g("self_attn.linear_Q.bias"), g("self_attn.linear_K.bias"),
                                            g("self_attn.linear_V.bias")]), device)
            L["w_sa_proj"] = _bf(g("self_attn.projection.weight"), device)
            L["b_sa_proj"] = _f32(g("self_attn.projection.bias"), device)
            L["w_q"], L["b_q"] = _bf(g("linear_Q.weight"), device), _f32(g("linear_Q.bias"), device)
            # linear(x + id_emb) = [x | id_emb] @ [W | W]^T : the add happens in the fp32 accumulator
            L["w_v2"] = _bf(torch.cat([g("linear_V.weight")] * 2, 1), device)
            L["b_v"] = _f32(g("linear_V.bias"), device)
            L["w_vmem2"] = _bf(torch.cat([g("linear_VMem.weight")] * 2, 1), device)
            L["b_vmem"] = _f32(g("linear_VMem.bias"), device)
            L["w_qmem"], L["b_qmem"] = _bf(g("linear_QMem.weight"), device), _f32(g("linear_QMem.bias"), device)
            # [lt_out | st_out] @ W^T: columns 0..255 = tgt2 + tgt3 (residual), columns 256..511 = tgt3 alone
            wl, wst = g("long_term_attn.projection.weight"), g("short_term_attn.projection.weight")
            L["w_proj_ls"] = _bf(torch.cat([torch.cat([wl, wst], 1), torch.cat([z, wst], 1)], 0), device)
            L["b_proj_ls"] = _f32(torch.cat([g("long_term_attn.projection.bias") + g("short_term_attn.projection.bias"),
                                             g("short_term_attn.projection.bias")]), device)
            L["w_ff1"], L["b_ff1"] = _bf(g("linear1.weight"), device), _f32(g("linear1.bias"), device)
            L["gn"] = nb("activation.gn")
            L["dw"] = _f32(g("activation.conv.weight").reshape(-1, 25).t(), device)        # tap-major [25][1024]
            L["w_ff2"], L["b_ff2"] = _bf(g("linear2.weight"), device), _f32(g("linear2.bias"), device)
            L["dec_norm"] = (_f32(sd["LSTT.decoder_norms.%d.weight" % i], device),
                             _f32(sd["LSTT.decoder_norms.%d.bias" % i], device))
            self.layers.append(L)
        self.ff_dim = self.layers[0]["w_ff1"].shape[0]
        self.cur_pos_emb = _f32(sd["cur_pos_emb"].reshape(-1), device)
        self.mem_pos_emb = _f32(sd["mem_pos_emb"], device)
        wid = sd["patch_wise_id_bank.weight"].detach().float().cpu()            # [256, n_cls, 17, 17]
        self.n_id_cls = wid.shape[1]
        self._id_w, self._id_taps = wid, None
        self.id_prefix = ops.id_prefix_table(wid).to(device)
        self.id_bias = _f32(sd["patch_wise_id_bank.bias"], device)
        self.id_norm = (None, None)                                             # AOT: no LayerNorm on the id embedding
        self.encoder = encoder_cls(model.encoder, model.encoder_projector, device, enc_dtype)
        self.decoder = decoder_cls(sd, device, intermediate=True, shortcut_dtype=self.encoder.dtype)

    def id_taps(self):
        """patch_wise_id_bank.weight as [n_cls, 17, 17, 256] fp32 on the device (soft masks), built on first use."""
        if self._id_taps is None:
            self._id_taps = ops.id_tap_table(self._id_w).to(self.device)
        return self._id_taps


class WorkspaceAOT:
    """Per-sequence device buffers for one token grid (h, w); allocated once, reused every frame."""

    def __init__(self, prep, h, w, n_slots, device, chunks_per_frame_fn):
        L, nl = h * w, prep.num_layers
        f32 = torch.float32

        def z(*s, dt=BF16):
            return torch.zeros(*s, device=device, dtype=dt)

        self.h, self.w, self.L, self.n_slots = h, w, L, n_slots
        self.tgt = z(L, 256, dt=f32)
        self.pos = position_embedding_sine(h, w).to(device)        # self_pos [L, 256] f32
        self.xn = z(L, 512)                                        # [norm1(x) + pos | norm1(x)]; norm3 reuses cols 0..255
        self.sa_qkv = z(L, 768)
        self.attn_out = z(L, 512)                                  # [long-term (or self) | short-term] attention outputs
        self.mem = [z(L, 768) for _ in range(nl)]                  # [curr_V | id_emb | tgt3]
        self.kcur = [z(L, 256) for _ in range(nl)]                 # curr_Q = curr_K
        self.vcur = [z(L, 256) for _ in range(nl)]                 # linear_V(curr_V + id_emb): the bank candidate
        self.knew = [z(L, 256) for _ in range(nl)]                 # linear_QMem(tgt3)
        self.kprev = [z(L, 256) for _ in range(nl)]                # short-term memory (previous frame)
        self.vprev = [z(L, 256) for _ in range(nl)]
        self.kbank = [z(n_slots, L, 256) for _ in range(nl)]       # long-term bank
        self.vbank = [z(n_slots, L, 256) for _ in range(nl)]
        self.qt, self.tpe_bias = z(L, 256), z(L, 8, 4, dt=f32)
        self.st_kv = z(L, 512)                                     # norm4(local_K + curr_K) | norm4(local_V + curr_V)
        self.ff = z(L, prep.ff_dim, dt=f32)
        self.ffa, self.ffd = z(L, prep.ff_dim), z(L, prep.ff_dim)
        n_chunks = max(T * chunks_per_frame_fn(T, L) for T in range(1, n_slots + 1))
        self.part_o = z(n_chunks, L, 256, dt=f32)
        self.part_m, self.part_l = z(n_chunks, 8, L, dt=f32), z(n_chunks, 8, L, dt=f32)
        self.mass = z(L * _lib.MAX_BANK, dt=f32)                   # layer-0 record_attn_weight [L, T]
        self.dec_in = z(L, 256 * (nl + 1), dt=f32)                 # channels-last decoder input [emb | out_0 | out_1 | out_2]
        self.gn_ws = z(65, dt=torch.float64)                  # 2 x 32 group sums + arrival counter
        self.fg = z(L, dt=f32)
        self.S = 1                                                  # the AOT engine runs one sequence per engine
        self.bank_state = z(1, _lib.BANK_STATE_WORDS, dt=torch.int32)   # [sequence][words], as the DeAOT workspace
        self.zero_order = z(1, dt=torch.int32)
        self.pred_logits = None

    def bind(self, st):
        pass


def mha_chunks_per_frame(T, L, n_sms=148):
    """Split-KV granularity of the multi-head attention grid (q tiles x 4 head pairs x T x cpf CTAs, one CTA
    per SM): the cpf in 1..4 that minimises waves x (64-key tiles per chunk + fixed cost)."""
    q_tiles = -(-L // 128)
    best, best_cost = 1, None
    for cpf in range(1, _lib.MAX_CHUNKS_PER_FRAME + 1):
        ctas = q_tiles * 4 * T * cpf
        tiles = -(-(-(-L // cpf)) // 64)
        cost = -(-ctas // n_sms) * (tiles + 3.0)
        if best_cost is None or cost < best_cost:
            best, best_cost = cpf, cost
    return best


class AOTBodies:
    """Device work of the AOT engine (mixed into engine.AOTEngine): LSTT forward, memory write, id embedding."""

    def _mha(self, q, k, v, order, T, tpe, out_col0, mass):
        ws = self.ws
        cpf = mha_chunks_per_frame(T, ws.L)
        ops.mha_partial(q, k, v, order, T, cpf, ws.part_o, ws.part_m, ws.part_l)
        ops.mha_combine(ws.part_o, ws.part_m, ws.part_l, ws.tpe_bias if tpe else None,
                        self._tpe_rows(T) if tpe else None, T, ws.L, cpf, ws.attn_out, out_col0, mass)

    def _id_embed(self, label_hw):
        """assign_identity + get_id_emb (aot_engine.py:208-232, models/aot.py:111-114): the gather kernel writes
        id_emb next to curr_V / tgt3 of every layer (the A operand of linear_V / linear_VMem)."""
        ws, prep = self.ws, self.prep
        dsts = [(ws.mem[i], 256) for i in range(prep.num_layers)]
        if label_hw.dim() == 4:                 # one-hot / probability masks [1, n_ch, H, W]
            ops.id_embed_soft(label_hw, None, prep.id_taps(), prep.id_bias, None, None, prep.n_id_cls, dsts, ws.h, ws.w)
            return
        ops.id_embed(label_hw[0], prep.id_prefix, prep.id_bias, None, None, prep.n_id_cls, dsts, ws.h, ws.w)

    def _lstt_forward(self, curr_emb16, is_reference, save_atten_weights, T):
        """LongShortTermTransformer.forward + SimplifiedTransformerBlock.forward (transformer.py:199-267, 553-692)."""
        ws, prep = self.ws, self.prep
        h, w, L = ws.h, ws.w, ws.L
        emb = self._tokens(curr_emb16)
        ws.tgt.copy_(emb)
        ws.dec_in[:, :256].copy_(emb)
        for i in range(prep.num_layers):
            self._block_forward(i, is_reference, save_atten_weights and i == 0, T)
        return ws.dec_in

    def _block_forward(self, i, is_reference, record, T):
        """One SimplifiedTransformerBlock (transformer.py:553-692) on ws.tgt, in place; layer i's memories."""
        ws, prep = self.ws, self.prep
        h, w = ws.h, ws.w
        tr = self.trace
        A, BF, F32, ACC = ops.ACT_NONE, ops.OUT_BF16, ops.OUT_F32, ops.OUT_F32_ACCUM
        Lw = prep.layers[i]
        mem = ws.mem[i]
        # self-attention: q = k = norm1(x) + pos, v = norm1(x)
        ops.layernorm_ex(ws.tgt, *Lw["norm1"], out=ws.xn, col_out=256, post_add=ws.pos, out_plus=ws.xn,
                         col_plus=0)
        ops.linear(ws.xn, Lw["w_sa_qkv"], Lw["b_sa_qkv"], [(0, 768, A, BF, ws.sa_qkv, 0)])
        self._mha(ws.sa_qkv[:, :256], ws.sa_qkv[:, 256:512], ws.sa_qkv[:, 512:], ws.zero_order, 1, False, 0,
                  None)
        ops.linear(ws.attn_out[:, :256], Lw["w_sa_proj"], Lw["b_sa_proj"], [(0, 256, A, ACC, ws.tgt, 0)])
        # curr_V = norm2(x); curr_Q = curr_K = linear_Q(curr_V)
        ops.layernorm_ex(ws.tgt, *Lw["norm2"], out=mem, col_out=0)
        ops.linear(mem[:, :256], Lw["w_q"], Lw["b_q"], [(0, 256, A, BF, ws.kcur[i], 0)])
        if is_reference:
            # the reference frame is its own long- and short-term memory (transformer.py:583-592)
            ops.linear(mem[:, :512], Lw["w_v2"], Lw["b_v"], [(0, 256, A, BF, ws.vcur[i], 0)])
            ws.kprev[i].copy_(ws.kcur[i]); ws.vprev[i].copy_(ws.vcur[i])
            ws.kbank[i][0].copy_(ws.kcur[i]); ws.vbank[i][0].copy_(ws.vcur[i])
        # long-term attention over the bank (TPE as a per-(head, frame) score bias)
        ops.tpe_prepare_mha(ws.kcur[i], prep.cur_pos_emb, prep.mem_pos_emb, ws.qt, ws.tpe_bias)
        mass = ws.mass if record else None
        self._mha(ws.qt, ws.kbank[i], ws.vbank[i], ws.bank_state[0], T, True, 0, mass)
        # short-term attention over norm4(previous frame's K / V + current K / V)
        ops.layernorm_ex(ws.kprev[i], *Lw["norm4"], xb=ws.kcur[i], out=ws.st_kv, col_out=0)
        ops.layernorm_ex(ws.vprev[i], *Lw["norm4"], xb=mem[:, :256], out=ws.st_kv, col_out=256)
        self._mha(ws.kcur[i], ws.st_kv[:, :256], ws.st_kv[:, 256:], ws.zero_order, 1, False, 256, None)
        # x += tgt2 + tgt3 (both projections in one GEMM); tgt3 also kept for the short-term memory
        ops.linear(ws.attn_out, Lw["w_proj_ls"], Lw["b_proj_ls"],
                   [(0, 256, A, ACC, ws.tgt, 0), (256, 512, A, BF, mem, 512)])
        ops.linear(mem[:, 512:], Lw["w_qmem"], Lw["b_qmem"], [(0, 256, A, BF, ws.knew[i], 0)])
        if is_reference:
            ws.kprev[i].copy_(ws.knew[i])
            ops.linear(mem[:, 256:], Lw["w_vmem2"], Lw["b_vmem"], [(0, 256, A, BF, ws.vprev[i], 0)])
        # feed-forward: linear1 -> GroupNorm(32) + GELU -> dw5x5 -> linear2
        ops.layernorm_ex(ws.tgt, *Lw["norm3"], out=ws.xn, col_out=0)
        ops.linear(ws.xn[:, :256], Lw["w_ff1"], Lw["b_ff1"], [(0, prep.ff_dim, A, F32, ws.ff, 0)])
        ops.groupnorm_gelu_tokens(ws.ff, *Lw["gn"], ws.ffa, 32, ws.gn_ws)
        ops.dwconv5(ws.ffa, 0, Lw["dw"], ws.ffd, 0, h, w, prep.ff_dim)
        ops.linear(ws.ffd, Lw["w_ff2"], Lw["b_ff2"], [(0, 256, A, ACC, ws.tgt, 0)])
        # output norm of this layer straight into the decoder input
        ops.layernorm_ex(ws.tgt, *Lw["dec_norm"], out_f32=ws.dec_in, col_f32=256 * (i + 1))
        if tr is not None:
            tr.setdefault("layers", []).append(dict(
                curr_Q=ws.kcur[i].float().cpu(), tgt3=mem[:, 512:].float().cpu(), tgt=ws.tgt.cpu().clone()))

    def _write_body(self, label_hw, T, update_long, do_evict, st=0):
        """update_short_memories + update_long_term_memory + restrict_long_memories (transformer.py:269-436).
        (`st` is the DeAOT workspace's double-buffer index; the AOT short-term memory is produced by a GEMM.)"""
        ws = self.ws
        A, BF = ops.ACT_NONE, ops.OUT_BF16
        self._id_embed(label_hw)
        for i, Lw in enumerate(self.prep.layers):
            ws.kprev[i].copy_(ws.knew[i])
            ops.linear(ws.mem[i][:, 256:], Lw["w_vmem2"], Lw["b_vmem"], [(0, 256, A, BF, ws.vprev[i], 0)])
            if update_long:
                ops.linear(ws.mem[i][:, :512], Lw["w_v2"], Lw["b_v"], [(0, 256, A, BF, ws.vcur[i], 0)])
        if update_long:
            srcs, dsts = [], []
            for i in range(self.prep.num_layers):
                srcs += [ws.kcur[i], ws.vcur[i]]
                dsts += [ws.kbank[i], ws.vbank[i]]
            ops.bank_append(ws.bank_state[0], T, -1, srcs, dsts)
            if do_evict:
                ops.foreground_prob(self.pred_id_logits[0], ws.fg, ws.h, ws.w)
                ops.bank_score_evict(ws.bank_state[0], ws.mass, ws.fg, ws.L, T, self.capacity)

    def _decoder_inputs(self, embs, lstt_out):
        h, w = embs[-1].shape[-2:]
        return [lstt_out.view(1, h, w, -1).permute(0, 3, 1, 2)]               # channels-last view, no copy
