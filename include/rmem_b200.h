/* rmem_b200.h -- C ABI of librmem_b200.so: the B200 (sm_100a) kernels behind the RMem per-frame
 * memory-reading hot path of AOT / DeAOT (reference: Bardli/RMem_ocu, aot_plus/networks/...).
 *
 * The reference has no FFI of its own (it is pure PyTorch); these entry points are what a binding
 * for this path binds -- one per operator the reference evaluates on the path -- and each cites
 * the reference code it replaces.  INTEGRATION.md shows the ctypes stub a maintainer would add.
 *
 * Conventions
 *   - every pointer is a DEVICE pointer unless its name ends in _host; the caller owns all buffers,
 *     the library never allocates device memory; workspaces are passed in;
 *   - every call is asynchronous on `stream` (a cudaStream_t passed as void*), launches on the CURRENT device
 *     (cudaSetDevice is the caller's business) and is re-entrant across streams and host threads.  Process-wide
 *     state: a thread-local error string; per (kernel, device) "dynamic shared memory opt-in granted" flags
 *     (atomics); the SM count per device; a few switches read once from the environment (RMEM_PDL,
 *     RMEM_LOCAL_AGG_VARIANT); and the 256-byte device scratch of rmem_groupnorm_tokens_static (see there);
 *   - return value: RMEM_OK (0) or a negative RMEM_ERR_* code; rmem_last_error() explains it;
 *   - bf16 = IEEE bfloat16 (uint16 storage), f32 = IEEE binary32; row-major, innermost dim last.
 */
#ifndef RMEM_B200_H_
#define RMEM_B200_H_

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define RMEM_ABI_VERSION 11

#define RMEM_OK 0
#define RMEM_ERR_ARG (-1)     /* invalid argument (shape / capacity)           */
#define RMEM_ERR_CUDA (-2)    /* a CUDA runtime call or kernel launch failed  */
#define RMEM_ERR_DRIVER (-3)  /* driver entry point (tensor map encode) failed */

#define RMEM_MAX_BANK 16             /* max frames in the long-term bank (former + latter + 1) */
#define RMEM_MAX_CHUNKS_PER_FRAME 4  /* split-KV granularity inside one memory frame          */
#define RMEM_TPE_ROWS 4              /* rows of mem_pos_emb (TEMPORAL_POSITIONAL_EMBEDDING_SLOT_4) */

int rmem_abi_version(void);
const char* rmem_last_error(void);

/* ---------------------------------------------------------------------------------------------
 * Long-term gated attention, DeAOT (1 head, d_att 128, value width 1024 = V | ID_V).
 * Replaces GatedPropagation.forward explicit path, layers/attention.py:174-193, as called from
 * GatedPropagationModule.forward, layers/transformer.py:1183-1184, with the bank layout of
 * DualBranchGPM.long_term_memories (transformer.py:859-878) turned into fixed slots.
 *
 *   q        bf16 [L][128]            current-frame queries, cur_pos_emb already added
 *   kbank    bf16 [n_slots][L][128]   memory keys, one slot per stored frame (no TPE added)
 *   vbank    bf16 [n_slots][L][1024]  memory values (V | ID_V)
 *   order    i32  [T]                 logical bank position (oldest first) -> slot
 *   part_o   f32  [T*cpf][L][1024]    un-normalised partial outputs per key chunk
 *   part_m   f32  [T*cpf][L]          running max per chunk (log2 units, scaled by 1/sqrt(128))
 *   part_l   f32  [T*cpf][L]          softmax denominators per chunk
 */
int rmem_lt_attn_deaot_partial(const void* q, const void* kbank, const void* vbank,
                               const int32_t* order, int T, int L, int n_slots,
                               int chunks_per_frame, float* part_o, float* part_m, float* part_l,
                               void* stream);

/* Merge of the chunk partials + temporal positional embedding + recording + gate.
 * Replaces the TPE add of transformer.py:1140-1172 (as a per-frame score bias q'.e_row/sqrt(d)),
 * the softmax normalisation of attention.py:187-193, the attention-weight recording of
 * transformer.py:1185-1192 and the gate `outputs * U` of attention.py:206.
 *
 *   tpe_bias       f32 [L][4]    log2e/sqrt(128) * (q' . mem_pos_emb[r]), or NULL (no TPE)
 *   tpe_rows_host  i32 [T]       HOST array: mem_pos_emb row used by bank position t
 *   gate           bf16 [L][1024] cat_curr_U, or NULL
 *   out            bf16 [L][1024]
 *   attn_mass      f32 [L][T]    record_attn_weight, or NULL
 */
int rmem_lt_attn_deaot_combine(const float* part_o, const float* part_m, const float* part_l,
                               const float* tpe_bias, const int32_t* tpe_rows_host, int T, int L,
                               int chunks_per_frame, const void* gate, void* out,
                               float* attn_mass, void* stream);

/* The same operator when the keys are ONE frame (a layer's self-attention, attention.py:174-206 called from
 * transformer.py:1208-1216; or a bank that holds one frame and needs no recording): every key sits in one chunk, so
 * the softmax completes inside the kernel and its epilogue normalises, applies the gate and writes bf16 -- one launch,
 * no partials.  A temporal-PE bias would be a per-row constant here and is therefore not an argument.
 *
 *   q     bf16 [L][128]     k  bf16 [L][128]     v  bf16 [L][1024]
 *   gate  bf16 [L][1024] or NULL                 out bf16 [L][1024]
 */
int rmem_lt_attn_deaot_single(const void* q, const void* k, const void* v, int L, const void* gate, void* out,
                              void* stream);

/* The same operator, second generation (csrc/lt_attn_sk.cu): one persistent, stream-K scheduled kernel on 2-CTA
 * clusters that share every K / V tile by TMA multicast, followed by the merge of its segment partials.  It replaces
 * the same reference code as the pair above (attention.py:174-193, 206; transformer.py:1140-1172, 1183-1192); the
 * temporal-PE bias is added to the scores inside the kernel, so the combine step needs neither it nor T-many
 * per-frame partials.  Results equal the first generation's to bf16 rounding.
 *
 *   S           independent sequences in one launch (grid dimension): q [S][L][128], kbank [S][n_slots][L][128], vbank
 *               [S][n_slots][L][1024], tpe_bias [S][L][4], gate / out [S][L][1024], attn_mass [S][L][T]; `order` of
 *               sequence s starts at order + s * order_stride (e.g. RMEM_BANK_STATE_WORDS when the bank states are
 *               stacked); all sequences share T and the temporal rows.  S = 1 is the single-sequence call.
 *   n_clusters  0: one 2-CTA cluster per SM pair (the kernel then owns every SM while it runs); k > 0: at most k
 *               clusters, which leaves SMs to kernels running on other streams.  The three calls of one attention
 *               must be given the same value.
 *   workspace   rmem_lt_attn_deaot_sk_workspace_bytes(L, T, n_clusters) bytes (take the maximum over the bank
 *               lengths T in use), 256-byte aligned, owned by the caller, contents irrelevant on entry; partial
 *               writes it, combine reads it -- one workspace per attention call in flight
 *   tpe_bias    f32 [L][4] (log2 units, from rmem_tpe_prepare) or NULL;  tpe_rows_host  HOST i32 [T]
 *   record      != 0: also write the per-(segment, frame) sums that attn_mass is built from
 *   gate / out  bf16 [L][1024];  attn_mass f32 [L][T] or NULL (needs record != 0 in the partial call)
 */
int64_t rmem_lt_attn_deaot_sk_workspace_bytes(int L, int T, int S, int n_clusters);
int rmem_lt_attn_deaot_sk_partial(const void* q, const void* kbank, const void* vbank, const int32_t* order,
                                  int order_stride, int T, int L, int n_slots, int S, const float* tpe_bias,
                                  const int32_t* tpe_rows_host, int record, int n_clusters, void* workspace,
                                  int64_t workspace_bytes, void* stream);
int rmem_lt_attn_deaot_sk_combine(int T, int L, int S, const void* gate, void* out, float* attn_mass, int n_clusters,
                                  const void* workspace, int64_t workspace_bytes, void* stream);


/* ---------------------------------------------------------------------------------------------
 * Multi-head attention, AOT (8 heads x 32, d_model 256): MultiheadAttention.forward
 * (layers/attention.py:28-81, explicit path :45-64) in its three roles inside
 * SimplifiedTransformerBlock.forward (layers/transformer.py:553-692): long-term attention over the
 * restricted bank (:632-635), short-term attention over the previous frame (:657-662, T = 1) and
 * self-attention (:565-571, T = 1).  Same split-KV scheme as the DeAOT kernel.
 *
 *   q        bf16 [L][ldq]  (256 channels used)  queries (for the bank: cur_pos_emb already added)
 *   k, v     bf16 [n_slots][L][ld]  slot s starts at element s * slot_stride (ignored if n_slots == 1)
 *   order    i32  [T]       logical bank position -> slot
 *   part_o   f32  [T*cpf][L][256];  part_m, part_l  f32 [T*cpf][8][L]
 */
int rmem_mha_partial(const void* q, int ldq, const void* k, int ldk, int64_t k_slot_stride,
                     const void* v, int ldv, int64_t v_slot_stride, const int32_t* order, int T, int L,
                     int n_slots, int chunks_per_frame, float* part_o, float* part_m, float* part_l,
                     void* stream);

/* Merge of the chunk partials per head + temporal positional embedding (transformer.py:594-629, as a
 * per-(head, frame) score bias) + attention-weight recording averaged over heads (:636-643).
 *   tpe_bias  f32 [L][8][4] (log2 units) or NULL;  out bf16 [L][ld_out] at column out_col0;
 *   attn_mass f32 [L][T] or NULL
 */
int rmem_mha_combine(const float* part_o, const float* part_m, const float* part_l,
                     const float* tpe_bias, const int32_t* tpe_rows_host, int T, int L,
                     int chunks_per_frame, void* out, int ld_out, int out_col0, float* attn_mass,
                     void* stream);

/* q_time = bf16(q + cur_pos_emb) [L][256];  tpe_bias[q][head][r] = log2e/sqrt(32) * (q_time_head . mem_pos_emb[r]_head)
 * (transformer.py:594-629);  cur_pos_emb f32 [256], mem_pos_emb f32 [4][256]. */
int rmem_tpe_prepare_mha(const void* q, int ldq, const float* cur_pos_emb, const float* mem_pos_emb,
                         void* q_time, float* tpe_bias, int L, void* stream);

/* LayerNorm(256) of xa (+ xb), f32 or bf16 inputs, with up to three outputs: bf16 `out`, bf16
 * `out_plus` = LN(.) + post_add (q = k = norm1(x) + self_pos, transformer.py:563-564) and f32 `out_f32`
 * (decoder_norms, :248-259).  The two-input form is norm4(local_K + curr_K) of :657-660.  Unused
 * outputs / xb / post_add are NULL. */
int rmem_layernorm_ex(const void* xa, int lda, const void* xb, int ldb, int in_bf16, const float* gamma,
                      const float* beta, void* out, int ld_out, int col_out, const float* post_add,
                      int ld_add, void* out_plus, int ld_plus, int col_plus, float* out_f32, int ld_f32,
                      int col_f32, int L, int C, void* stream);

/* GroupNorm(groups, C = 32 * groups) + exact GELU over a token matrix x f32 [L][C] -> bf16 [L][C]
 * (GNActDWConv2d front half, layers/basic.py:15-35).  workspace: 16 * groups + 8 bytes, 8-byte
 * aligned, ZERO on first use; the call leaves it zero again (the last CTA re-zeroes it: no memset node per call). */
int rmem_groupnorm_gelu_tokens(const float* x, const float* gamma, const float* beta, void* out, int L,
                               int C, int groups, void* workspace, void* stream);

/* ---------------------------------------------------------------------------------------------
 * Linear layer with fused epilogue: C[M,N] = A[M,K] * W[N,K]^T + bias, then per column segment an
 * activation and a destination (so one launch produces e.g. K | silu(V) | silu(U) into three
 * buffers, or adds the projection straight into the fp32 residual streams).
 * Replaces nn.Linear + silu / split / cat / residual adds in GatedPropagationModule.forward
 * (layers/transformer.py:1104-1123, 1212-1232), GatedPropagation (attention.py:152-172, 209) and
 * fuse_key_value_id (transformer.py:1238-1244).
 *   a   bf16 [M][lda]   w  bf16 [N][ldw]   bias f32 [N] or NULL
 */
#define RMEM_GEMM_MAX_SEGMENTS 4
#define RMEM_ACT_NONE 0
#define RMEM_ACT_SILU 1
#define RMEM_OUT_BF16 0       /* dst bf16, store          */
#define RMEM_OUT_F32 1        /* dst f32, store           */
#define RMEM_OUT_F32_ACCUM 2  /* dst f32, dst += value    */
typedef struct rmem_gemm_segment {
  int32_t col_begin;  /* first output column of the segment, multiple of 32 */
  int32_t col_end;    /* one past the last output column                    */
  int32_t act;        /* RMEM_ACT_*                                         */
  int32_t mode;       /* RMEM_OUT_*                                         */
  void* dst;          /* destination matrix                                  */
  int32_t ld;         /* destination leading dimension (elements)           */
  int32_t dst_col0;   /* destination column of col_begin                    */
} rmem_gemm_segment;
int rmem_linear(const void* a, int lda, const void* w, int ldw, const float* bias, int M, int N,
                int K, const rmem_gemm_segment* segments_host, int n_segments, void* stream);

/* rmem_linear with IEEE fp16 operands when ab_f16 != 0 (a and w fp16 instead of bf16; everything else as above):
 * the decoder's 1x1 adapter convolutions (decoders/fpn.py:49-60) read the backbone's feature maps, which are fp16
 * in the default configuration (fp16 keeps the masks within 99.9 % of the fp32 reference, bf16 does not). */
int rmem_linear_ex(const void* a, int lda, const void* w, int ldw, const float* bias, int M, int N, int K,
                   const rmem_gemm_segment* segments_host, int n_segments, int ab_f16, void* stream);

/* LayerNorm over the last dim (C = 256) of up to two [L][C] streams in one launch, bf16 output
 * written at a column offset of a wider buffer (so norm2 | id_norm2 land concatenated).
 * Replaces nn.LayerNorm calls norm1 / id_norm1 / norm2 / id_norm2 (transformer.py:1104, 1120,
 * 1223-1225) and DeAOT.get_id_emb's id_norm (models/deaot.py:65-69).
 *   x{0,1}  f32 (in_bf16 = 0) or bf16 (in_bf16 = 1) [L][C];  gamma/beta f32 [C];  x1 may be NULL
 */
int rmem_layernorm(const void* x0, const float* gamma0, const float* beta0, void* out0, int ld0,
                   int col0, const void* x1, const float* gamma1, const float* beta1, void* out1,
                   int ld1, int col1, int L, int C, int in_bf16, void* stream);

/* The first LayerNorm of the first block (norm1 of layer 0, transformer.py:1104) reads the encoder's 16-bit embedding
 * and starts the fp32 residual stream: LN(x) as bf16 at out[:, col0 : col0 + 256] and x itself as f32 in copy_f32
 * [L][256] (the reference's `tgt = curr_embs`, transformer.py:785-790), one kernel instead of a copy + a LayerNorm.
 *   in_kind 0: x f32, 1: bf16, 2: IEEE fp16 (the fp16 backbone); copy_f32 may be NULL. */
int rmem_layernorm_copy(const void* x, int in_kind, const float* gamma, const float* beta, void* out, int ld, int col0,
                        float* copy_f32, int L, int C, void* stream);

/* 5x5 depth-wise convolution over the (h, w) token grid, channels-last bf16, zero padding 2.
 * Replaces DWConv2d.forward (layers/basic.py:38-57) behind the gate of GatedPropagation /
 * LocalGatedPropagation (attention.py:208, 357).
 *   x bf16 [h*w][ldx] (channels at col xcol0..xcol0+C)   wt f32 [25][C] (tap-major repack of
 *   conv.weight [C,1,5,5])   y bf16 [h*w][ldy] at column ycol0
 */
int rmem_dwconv5(const void* x, int ldx, int xcol0, const float* wt, void* y, int ldy, int ycol0,
                 int h, int w, int C, void* stream);

/* GroupNorm over an [L][C] token matrix viewed as [1, C, L] (GroupNorm1D, layers/basic.py:6-12):
 * statistics per group over (C/groups channels x L tokens).  Replaces DualBranchGPM's
 * decoder_norms[-1] (transformer.py:806-808).  x = cat(x_a[L][Ca], x_b[L][C-Ca]) f32; out [L][C] bf16, or
 * f32 when out_f32 == 1, IEEE fp16 when out_f32 == 2 (whatever the decoder's convolutions take).  Ca, C and C / groups
 * are multiples of 4.
 *   workspace: 16 * groups + 8 bytes, 8-byte aligned (fp64 sum / sum of squares + an arrival counter); the
 *   caller zero-fills it once, every call leaves it zero again
 */
int rmem_groupnorm_tokens(const float* x_a, int Ca, const float* x_b, const float* gamma,
                          const float* beta, void* out, int out_f32, int L, int C, int groups,
                          void* workspace, void* stream);

/* rmem_groupnorm_tokens without a caller workspace, for callers that cannot allocate device memory (code under
 * CUDA-graph capture, tight serving loops, harnesses that forbid cudaMalloc).  This entry point has NO counterpart
 * in the reference.  Same arguments and results as
 * rmem_groupnorm_tokens minus `workspace`: the statistics scratch is a 256-byte file-scope `static __device__`
 * array inside the library (csrc/fused.cu), resident on every GPU that loads the library until the process exits;
 * its device address is resolved once per device (cudaGetSymbolAddress, on the first call with that device
 * current).  The call allocates nothing (no cudaMalloc / cudaFree, no allocating library routine); it issues one
 * cudaMemsetAsync of the scratch on `stream` and the two GroupNorm kernels.
 *   Limit: groups <= RMEM_GN_STATIC_MAX_GROUPS (the scratch holds 16 * groups + 8 bytes), else RMEM_ERR_ARG with
 *   an explanatory rmem_last_error().
 *   Sharing: every call on a device uses the SAME array, which keeps the last call's contents (each call clears it
 *   first).  Calls that overlap on the GPU (different streams, or graph branches) race on it and give wrong
 *   results: serialise them, or use rmem_groupnorm_tokens with one workspace per concurrent call. */
#define RMEM_GN_STATIC_MAX_GROUPS 8
int rmem_groupnorm_tokens_static(const float* x_a, int Ca, const float* x_b, const float* gamma,
                                 const float* beta, void* out, int out_f32, int L, int C, int groups,
                                 void* stream);

/* Per-query temporal-PE score bias and the time-shifted query (transformer.py:1142-1172):
 *   q_time = bf16(q + cur_pos_emb);  tpe_bias[q][r] = log2e/sqrt(C) * ((q + cur_pos_emb) . mem_pos_emb[r])
 *   q bf16 [L][C] (C = 128);  cur_pos_emb f32 [C];  mem_pos_emb f32 [4][C]
 */
int rmem_tpe_prepare(const void* q, const float* cur_pos_emb, const float* mem_pos_emb,
                     void* q_time, float* tpe_bias, int L, int C, void* stream);

/* ---------------------------------------------------------------------------------------------
 * Short-term local attention, DeAOT: 15x15 window (max_dis 7, dilation 1), 1 head, d_att 128,
 * value width 1024.  Replaces LocalGatedPropagation.forward (layers/attention.py:289-361):
 * F.unfold correlation (:331-335), relative-position bias (:314, 342), border mask (:305-312,
 * 344), softmax over the 225 taps (:346), local2global scatter + dense matmul (:350-353, 363-402)
 * and the gate (:355).
 *
 * Stage 1 (scores): probs[q][t] = softmax_t( q.k[nbr(q,t)]/sqrt(128) + rel_bias[q][t] ) over the
 *   in-image taps (out-of-image taps get probability 0, as exp(-1e8) does in fp32).
 *   q, k bf16 [h*w][128];  rel_bias f32 [h*w][ld_rel] (= q W_rel^T + b_rel, from rmem_linear);
 *   probs f32 [h*w][225] or NULL (this is local_attn[0,0,:,q] of the reference, transposed);
 *   banded bf16 [rmem_local_banded_elems(h, w)] or NULL: the same probabilities in the per-tile banded
 *   layout stage 2 multiplies with (8x16 query tiles, 11 blocks of 2 halo rows x 32 key slots; see
 *   csrc/local_attn.cu).  Only in-window slots are written: the CALLER zero-fills the buffer once
 *   (per sequence / geometry); it stays valid across frames.
 * Stage 2 (aggregate): out[q][c] = gate[q][c] * sum_t probs[q][t] * v[nbr(q,t)][c]
 *   banded from stage 1;  v bf16 [h*w][1024];  gate bf16 [h*w][1024] or NULL;  out bf16 [h*w][1024]
 */
int64_t rmem_local_banded_elems(int h, int w);
int rmem_local_attn_scores(const void* q, const void* k, const float* rel_bias, int ld_rel,
                           float* probs, void* banded, int h, int w, void* stream);
int rmem_local_attn_aggregate(const void* banded, const void* v, const void* gate, void* out,
                              int h, int w, void* stream);

/* The same operator in ONE kernel (flash-attention form, csrc/local_attn_fused.cu): scores, relative-position
 * bias, border / window mask, softmax, aggregation and gate without materialising the probabilities --
 *   out[q][c] = gate[q][c] * sum_t softmax_t( q.k[nbr(q,t)]/sqrt(128) + rel_bias[q][t] ) v[nbr(q,t)][c]
 * q, k bf16 [h*w][128]; v, gate (or NULL), out bf16 [h*w][1024]; rel_bias f32 [h*w][ld_rel], ld_rel >= 225.
 * This is what the engines call; the two-stage form above remains for callers that want local_attn itself. */
int rmem_local_attn(const void* q, const void* k, const void* v, const float* rel_bias, int ld_rel,
                    const void* gate, void* out, int h, int w, void* stream);

/* ---------------------------------------------------------------------------------------------
 * Sequence-batched forms: S independent video sequences of the same frame size step in lock-step and share every
 * launch (the reference asserts batch size 1 on this path, transformer.py:641; one sequence is 1620 tokens -- a
 * fraction of a wave on 148 SMs -- so independent sequences are what fills the chip).  Each tensor of the
 * single-sequence call becomes a contiguous stack [S][...] of the same layout; the sequence is a grid dimension.
 * rmem_linear, rmem_layernorm, rmem_tpe_prepare are row-wise and take the stack as M = S * L rows unchanged.
 *   rmem_dwconv5_batched               x, y: [S * h * w][ld] rows
 *   rmem_groupnorm_tokens_batched      statistics per sequence; workspace: S consecutive (16 * groups + 8)-byte blocks
 *   rmem_local_attn_scores_batched     q, k [S][h*w][128]; rel_bias [S][h*w][ld_rel]; probs [S][h*w][225]; banded
 *                                      [S][rmem_local_banded_elems(h, w)]
 *   rmem_local_attn_aggregate_batched  v, gate, out [S][h*w][1024]
 *   rmem_bank_tick_batched             bank_states [S][RMEM_BANK_STATE_WORDS]
 *   rmem_id_embed_batched              labels [S][H][W]; destination rows [S * h * w][ld]
 *   rmem_groupnorm_nhwc_batched        x, y [S][P][C], statistics per image; workspace: S consecutive blocks
 *   rmem_upsample_add_nhwc_batched     x [S][hi*wi][C]; a, out [S][ho*wo][C]
 *   rmem_finish_logits_batched         conv image s at conv + s * sn; pred [S][C][Hs][Ws]; out [S][C][Ho][Wo];
 *                                      n_valid_host: HOST i32 [S] valid channels per image; S <= 64
 *   rmem_pack_image_s2d_batched, rmem_maxpool3s2_nhwc_batched, rmem_argmax_label_batched   image stacks
 */
int rmem_bank_tick_batched(void* bank_states, int S, void* stream);
int rmem_id_embed_batched(const float* labels, int H, int W, const float* prefix, const float* bias,
                          const float* gamma, const float* beta, int n_cls, void* const* dst_host,
                          const int32_t* ld_host, const int32_t* col0_host, int n_dst, int h, int w, int S,
                          void* stream);
int rmem_groupnorm_nhwc_batched(const float* x, const float* pre_bias, const float* gamma, const float* beta, float* y,
                                int P, int C, int groups, int relu, int S, void* workspace, void* stream);
int rmem_upsample_add_nhwc_batched(const float* x, int hi, int wi, const float* a, const float* a_bias, float* out,
                                   int ho, int wo, int C, int S, void* stream);
int rmem_finish_logits_batched(const float* conv, int64_t sn, int64_t sc, int64_t sy, int64_t sx, const float* bias,
                               float* pred, int C, int Hs, int Ws, const int32_t* n_valid_host, int S, float* out,
                               int Ho, int Wo, void* stream);
int rmem_pack_image_s2d_batched(const void* img, int img_u8_hwc, const float* mean_host, const float* inv_std_host,
                                int H, int W, void* out, int out_f16, int S, void* stream);
int rmem_maxpool3s2_nhwc_batched(const void* x, int H, int W, int C, const float* bias, int relu, void* y, int f16,
                                 int S, void* stream);
int rmem_argmax_label_batched(const float* logits, int C, int HW, float* label, int S, void* stream);

/* The decoder's last stage in one pass (decoders/fpn.py:63-68: conv_4x's GroupNorm(8) + ReLU, then the 1x1 conv_out;
 * aot_engine.py:451-463: classes beyond the sequence's objects -> -1e10, bilinear align_corners resize to the output
 * size; and the evaluator's argmax, networks/managers/evaluator.py:430-441).  The normalised feature map is never
 * written: statistics, then GroupNorm + ReLU + the C -> Co dot products + bias + mask straight into pred.
 *   x        f32 (x_f16 = 0) or fp16 (x_f16 = 1) [S][Hs*Ws][C] raw output of the 3x3 conv_4x (channels-last), C = 64 or 128
 *   pre_bias f32 [C] conv_4x bias or NULL; gamma, beta f32 [C]; groups (8)
 *   w_out    f32 [Co][C], b_out f32 [Co], Co <= 16
 *   n_valid_host  HOST i32 [S]: objects + 1 per image;  S <= 64
 *   pred     f32 [S][Co][Hs][Ws]  (= pred_id_logits)
 *   out      f32 [S][Co][Ho][Wo] or NULL;  label f32 [S][Ho][Wo] or NULL (argmax over the classes of out, first maximum)
 *   workspace: as rmem_groupnorm_nhwc_batched (S blocks of 16 * groups + 8 bytes, zero before the first call)
 */
int rmem_logits_head_batched(const void* x, int x_f16, const float* pre_bias, const float* gamma, const float* beta, int Hs,
                             int Ws, int C, int groups, const float* w_out, const float* b_out, int Co,
                             const int32_t* n_valid_host, int S, float* pred, float* out, int Ho, int Wo, float* label,
                             void* workspace, void* stream);

/* rmem_groupnorm_nhwc_batched / rmem_upsample_add_nhwc_batched for a decoder that keeps its activations in fp16
 * (decoders/fpn.py:44-63 with 16-bit cuDNN convolutions: 62 -> 38 us for the four 3x3 convolutions at 480p, half the
 * bytes in these kernels; statistics, normalisation and interpolation are computed in fp32 either way).
 *   f16 = 1: x / y (GroupNorm) and x / out (upsample) are fp16, channels-last; 0: fp32 as the plain entry points.
 *   the adapter branch `a` (+ a_bias) stays fp32: it comes from a GEMM epilogue and is read once.
 */
/* One decoder stage in one pass (decoders/fpn.py:44-63): GroupNorm(groups) + ReLU of a convolution's raw output x
 * [S][hi*wi][C] (+ pre_bias), bilinear align_corners resize to ho x wo, plus the next shortcut's adapter branch a
 * [S][ho*wo][C] f32 (+ a_bias): two launches (statistics, then everything else); the normalised map is not stored.
 * f16: x and out are fp16 (else fp32).  workspace as rmem_groupnorm_nhwc_batched. */
int rmem_gn_upsample_add_nhwc(const void* x, const float* pre_bias, const float* gamma, const float* beta, int hi, int wi,
                              int C, int groups, const float* a, const float* a_bias, void* out, int ho, int wo, int S,
                              int f16, void* workspace, void* stream);
int rmem_groupnorm_nhwc_ex(const void* x, const float* pre_bias, const float* gamma, const float* beta, void* y, int P,
                           int C, int groups, int relu, int S, int f16, void* workspace, void* stream);
int rmem_upsample_add_nhwc_ex(const void* x, int hi, int wi, const float* a, const float* a_bias, void* out, int ho, int wo,
                              int C, int S, int f16, void* stream);
int rmem_dwconv5_batched(const void* x, int ldx, int xcol0, const float* wt, void* y, int ldy, int ycol0, int h, int w,
                         int C, int S, void* stream);
int rmem_groupnorm_tokens_batched(const float* x_a, int Ca, const float* x_b, const float* gamma, const float* beta,
                                  void* out, int out_f32, int L, int C, int groups, int S, void* workspace,
                                  void* stream);
int rmem_local_attn_scores_batched(const void* q, const void* k, const float* rel_bias, int ld_rel, float* probs,
                                   void* banded, int h, int w, int S, void* stream);
int rmem_local_attn_aggregate_batched(const void* banded, const void* v, const void* gate, void* out, int h, int w,
                                      int S, void* stream);

/* ---------------------------------------------------------------------------------------------
 * Restricted long-term bank.  State lives on the device so the frame loop never synchronises:
 *   order     i32 [RMEM_MAX_BANK]  logical position (oldest first) -> slot; order[T..] = free slots
 *   frame_id  i32 [RMEM_MAX_BANK]  per SLOT: frame_step stored there
 *   ema       f32 [RMEM_MAX_BANK]  per SLOT: moving-mean relevance (stored_attn_weight_dict)
 *   ema_valid i32 [RMEM_MAX_BANK]  per SLOT: 1 once the slot has a stored relevance
 *   times     i32 [RMEM_MAX_BANK]  per SLOT: UCB visit count (stored_frame_times)
 * laid out consecutively in one int32/f32 block of RMEM_BANK_STATE_WORDS words.
 */
#define RMEM_BANK_STATE_WORDS (5 * RMEM_MAX_BANK + 8)
/* words [5*RMEM_MAX_BANK ..): 0 = last drop index (logical, -1 none), 1 = dropped frame id,
 * 2 = decision margin (f32 bits), 3 = number of evictions so far, 4 = argmin of the last scoring
 * pass (set even when nothing had to be dropped), 5 = device-side frame_step */

/* Reset the state for a new sequence: order = identity, slot 0 holds the reference frame
 * (init_memory, transformer.py:438-443 / 993-998 + aot_engine.py:321-323). */
int rmem_bank_init(void* bank_state, int frame_step, void* stream);

/* frame_step += 1 on the device (AOTEngine.match_propogate_one_frame, aot_engine.py:399); keeps the
 * frame counter in the state block so a captured CUDA graph needs no per-frame host scalar. */
int rmem_bank_tick(void* bank_state, void* stream);

/* Append: copy the candidate rows into the free slot order[T] of every bank tensor and stamp
 * frame_id (frame_step < 0: use the device-side counter). Replaces update_long_term_memory (transformer.py:306-322, 859-878) -- torch.cat
 * re-allocation becomes one strided copy.  n_tensors <= 8.
 *   src_host / dst_host: HOST arrays of device pointers; row_bytes_host: bytes per slot.
 */
int rmem_bank_append(void* bank_state, int T, int frame_step, const void* const* src_host,
                     void* const* dst_host, const int64_t* row_bytes_host, int n_tensors,
                     void* stream);

/* Relevance + freshness scoring and eviction, in place (restrict_long_memories,
 * transformer.py:324-436 AOT / 880-991 DeAOT, use_atten_weight=True):
 *   s_t = sum_q W[q,t] fg[q]; r = s / sum s; EMA 0.2/0.8 with the stored value of the same frame;
 *   times += 1 for all T_old+1 frames; n_t = times (n_0 = T_old); bonus = 1.5 sqrt(log(sum n)/(n+8));
 *   drop = 1 + argmin_{t>=1}(r_t + bonus_t); if T_old + 1 > capacity: remove position `drop`
 *   from `order` (later positions shift down, the slot becomes free) and clear its statistics.
 * attn_mass f32 [L][T_old] (layer-0 record_attn_weight), fg f32 [L] or NULL.
 * The AOT variant returns before any statistics while the bank is filling (transformer.py:332-334):
 * its caller simply does not call this function until T_old + 1 > capacity.
 */
int rmem_bank_score_evict(void* bank_state, const float* attn_mass, const float* fg, int L,
                          int T_old, int capacity, void* stream);

/* Identity embedding of a label map in one gather kernel: one_hot_mask + ignore channel
 * (utils/image.py:69-74, aot_engine.py:208-225) + the 17x17 stride-16 pad-8 patch_wise_id_bank conv
 * (models/aot.py:64-83, 111-114) + optionally DeAOT's id_norm LayerNorm (models/deaot.py:65-69).
 *   label  f32 [H][W]: 0..n_cls-2 object ids, 255 = ignore (class n_cls-1), anything else = no class
 *   prefix f32: [n_cls][17][18][256] row-prefix sums of conv.weight, prefix[k][ky][j][c] =
 *          sum_{kx<j} weight[c][k][ky][kx], followed by [n_cls][18][18][256] 2-D prefix sums
 *          prefix2[k][i][j][c] = sum_{ky<i, kx<j} weight[c][k][ky][kx] (window rows with the same run structure
 *          are summed as rectangles);   bias f32 [256];  gamma/beta f32 [256] or NULL
 *   dst_host: HOST array of n_dst (<= 4) bf16 destinations [h*w][ld_host[i]], written at col0_host[i]
 */
int rmem_id_embed(const float* label, int H, int W, const float* prefix, const float* bias,
                  const float* gamma, const float* beta, int n_cls, void* const* dst_host,
                  const int32_t* ld_host, const int32_t* col0_host, int n_dst, int h, int w,
                  void* stream);

/* The same embedding from a one-hot or probability mask that already has one channel per identity
 * (update_short_term_memory's second branch, aot_engine.py:330-336 -> assign_identity :208-232): channel 0 is
 * multiplied by (ignore == 0) and the ignore map becomes the last class (:214-224), then the 17x17 stride-16 conv
 * (models/aot.py:111-114) as a weighted sum and DeAOT's id_norm (models/deaot.py:65-69).
 *   prob   f32 [S][n_ch][H][W], n_ch <= n_cls <= 16 (missing channels count as zero)
 *   ignore f32 [S][H][W] or NULL
 *   taps   f32 [n_cls][17][17][256] = conv.weight[c][k][ky][kx] with the output channel innermost
 *   destinations as rmem_id_embed, rows [S * h * w]
 */
int rmem_id_embed_soft(const float* prob, int n_ch, const float* ignore, int H, int W, const float* taps,
                       const float* bias, const float* gamma, const float* beta, int n_cls, void* const* dst_host,
                       const int32_t* ld_host, const int32_t* col0_host, int n_dst, int h, int w, int S, void* stream);

/* ---------------------------------------------------------------------------------------------
 * Decoding head (FPNSegmentationHead, decoders/fpn.py:36-68): the convolutions stay cuDNN calls; the
 * GroupNorm(8) + ReLU of every ConvGN (layers/basic.py:60-70, fpn.py:44-63) and the bilinear
 * (align_corners=True) upsampling + shortcut add (fpn.py:49-60) run here on the channels-last
 * [H*W][C] view.
 *   rmem_groupnorm_nhwc: x, y f32 [P][C] (C = 64 / 128 / 256, C/groups a power of two in [4,32]);
 *       pre_bias f32 [C] or NULL is added to x first (the conv bias of ConvGN, so the convolution itself runs
 *       without a separate bias pass); gamma/beta f32 [C]; relu != 0 applies max(.,0); workspace 16*groups + 8
 *       bytes, zero on first use, left zero by the call.
 *   rmem_upsample_add_nhwc: out[ho*wo][C] = a[ho*wo][C] (or 0 when NULL) + a_bias[C] (or 0) + bilinear(x[hi*wi][C]).
 */
int rmem_groupnorm_nhwc(const float* x, const float* pre_bias, const float* gamma, const float* beta, float* y,
                        int P, int C, int groups, int relu, void* workspace, void* stream);
int rmem_upsample_add_nhwc(const float* x, int hi, int wi, const float* a, const float* a_bias, float* out,
                           int ho, int wo, int C, void* stream);

/* Tail of decode_current_logits (aot_engine.py:451-463): pred[c][y][x] = conv[c*sc + y*sy + x*sx] + bias[c] for
 * c < n_valid, -1e10 for the channels beyond the sequence's objects (:455-457); then, when out != NULL, the
 * bilinear align_corners=True resize pred [C][Hs][Ws] -> out [C][Ho][Wo] (:459-463).  Strides in elements. */
int rmem_finish_logits(const float* conv, int64_t sc, int64_t sy, int64_t sx, const float* bias, float* pred,
                       int C, int Hs, int Ws, int n_valid, float* out, int Ho, int Wo, void* stream);

/* Encoder input packing (AOT.encode_image front, networks/models/aot.py:116-119 + encoders/resnet.py:150-153):
 * img f32 [3][H][W] -> out bf16 [H][W][8] channels-last, channels 3..7 zero. */
int rmem_pack_image(const float* img, int H, int W, void* out, void* stream);

/* Encoder input packing for the space-to-depth form of the stem convolution (resnet.py:150-151: the 7x7 stride-2
 * conv on 3 channels == a 4x4 stride-1 conv on the 2x2 space-to-depth image): img f32 [3][H][W] -> out bf16
 * channels-last [Hs+3][Ws+3][16], Hs = (H+1)/2, Ws = (W+1)/2, out[y'+2][x'+2][(dy*2+dx)*3+c] = img[c][2y'+dy][2x'+dx],
 * zero elsewhere (conv padding, out-of-image pixels, channels 12..15). */
int rmem_pack_image_s2d(const float* img, int H, int W, void* out, void* stream);

/* The same packing with a choice of input and output types.  img_u8_hwc != 0: img is u8 interleaved [H][W][3]
 * (a decoded frame as it comes off the host) and the normalisation of the reference's dataloader
 * (dataloaders/video_transforms.py ToTensor: (x / 255 - mean) / std) happens here, (u8 / 255 - mean_host[c]) *
 * inv_std_host[c] with mean_host / inv_std_host HOST arrays of 3 floats -- a quarter of the upload bytes of the f32
 * planes; img_u8_hwc == 0: f32 planes as above.  out_f16 != 0 writes IEEE fp16 instead of bf16. */
int rmem_pack_image_s2d_ex(const void* img, int img_u8_hwc, const float* mean_host, const float* inv_std_host,
                           int H, int W, void* out, int out_f16, void* stream);

/* Encoder stem max-pool, MaxPool2d(kernel 3, stride 2, padding 1) (networks/encoders/resnet.py:154):
 * x bf16 channels-last [H][W][C] -> y bf16 [(H-1)/2+1][(W-1)/2+1][C]; C a multiple of 8.  bias f32 [C] (or NULL)
 * and relu != 0 give y = relu(maxpool(x) + bias) = maxpool(relu(x + bias)): the stem's conv bias (folded BN,
 * resnet.py:151-153) and ReLU applied after the pool, on a quarter of the pixels. */
int rmem_maxpool3s2_nhwc(const void* x, int H, int W, int C, const float* bias, int relu, void* y, void* stream);

/* rmem_maxpool3s2_nhwc on IEEE fp16 maps when f16 != 0 (bf16 otherwise). */
int rmem_maxpool3s2_nhwc_ex(const void* x, int H, int W, int C, const float* bias, int relu, void* y, int f16,
                            void* stream);

/* label[i] = (f32) argmax_c logits[c][i], first maximum -- the evaluator's softmax + argmax over the object
 * channel (networks/managers/evaluator.py:430-441); logits f32 [C][HW], label f32 [HW]. */
int rmem_argmax_label(const float* logits, int C, int HW, float* label, void* stream);

/* fg[q] = 1 - softmax_c(bilinear_align_corners(pred_id_logits -> (h,w)))[0]
 * (aot_engine.py:355-362).  logits f32 [C][Hs][Ws] -> fg f32 [h*w]. */
int rmem_foreground_prob(const float* logits, int C, int Hs, int Ws, float* fg, int h, int w,
                         void* stream);

#ifdef __cplusplus
}
#endif
#endif /* RMEM_B200_H_ */
