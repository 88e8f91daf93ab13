#include <cstdlib>
// extern "C" surface of librmem_b200.so (declared in include/rmem_b200.h) + error plumbing.
#include <stdio.h>
#include <string.h>
#include "rmem_internal.h"

namespace rmem {

static thread_local char g_err[512] = "";

int set_error(int code, const char* msg) {
  snprintf(g_err, sizeof(g_err), "%s", msg);
  return code;
}
int set_cuda_error(cudaError_t e, const char* where) {
  snprintf(g_err, sizeof(g_err), "%s: %s", where, cudaGetErrorString(e));
  return RMEM_ERR_CUDA;
}
int check_launch(const char* kernel_name) {
  cudaError_t e = cudaPeekAtLastError();
  if (e != cudaSuccess) {
    cudaGetLastError();
    return set_cuda_error(e, kernel_name);
  }
  return RMEM_OK;
}

TensorMapEncodeFn get_tensor_map_encoder() {
  // function-local static with an initialiser: resolved once, thread-safe (C++11)
  static const TensorMapEncodeFn fn = []() -> TensorMapEncodeFn {
    void* p = nullptr;
    cudaDriverEntryPointQueryResult qres;
    cudaError_t e = cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &qres);
    if (e == cudaSuccess && qres == cudaDriverEntryPointSuccess) return reinterpret_cast<TensorMapEncodeFn>(p);
    return nullptr;
  }();
  return fn;
}

int ensure_dynamic_smem(const void* kernel, SmemOptIn* rec, int bytes, const char* what) {
  int dev = -1;
  cudaError_t e = cudaGetDevice(&dev);
  if (e != cudaSuccess) return set_cuda_error(e, what);
  if (dev < 0 || dev >= kMaxDevices) return set_error(RMEM_ERR_ARG, "device ordinal >= 64");
  if (rec->granted[dev].load(std::memory_order_acquire) >= bytes) return RMEM_OK;
  e = cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, bytes);
  if (e != cudaSuccess) return set_cuda_error(e, what);
  // racing threads both set the attribute (idempotent); the record only ever grows
  int seen = rec->granted[dev].load(std::memory_order_relaxed);
  while (seen < bytes && !rec->granted[dev].compare_exchange_weak(seen, bytes, std::memory_order_release)) {}
  return RMEM_OK;
}

}  // namespace rmem

using namespace rmem;

namespace rmem {
bool pdl_enabled() {
  static const bool on = [] {
    const char* e = getenv("RMEM_PDL");
    return e && e[0] == '1';
  }();
  return on;
}
}  // namespace rmem

extern "C" {

int rmem_abi_version(void) { return RMEM_ABI_VERSION; }
const char* rmem_last_error(void) { return g_err; }

int rmem_lt_attn_deaot_partial(const void* q, const void* kbank, const void* vbank,
                               const int32_t* order, int T, int L, int n_slots,
                               int chunks_per_frame, float* part_o, float* part_m, float* part_l,
                               void* stream) {
  return lt_attn_partial_launch(q, kbank, vbank, order, T, L, n_slots, chunks_per_frame, part_o,
                                part_m, part_l, static_cast<cudaStream_t>(stream));
}

int rmem_lt_attn_deaot_combine(const float* part_o, const float* part_m, const float* part_l,
                               const float* tpe_bias, const int32_t* tpe_rows_host, int T, int L,
                               int chunks_per_frame, const void* gate, void* out,
                               float* attn_mass, void* stream) {
  return lt_attn_combine_launch(part_o, part_m, part_l, tpe_bias, tpe_rows_host, T, L,
                                chunks_per_frame, gate, out, attn_mass,
                                static_cast<cudaStream_t>(stream));
}

int rmem_lt_attn_deaot_single(const void* q, const void* k, const void* v, int L, const void* gate, void* out,
                              void* stream) {
  return lt_attn_direct_launch(q, k, v, L, gate, out, static_cast<cudaStream_t>(stream));
}

#define ST(s) static_cast<cudaStream_t>(s)

int64_t rmem_lt_attn_deaot_sk_workspace_bytes(int L, int T, int S, int n_clusters) {
  return lt_attn_sk_workspace_bytes(L, T, S, n_clusters);
}
int rmem_lt_attn_deaot_sk_partial(const void* q, const void* kbank, const void* vbank, const int32_t* order,
                                  int order_stride, int T, int L, int n_slots, int S, const float* tpe_bias,
                                  const int32_t* tpe_rows_host, int record, int n_clusters, void* workspace,
                                  int64_t workspace_bytes, void* stream) {
  return lt_attn_sk_partial_launch(q, kbank, vbank, order, order_stride, T, L, n_slots, S, tpe_bias, tpe_rows_host,
                                   record, n_clusters, workspace, workspace_bytes, ST(stream));
}
int rmem_lt_attn_deaot_sk_combine(int T, int L, int S, const void* gate, void* out, float* attn_mass, int n_clusters,
                                  const void* workspace, int64_t workspace_bytes, void* stream) {
  return lt_attn_sk_combine_launch(T, L, S, gate, out, attn_mass, n_clusters, workspace, workspace_bytes, ST(stream));
}
int rmem_bank_tick_batched(void* bank_states, int S, void* stream) {
  return bank_tick_batched_launch(bank_states, S, ST(stream));
}
int rmem_id_embed_batched(const float* labels, int H, int W, const float* prefix, const float* bias,
                          const float* gamma, const float* beta, int n_cls, void* const* dst_host,
                          const int32_t* ld_host, const int32_t* col0_host, int n_dst, int h, int w, int S,
                          void* stream) {
  return id_embed_batched_launch(labels, H, W, prefix, bias, gamma, beta, n_cls, dst_host, ld_host, col0_host, n_dst,
                                 h, w, S, ST(stream));
}
int rmem_logits_head_batched(const void* x, int x_f16, const float* pre_bias, const float* gamma, const float* beta, int Hs,
                             int Ws, int C, int groups, const float* w_out, const float* b_out, int Co,
                             const int32_t* n_valid_host, int S, float* pred, float* out, int Ho, int Wo, float* label,
                             void* workspace, void* stream) {
  return logits_head_batched_launch(x, x_f16, pre_bias, gamma, beta, Hs, Ws, C, groups, w_out, b_out, Co, n_valid_host, S,
                                    pred, out, Ho, Wo, label, workspace, ST(stream));
}
int rmem_layernorm_copy(const void* x, int in_kind, const float* gamma, const float* beta, void* out, int ld, int col0,
                        float* copy_f32, int L, int C, void* stream) {
  return layernorm_copy_launch(x, gamma, beta, out, ld, col0, nullptr, nullptr, nullptr, nullptr, 0, 0, L, C, in_kind,
                               copy_f32, ST(stream));
}
int rmem_gn_upsample_add_nhwc(const void* x, const float* pre_bias, const float* gamma, const float* beta, int hi, int wi,
                              int C, int groups, const float* a, const float* a_bias, void* out, int ho, int wo, int S,
                              int f16, void* workspace, void* stream) {
  return gn_upsample_add_nhwc_launch(x, pre_bias, gamma, beta, hi, wi, C, groups, a, a_bias, out, ho, wo, S, f16, workspace,
                                     ST(stream));
}
int rmem_groupnorm_nhwc_ex(const void* x, const float* pre_bias, const float* gamma, const float* beta, void* y, int P,
                           int C, int groups, int relu, int S, int f16, void* workspace, void* stream) {
  return groupnorm_nhwc_ex_launch(x, pre_bias, gamma, beta, y, P, C, groups, relu, S, f16, workspace, ST(stream));
}
int rmem_upsample_add_nhwc_ex(const void* x, int hi, int wi, const float* a, const float* a_bias, void* out, int ho, int wo,
                              int C, int S, int f16, void* stream) {
  return upsample_add_nhwc_ex_launch(x, hi, wi, a, a_bias, out, ho, wo, C, S, f16, ST(stream));
}
int rmem_id_embed_soft(const float* prob, int n_ch, const float* ignore, int H, int W, const float* taps,
                       const float* bias, const float* gamma, const float* beta, int n_cls, void* const* dst_host,
                       const int32_t* ld_host, const int32_t* col0_host, int n_dst, int h, int w, int S, void* stream) {
  return id_embed_soft_launch(prob, n_ch, ignore, H, W, taps, bias, gamma, beta, n_cls, dst_host, ld_host, col0_host,
                              n_dst, h, w, S, ST(stream));
}
int rmem_groupnorm_nhwc_batched(const float* x, const float* pre_bias, const float* gamma, const float* beta, float* y,
                                int P, int C, int groups, int relu, int S, void* workspace, void* stream) {
  return groupnorm_nhwc_batched_launch(x, pre_bias, gamma, beta, y, P, C, groups, relu, S, workspace, ST(stream));
}
int rmem_upsample_add_nhwc_batched(const float* x, int hi, int wi, const float* a, const float* a_bias, float* out,
                                   int ho, int wo, int C, int S, void* stream) {
  return upsample_add_nhwc_batched_launch(x, hi, wi, a, a_bias, out, ho, wo, C, S, ST(stream));
}
int rmem_finish_logits_batched(const float* conv, int64_t sn, int64_t sc, int64_t sy, int64_t sx, const float* bias,
                               float* pred, int C, int Hs, int Ws, const int32_t* n_valid_host, int S, float* out,
                               int Ho, int Wo, void* stream) {
  return finish_logits_batched_launch(conv, sn, sc, sy, sx, bias, pred, C, Hs, Ws, n_valid_host, S, out, Ho, Wo,
                                      ST(stream));
}
int rmem_pack_image_s2d_batched(const void* img, int img_u8_hwc, const float* mean_host, const float* inv_std_host,
                                int H, int W, void* out, int out_f16, int S, void* stream) {
  return pack_image_s2d_batched_launch(img, img_u8_hwc, mean_host, inv_std_host, H, W, out, out_f16, S, ST(stream));
}
int rmem_maxpool3s2_nhwc_batched(const void* x, int H, int W, int C, const float* bias, int relu, void* y, int f16,
                                 int S, void* stream) {
  return maxpool3s2_nhwc_batched_launch(x, H, W, C, bias, relu, y, f16, S, ST(stream));
}
int rmem_argmax_label_batched(const float* logits, int C, int HW, float* label, int S, void* stream) {
  return argmax_label_batched_launch(logits, C, HW, label, S, ST(stream));
}
int rmem_dwconv5_batched(const void* x, int ldx, int xcol0, const float* wt, void* y, int ldy, int ycol0, int h, int w,
                         int C, int S, void* stream) {
  return dwconv5_batched_launch(x, ldx, xcol0, wt, y, ldy, ycol0, h, w, C, S, ST(stream));
}
int rmem_groupnorm_tokens_batched(const float* x_a, int Ca, const float* x_b, const float* gamma, const float* beta,
                                  void* out, int out_f32, int L, int C, int groups, int S, void* workspace,
                                  void* stream) {
  return groupnorm_tokens_batched_launch(x_a, Ca, x_b, gamma, beta, out, out_f32, L, C, groups, S, workspace, ST(stream));
}
int rmem_local_attn_scores_batched(const void* q, const void* k, const float* rel_bias, int ld_rel, float* probs,
                                   void* banded, int h, int w, int S, void* stream) {
  return local_scores_batched_launch(q, k, rel_bias, ld_rel, probs, banded, h, w, S, ST(stream));
}
int rmem_local_attn_aggregate_batched(const void* banded, const void* v, const void* gate, void* out, int h, int w,
                                      int S, void* stream) {
  return local_aggregate_batched_launch(banded, v, gate, out, h, w, S, ST(stream));
}

int rmem_mha_partial(const void* q, int ldq, const void* k, int ldk, int64_t k_slot_stride,
                     const void* v, int ldv, int64_t v_slot_stride, const int32_t* order, int T, int L,
                     int n_slots, int chunks_per_frame, float* part_o, float* part_m, float* part_l,
                     void* stream) {
  return mha_partial_launch(q, ldq, k, ldk, k_slot_stride, v, ldv, v_slot_stride, order, T, L, n_slots,
                            chunks_per_frame, part_o, part_m, part_l, ST(stream));
}
int rmem_mha_combine(const float* part_o, const float* part_m, const float* part_l,
                     const float* tpe_bias, const int32_t* tpe_rows_host, int T, int L,
                     int chunks_per_frame, void* out, int ld_out, int out_col0, float* attn_mass,
                     void* stream) {
  return mha_combine_launch(part_o, part_m, part_l, tpe_bias, tpe_rows_host, T, L, chunks_per_frame, out,
                            ld_out, out_col0, attn_mass, ST(stream));
}
int rmem_tpe_prepare_mha(const void* q, int ldq, const float* cur_pos_emb, const float* mem_pos_emb,
                         void* q_time, float* tpe_bias, int L, void* stream) {
  return tpe_prepare_mha_launch(q, ldq, cur_pos_emb, mem_pos_emb, q_time, tpe_bias, L, ST(stream));
}
int rmem_layernorm_ex(const void* xa, int lda, const void* xb, int ldb, int in_bf16, const float* gamma,
                      const float* beta, void* out, int ld_out, int col_out, const float* post_add,
                      int ld_add, void* out_plus, int ld_plus, int col_plus, float* out_f32, int ld_f32,
                      int col_f32, int L, int C, void* stream) {
  return layernorm_ex_launch(xa, lda, xb, ldb, in_bf16, gamma, beta, out, ld_out, col_out, post_add, ld_add,
                             out_plus, ld_plus, col_plus, out_f32, ld_f32, col_f32, L, C, ST(stream));
}
int rmem_groupnorm_gelu_tokens(const float* x, const float* gamma, const float* beta, void* out, int L,
                               int C, int groups, void* workspace, void* stream) {
  return groupnorm_gelu_tokens_launch(x, gamma, beta, out, L, C, groups, workspace, ST(stream));
}

int rmem_linear(const void* a, int lda, const void* w, int ldw, const float* bias, int M, int N,
                int K, const rmem_gemm_segment* segments_host, int n_segments, void* stream) {
  return gemm_launch(a, lda, w, ldw, bias, M, N, K, segments_host, n_segments, ST(stream));
}
int rmem_linear_ex(const void* a, int lda, const void* w, int ldw, const float* bias, int M, int N, int K,
                   const rmem_gemm_segment* segments_host, int n_segments, int ab_f16, void* stream) {
  return gemm_ex_launch(a, lda, w, ldw, bias, M, N, K, segments_host, n_segments, ab_f16, ST(stream));
}
int rmem_layernorm(const void* x0, const float* gamma0, const float* beta0, void* out0, int ld0,
                   int col0, const void* x1, const float* gamma1, const float* beta1, void* out1,
                   int ld1, int col1, int L, int C, int in_bf16, void* stream) {
  return layernorm_launch(x0, gamma0, beta0, out0, ld0, col0, x1, gamma1, beta1, out1, ld1, col1, L, C,
                          in_bf16, ST(stream));
}
int rmem_dwconv5(const void* x, int ldx, int xcol0, const float* wt, void* y, int ldy, int ycol0,
                 int h, int w, int C, void* stream) {
  return dwconv5_launch(x, ldx, xcol0, wt, y, ldy, ycol0, h, w, C, ST(stream));
}
int rmem_groupnorm_tokens(const float* x_a, int Ca, const float* x_b, const float* gamma,
                          const float* beta, void* out, int out_f32, int L, int C, int groups,
                          void* workspace, void* stream) {
  return groupnorm_tokens_launch(x_a, Ca, x_b, gamma, beta, out, out_f32, L, C, groups, workspace, ST(stream));
}
int rmem_groupnorm_tokens_static(const float* x_a, int Ca, const float* x_b, const float* gamma,
                                 const float* beta, void* out, int out_f32, int L, int C, int groups,
                                 void* stream) {
  return groupnorm_tokens_static_launch(x_a, Ca, x_b, gamma, beta, out, out_f32, L, C, groups, ST(stream));
}
int rmem_tpe_prepare(const void* q, const float* cur_pos_emb, const float* mem_pos_emb,
                     void* q_time, float* tpe_bias, int L, int C, void* stream) {
  return tpe_prepare_launch(q, cur_pos_emb, mem_pos_emb, q_time, tpe_bias, L, C, ST(stream));
}
int64_t rmem_local_banded_elems(int h, int w) { return static_cast<int64_t>(local_banded_elems(h, w)); }
int rmem_local_attn_scores(const void* q, const void* k, const float* rel_bias, int ld_rel,
                           float* probs, void* banded, int h, int w, void* stream) {
  return local_scores_launch(q, k, rel_bias, ld_rel, probs, banded, h, w, ST(stream));
}
int rmem_local_attn(const void* q, const void* k, const void* v, const float* rel_bias, int ld_rel,
                    const void* gate, void* out, int h, int w, void* stream) {
  return local_attn_fused_launch(q, k, v, rel_bias, ld_rel, gate, out, h, w, ST(stream));
}
int rmem_local_attn_aggregate(const void* banded, const void* v, const void* gate, void* out,
                              int h, int w, void* stream) {
  return local_aggregate_launch(banded, v, gate, out, h, w, ST(stream));
}
int rmem_groupnorm_nhwc(const float* x, const float* pre_bias, const float* gamma, const float* beta, float* y,
                        int P, int C, int groups, int relu, void* workspace, void* stream) {
  return groupnorm_nhwc_launch(x, pre_bias, gamma, beta, y, P, C, groups, relu, workspace, ST(stream));
}
int rmem_upsample_add_nhwc(const float* x, int hi, int wi, const float* a, const float* a_bias, float* out,
                           int ho, int wo, int C, void* stream) {
  return upsample_add_nhwc_launch(x, hi, wi, a, a_bias, out, ho, wo, C, ST(stream));
}
int rmem_finish_logits(const float* conv, int64_t sc, int64_t sy, int64_t sx, const float* bias, float* pred,
                       int C, int Hs, int Ws, int n_valid, float* out, int Ho, int Wo, void* stream) {
  return finish_logits_launch(conv, sc, sy, sx, bias, pred, C, Hs, Ws, n_valid, out, Ho, Wo, ST(stream));
}
int rmem_pack_image(const float* img, int H, int W, void* out, void* stream) {
  return pack_image_launch(img, H, W, out, ST(stream));
}
int rmem_pack_image_s2d(const float* img, int H, int W, void* out, void* stream) {
  return pack_image_s2d_launch(img, H, W, out, ST(stream));
}
int rmem_pack_image_s2d_ex(const void* img, int img_u8_hwc, const float* mean_host, const float* inv_std_host,
                           int H, int W, void* out, int out_f16, void* stream) {
  return pack_image_s2d_ex_launch(img, img_u8_hwc, mean_host, inv_std_host, H, W, out, out_f16, ST(stream));
}
int rmem_maxpool3s2_nhwc_ex(const void* x, int H, int W, int C, const float* bias, int relu, void* y, int f16,
                            void* stream) {
  return maxpool3s2_nhwc_ex_launch(x, H, W, C, bias, relu, y, f16, ST(stream));
}
int rmem_maxpool3s2_nhwc(const void* x, int H, int W, int C, const float* bias, int relu, void* y, void* stream) {
  return maxpool3s2_nhwc_launch(x, H, W, C, bias, relu, y, ST(stream));
}
int rmem_argmax_label(const float* logits, int C, int HW, float* label, void* stream) {
  return argmax_label_launch(logits, C, HW, label, ST(stream));
}
int rmem_bank_tick(void* bank_state, void* stream) { return bank_tick_launch(bank_state, ST(stream)); }
int rmem_id_embed(const float* label, int H, int W, const float* prefix, const float* bias,
                  const float* gamma, const float* beta, int n_cls, void* const* dst_host,
                  const int32_t* ld_host, const int32_t* col0_host, int n_dst, int h, int w,
                  void* stream) {
  return id_embed_launch(label, H, W, prefix, bias, gamma, beta, n_cls, dst_host, ld_host, col0_host, n_dst,
                         h, w, ST(stream));
}
int rmem_bank_init(void* bank_state, int frame_step, void* stream) {
  return bank_init_launch(bank_state, frame_step, ST(stream));
}
int rmem_bank_append(void* bank_state, int T, int frame_step, const void* const* src_host,
                     void* const* dst_host, const int64_t* row_bytes_host, int n_tensors,
                     void* stream) {
  return bank_append_launch(bank_state, T, frame_step, src_host, dst_host, row_bytes_host, n_tensors,
                            ST(stream));
}
int rmem_bank_score_evict(void* bank_state, const float* attn_mass, const float* fg, int L,
                          int T_old, int capacity, void* stream) {
  return bank_score_evict_launch(bank_state, attn_mass, fg, L, T_old, capacity, ST(stream));
}
int rmem_foreground_prob(const float* logits, int C, int Hs, int Ws, float* fg, int h, int w,
                         void* stream) {
  return foreground_prob_launch(logits, C, Hs, Ws, fg, h, w, ST(stream));
}

}  // extern "C"
