// Internal declarations shared by the kernels and the C-ABI translation unit (capi.cu).
#pragma once
#include <cuda.h>
#include <cuda_runtime.h>
#include <stdint.h>
#include <atomic>
#include "../../include/rmem_b200.h"

namespace rmem {

// ---- error reporting (thread-local last-error string, never throws across the ABI)
int set_error(int code, const char* msg);
int set_cuda_error(cudaError_t e, const char* where);
int check_launch(const char* kernel_name);

// Programmatic dependent launch: a kernel launched through launch_k may begin (block scheduling, shared-memory
// carve-out, barrier / TMEM set-up, descriptor prefetch, weight loads) while its predecessor on the stream is still
// draining.  The contract every kernel launched this way keeps: pdl_wait() comes before its first access to any global
// memory another kernel of the frame writes or reads (activations, workspaces, outputs) -- only parameters and weights
// may be touched earlier -- and pdl_trigger() follows it, so at most one dependent is ever waiting beside a running grid.
// Opt-in (RMEM_PDL=1; without it launches are fully serialised and griddepcontrol.* are no-ops): on a chain of the
// path's small kernels it saves 0.6 us per launch (9.11 -> 8.49 us, tools/dev_pdl.py), but inside the frame, where
// the next frame's encoder already fills the idle SMs, it measured neutral (0.766-0.774 vs 0.759 ms per frame).
bool pdl_enabled();
#ifdef __CUDACC__
__device__ __forceinline__ void pdl_wait() { asm volatile("griddepcontrol.wait;" ::: "memory"); }
__device__ __forceinline__ void pdl_trigger() { asm volatile("griddepcontrol.launch_dependents;" ::: "memory"); }
template <typename... KArgs, typename... Args>
inline cudaError_t launch_k(void (*kernel)(KArgs...), dim3 grid, dim3 block, size_t smem, cudaStream_t stream,
                            Args&&... args) {
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = grid;
  cfg.blockDim = block;
  cfg.dynamicSmemBytes = smem;
  cfg.stream = stream;
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  attr[0].val.programmaticStreamSerializationAllowed = 1;
  cfg.attrs = attr;
  cfg.numAttrs = pdl_enabled() ? 1 : 0;
  return cudaLaunchKernelEx(&cfg, kernel, static_cast<KArgs>(args)...);
}
#endif

typedef CUresult (*TensorMapEncodeFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*,
                                      const cuuint64_t*, const cuuint64_t*, const cuuint32_t*,
                                      const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                                      CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
TensorMapEncodeFn get_tensor_map_encoder();

// cudaFuncAttributeMaxDynamicSharedMemorySize belongs to a (kernel, device) pair: one record per kernel holds the
// size already granted on every device, so engines on several GPUs of one process (and several host threads)
// all get the opt-in.  Zero-initialised static storage is the valid start state.
constexpr int kMaxDevices = 64;
struct SmemOptIn {
  std::atomic<int> granted[kMaxDevices];
};
int ensure_dynamic_smem(const void* kernel, SmemOptIn* rec, int bytes, const char* what);

// ---- launchers (one per kernel family); all asynchronous on `stream`
int lt_attn_partial_launch(const void* q, const void* kbank, const void* vbank,
                           const int32_t* order_dev, int T, int L, int n_slots,
                           int chunks_per_frame, float* part_o, float* part_m, float* part_l,
                           cudaStream_t stream);
int lt_attn_direct_launch(const void* q, const void* k, const void* v, int L, const void* gate, void* out,
                          cudaStream_t stream);
int lt_attn_combine_launch(const float* part_o, const float* part_m, const float* part_l,
                           const float* tpe_bias, const int32_t* tpe_rows_host, int T, int L,
                           int chunks_per_frame, const void* gate, void* out, float* attn_mass,
                           cudaStream_t stream);

int64_t lt_attn_sk_workspace_bytes(int L, int T, int S, int n_clusters);
int lt_attn_sk_partial_launch(const void* q, const void* kbank, const void* vbank, const int32_t* order_dev,
                              int order_stride, int T, int L, int n_slots, int S, const float* tpe_bias,
                              const int32_t* tpe_rows_host, int record, int n_clusters, void* workspace,
                              int64_t workspace_bytes, cudaStream_t stream);
int lt_attn_sk_combine_launch(int T, int L, int S, const void* gate, void* out, float* attn_mass, int n_clusters,
                              const void* workspace, int64_t workspace_bytes, cudaStream_t stream);

int mha_partial_launch(const void* q, int ldq, const void* k, int ldk, int64_t k_slot_stride, const void* v,
                       int ldv, int64_t v_slot_stride, const int32_t* order_dev, int T, int L, int n_slots,
                       int chunks_per_frame, float* part_o, float* part_m, float* part_l,
                       cudaStream_t stream);
int mha_combine_launch(const float* part_o, const float* part_m, const float* part_l, const float* tpe_bias,
                       const int32_t* tpe_rows_host, int T, int L, int chunks_per_frame, void* out, int ld_out,
                       int out_col0, float* attn_mass, cudaStream_t stream);
int tpe_prepare_mha_launch(const void* q, int ldq, const float* cur, const float* mem, void* q_time, float* bias,
                           int L, cudaStream_t stream);
int layernorm_ex_launch(const void* xa, int lda, const void* xb, int ldb, int in_bf16, const float* gamma,
                        const float* beta, void* out, int ld_o, int col_o, const float* post_add, int ld_add,
                        void* out_plus, int ld_p, int col_p, float* out_f32, int ld_f, int col_f, int L, int C,
                        cudaStream_t stream);
int groupnorm_gelu_tokens_launch(const float* x, const float* gamma, const float* beta, void* out, int L, int C,
                                 int groups, void* workspace, cudaStream_t stream);

int gemm_launch(const void* a, int lda, const void* w, int ldw, const float* bias, int M, int N, int K,
                const rmem_gemm_segment* segs, int n_seg, cudaStream_t stream);
int gemm_ex_launch(const void* a, int lda, const void* w, int ldw, const float* bias, int M, int N, int K,
                   const rmem_gemm_segment* segs, int n_seg, int ab_f16, cudaStream_t stream);
int layernorm_copy_launch(const void* x0, const float* g0, const float* b0, void* o0, int ld0, int c0,
                          const void* x1, const float* g1, const float* b1, void* o1, int ld1, int c1,
                          int L, int C, int in_kind, float* copy_f32, cudaStream_t stream);
int layernorm_launch(const void* x0, const float* g0, const float* b0, void* o0, int ld0, int c0,
                     const void* x1, const float* g1, const float* b1, void* o1, int ld1, int c1,
                     int L, int C, int in_bf16, cudaStream_t stream);
int dwconv5_launch(const void* x, int ldx, int xcol0, const float* wt, void* y, int ldy, int ycol0,
                   int h, int w, int C, cudaStream_t stream);
int groupnorm_tokens_launch(const float* xa, int Ca, const float* xb, const float* gamma,
                            const float* beta, void* out, int out_f32, int L, int C, int groups,
                            void* workspace, cudaStream_t stream);
int dwconv5_batched_launch(const void* x, int ldx, int xcol0, const float* wt, void* y, int ldy, int ycol0,
                           int h, int w, int C, int S, cudaStream_t stream);
int groupnorm_tokens_batched_launch(const float* xa, int Ca, const float* xb, const float* gamma,
                                    const float* beta, void* out, int out_f32, int L, int C, int groups, int S,
                                    void* workspace, cudaStream_t stream);
int groupnorm_tokens_static_launch(const float* xa, int Ca, const float* xb, const float* gamma, const float* beta,
                                   void* out, int out_f32, int L, int C, int groups, cudaStream_t stream);
int tpe_prepare_launch(const void* q, const float* cur, const float* mem, void* q_time, float* bias,
                       int L, int C, cudaStream_t stream);
int foreground_prob_launch(const float* logits, int C, int Hs, int Ws, float* fg, int h, int w,
                           cudaStream_t stream);
size_t local_banded_elems(int h, int w);
int local_scores_launch(const void* q, const void* k, const float* rel, int ld_rel, float* probs, void* banded,
                        int h, int w, cudaStream_t stream);
int local_scores_batched_launch(const void* q, const void* k, const float* rel, int ld_rel, float* probs, void* banded,
                                int h, int w, int S, cudaStream_t stream);
int local_aggregate_batched_launch(const void* banded, const void* v, const void* gate, void* out, int h, int w, int S,
                                   cudaStream_t stream);
int local_attn_fused_launch(const void* q, const void* k, const void* v, const float* rel, int ld_rel,
                            const void* gate, void* out, int h, int w, cudaStream_t stream);
int local_aggregate_launch(const void* banded, const void* v, const void* gate, void* out, int h, int w,
                           cudaStream_t stream);
int groupnorm_nhwc_launch(const float* x, const float* pre_bias, const float* gamma, const float* beta, float* y,
                          int P, int C, int groups, int relu, void* workspace, cudaStream_t stream);
int upsample_add_nhwc_launch(const float* xin, int hi, int wi, const float* a, const float* a_bias, float* out,
                             int ho, int wo, int C, cudaStream_t stream);
int finish_logits_launch(const float* conv, long long sc, long long sy, long long sx, const float* bias, float* pred,
                         int C, int Hs, int Ws, int n_valid, float* out, int Ho, int Wo, cudaStream_t stream);
int pack_image_launch(const float* img, int H, int W, void* out, cudaStream_t stream);
int pack_image_s2d_launch(const float* img, int H, int W, void* out, cudaStream_t stream);
int pack_image_s2d_ex_launch(const void* img, int img_u8_hwc, const float* mean_host, const float* inv_std_host,
                             int H, int W, void* out, int out_f16, cudaStream_t stream);
int maxpool3s2_nhwc_ex_launch(const void* x, int H, int W, int C, const float* bias, int relu, void* y, int f16,
                              cudaStream_t stream);
int maxpool3s2_nhwc_launch(const void* x, int H, int W, int C, const float* bias, int relu, void* y,
                           cudaStream_t stream);
int argmax_label_launch(const float* logits, int C, int HW, float* label, cudaStream_t stream);
int bank_tick_launch(void* state, cudaStream_t stream);
int id_embed_launch(const float* label, int H, int W, const float* prefix, const float* bias,
                    const float* gamma, const float* beta, int n_cls, void* const* dst, const int* ld,
                    const int* col0, int n_dst, int h, int w, cudaStream_t stream);
int bank_init_launch(void* state, int frame_step, cudaStream_t stream);
int bank_tick_batched_launch(void* state, int S, cudaStream_t stream);
int id_embed_soft_launch(const float* prob, int n_ch, const float* ignore, int H, int W, const float* taps,
                         const float* bias, const float* gamma, const float* beta, int n_cls, void* const* dst,
                         const int* ld, const int* col0, int n_dst, int h, int w, int S, cudaStream_t stream);
int id_embed_batched_launch(const float* label, int H, int W, const float* prefix, const float* bias,
                            const float* gamma, const float* beta, int n_cls, void* const* dst, const int* ld,
                            const int* col0, int n_dst, int h, int w, int S, cudaStream_t stream);
int groupnorm_nhwc_batched_launch(const float* x, const float* pre_bias, const float* gamma, const float* beta,
                                  float* y, int P, int C, int groups, int relu, int S, void* workspace,
                                  cudaStream_t stream);
int upsample_add_nhwc_batched_launch(const float* xin, int hi, int wi, const float* a, const float* a_bias, float* out,
                                     int ho, int wo, int C, int S, cudaStream_t stream);
int finish_logits_batched_launch(const float* conv, long long sn, long long sc, long long sy, long long sx,
                                 const float* bias, float* pred, int C, int Hs, int Ws, const int* n_valid_host, int S,
                                 float* out, int Ho, int Wo, cudaStream_t stream);
int pack_image_s2d_batched_launch(const void* img, int img_u8_hwc, const float* mean_host, const float* inv_std_host,
                                  int H, int W, void* out, int out_f16, int S, cudaStream_t stream);
int maxpool3s2_nhwc_batched_launch(const void* x, int H, int W, int C, const float* bias, int relu, void* y, int f16,
                                   int S, cudaStream_t stream);
int gn_upsample_add_nhwc_launch(const void* x, const float* pre_bias, const float* gamma, const float* beta, int hi, int wi,
                                int C, int groups, const float* a, const float* a_bias, void* out, int ho, int wo, int S,
                                int f16, void* workspace, cudaStream_t stream);
int groupnorm_nhwc_ex_launch(const void* x, const float* pre_bias, const float* gamma, const float* beta, void* y, int P,
                             int C, int groups, int relu, int S, int f16, void* workspace, cudaStream_t stream);
int upsample_add_nhwc_ex_launch(const void* xin, int hi, int wi, const float* a, const float* a_bias, void* out, int ho,
                                int wo, int C, int S, int f16, cudaStream_t stream);
int logits_head_batched_launch(const void* x, int x_f16, const float* pre_bias, const float* gamma, const float* beta, int Hs,
                               int Ws, int C, int groups, const float* w_out, const float* b_out, int Co,
                               const int* n_valid_host, int S, float* pred, float* out, int Ho, int Wo, float* label,
                               void* workspace, cudaStream_t stream);
int argmax_label_batched_launch(const float* logits, int C, int HW, float* label, int S, cudaStream_t stream);
int bank_append_launch(void* state, int T, int frame_step, const void* const* src, void* const* dst,
                       const int64_t* row_bytes, int n_tensors, cudaStream_t stream);
int bank_score_evict_launch(void* state, const float* attn_mass, const float* fg, int L, int T_old,
                            int capacity, cudaStream_t stream);


#ifdef __CUDACC__
// GroupNorm statistics workspace: 2 * groups fp64 accumulators followed by one 8-byte arrival counter.  The caller
// provides it zero-filled once; every apply kernel CTA calls this after its last read and the last CTA to arrive
// re-zeroes accumulators and counter, so no memset node sits in front of each GroupNorm (a memset node costs
// ~7 us inside a CUDA graph -- more than the statistics kernel).
// mean / rstd of every group from the fp64 sums, once per CTA into shared memory (instead of fp64 divisions in
// every thread, or -- in grid-stride loops -- in every iteration)
__device__ __forceinline__ void gn_group_stats(const double* ws, int groups, double cnt, float* mean_sh, float* rstd_sh) {
  if (static_cast<int>(threadIdx.x) < groups) {
    const double m = ws[2 * threadIdx.x] / cnt;
    const double var = ws[2 * threadIdx.x + 1] / cnt - m * m;
    mean_sh[threadIdx.x] = static_cast<float>(m);
    rstd_sh[threadIdx.x] = rsqrtf(static_cast<float>(var) + 1e-5f);
  }
  __syncthreads();
}
__device__ __forceinline__ void gn_ws_release(double* ws, int n_stats) {
  __syncthreads();
  if (threadIdx.x == 0) {
    unsigned int* counter = reinterpret_cast<unsigned int*>(ws + n_stats);
    __threadfence();
    if (atomicAdd(counter, 1u) == gridDim.x - 1) {
      for (int i = 0; i < n_stats; ++i) ws[i] = 0.0;
      *counter = 0u;
      __threadfence();
    }
  }
}
#endif

}  // namespace rmem
