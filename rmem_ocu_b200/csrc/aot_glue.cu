// Normalisation glue of the AOT propagation block (SimplifiedTransformerBlock.forward, reference
// aot_plus/networks/layers/transformer.py:553-692) that the DeAOT kernels of fused.cu do not cover:
//   * LayerNorm(256) of a sum of two streams (norm4(local_K + curr_K), :657-660), with an optional
//     second output LN(x) + pos (q = k = norm1(x) + self_pos, :563-564) and an optional fp32 output
//     (the three decoder_norms of LongShortTermTransformer.forward, :248-259, written straight into the
//     channels-last decoder input);
//   * GroupNorm(32) + GELU over the token matrix (GNActDWConv2d, layers/basic.py:15-35) in front of the
//     5x5 depth-wise conv of the feed-forward branch (:682-687).
// HBM/L2-bound, one pass per output, vectorised 16/32-byte accesses.
#include <cuda_bf16.h>
#include "rmem_internal.h"

namespace rmem {

namespace {
__device__ __forceinline__ float wsum(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}

__device__ __forceinline__ void load8(const void* base, size_t elem_off, int is_bf16, float (&v)[8]) {
  if (is_bf16) {
    const uint4 raw = *reinterpret_cast<const uint4*>(static_cast<const __nv_bfloat16*>(base) + elem_off);
    const __nv_bfloat162* h = reinterpret_cast<const __nv_bfloat162*>(&raw);
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      v[2 * i] = __bfloat162float(h[i].x);
      v[2 * i + 1] = __bfloat162float(h[i].y);
    }
  } else {
    const float4* px = reinterpret_cast<const float4*>(static_cast<const float*>(base) + elem_off);
    const float4 a = px[0], b = px[1];
    v[0] = a.x; v[1] = a.y; v[2] = a.z; v[3] = a.w; v[4] = b.x; v[5] = b.y; v[6] = b.z; v[7] = b.w;
  }
}

__device__ __forceinline__ void store8_bf16(__nv_bfloat16* dst, const float (&y)[8]) {
  uint4 pk;
  __nv_bfloat162* o = reinterpret_cast<__nv_bfloat162*>(&pk);
#pragma unroll
  for (int i = 0; i < 4; ++i) o[i] = __floats2bfloat162_rn(y[2 * i], y[2 * i + 1]);
  *reinterpret_cast<uint4*>(dst) = pk;
}
}  // namespace

struct LnExParams {
  const void* xa; int lda;
  const void* xb; int ldb;          // optional second addend (same dtype as xa)
  int in_bf16;
  const float* gamma; const float* beta;
  __nv_bfloat16* out; int ld_o, col_o;            // optional bf16 output
  const float* post_add; int ld_add;              // optional f32 [L][256] added after the norm ...
  __nv_bfloat16* out_plus; int ld_p, col_p;       // ... into this second bf16 output
  float* out_f32; int ld_f, col_f;                // optional f32 output
  int L;
};

// one warp per row, 8 channels per lane (C = 256)
__global__ void __launch_bounds__(256) layernorm256_ex_kernel(const LnExParams p) {
  const int row = blockIdx.x * 8 + (threadIdx.x >> 5);
  const int lane = threadIdx.x & 31;
  if (row >= p.L) return;
  float v[8];
  load8(p.xa, static_cast<size_t>(row) * p.lda + lane * 8, p.in_bf16, v);
  if (p.xb) {
    float w[8];
    load8(p.xb, static_cast<size_t>(row) * p.ldb + lane * 8, p.in_bf16, w);
#pragma unroll
    for (int i = 0; i < 8; ++i) v[i] += w[i];
  }
  float s = 0.f;
#pragma unroll
  for (int i = 0; i < 8; ++i) s += v[i];
  const float mean = wsum(s) * (1.f / 256.f);
  float q = 0.f;
#pragma unroll
  for (int i = 0; i < 8; ++i) { const float d = v[i] - mean; q += d * d; }
  const float rstd = rsqrtf(wsum(q) * (1.f / 256.f) + 1e-5f);
  float g[8], b[8], y[8];
  load8(p.gamma, lane * 8, 0, g);
  load8(p.beta, lane * 8, 0, b);
#pragma unroll
  for (int i = 0; i < 8; ++i) y[i] = (v[i] - mean) * rstd * g[i] + b[i];
  if (p.out) store8_bf16(p.out + static_cast<size_t>(row) * p.ld_o + p.col_o + lane * 8, y);
  if (p.out_f32) {
    float4* d = reinterpret_cast<float4*>(p.out_f32 + static_cast<size_t>(row) * p.ld_f + p.col_f + lane * 8);
    d[0] = make_float4(y[0], y[1], y[2], y[3]);
    d[1] = make_float4(y[4], y[5], y[6], y[7]);
  }
  if (p.out_plus) {
    float a[8];
    load8(p.post_add, static_cast<size_t>(row) * p.ld_add + lane * 8, 0, a);
#pragma unroll
    for (int i = 0; i < 8; ++i) a[i] += y[i];
    store8_bf16(p.out_plus + static_cast<size_t>(row) * p.ld_p + p.col_p + lane * 8, a);
  }
}

int layernorm_ex_launch(const void* xa, int lda, const void* xb, int ldb, int in_bf16, const float* gamma,
                        const float* beta, void* out, int ld_o, int col_o, const float* post_add, int ld_add,
                        void* out_plus, int ld_p, int col_p, float* out_f32, int ld_f, int col_f, int L, int C,
                        cudaStream_t stream) {
  if (C != 256 || L < 1 || (lda % 8) || (xb && (ldb % 8)) || (out && ((ld_o % 8) || (col_o % 8))) ||
      (out_plus && (!post_add || (ld_p % 8) || (col_p % 8) || (ld_add % 8))) ||
      (out_f32 && ((ld_f % 4) || (col_f % 4))) || (!out && !out_plus && !out_f32))
    return set_error(RMEM_ERR_ARG, "layernorm_ex: C must be 256, ld / col offsets multiples of 8, one output");
  LnExParams p;
  p.xa = xa; p.lda = lda; p.xb = xb; p.ldb = ldb; p.in_bf16 = in_bf16;
  p.gamma = gamma; p.beta = beta;
  p.out = static_cast<__nv_bfloat16*>(out); p.ld_o = ld_o; p.col_o = col_o;
  p.post_add = post_add; p.ld_add = ld_add;
  p.out_plus = static_cast<__nv_bfloat16*>(out_plus); p.ld_p = ld_p; p.col_p = col_p;
  p.out_f32 = out_f32; p.ld_f = ld_f; p.col_f = col_f;
  p.L = L;
  layernorm256_ex_kernel<<<(L + 7) / 8, 256, 0, stream>>>(p);
  return check_launch("layernorm256_ex_kernel");
}

// ------------------------------------------------------------------------------------------------
// GroupNorm(groups) + GELU over x f32 [L][C] viewed as [1, C, h, w]: statistics per group over
// (C/groups channels x L tokens).  C/groups == 32, C <= 1024: thread t of a 256-thread block covers
// channels t + 256k, so one warp's 32 lanes are exactly one group.
__global__ void __launch_bounds__(256)
gn32_stats_kernel(const float* __restrict__ x, int L, int C, double* __restrict__ ws) {
  float s[4] = {0.f, 0.f, 0.f, 0.f}, q[4] = {0.f, 0.f, 0.f, 0.f};
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  for (int r = 0; r < 8; ++r) {
    const int row = blockIdx.x * 8 + r;
    if (row >= L) break;
    int k = 0;
    for (int c = threadIdx.x; c < C; c += 256, ++k) {
      const float v = x[static_cast<size_t>(row) * C + c];
      s[k] += v; q[k] += v * v;
    }
  }
  int k = 0;
  for (int c0 = 0; c0 < C; c0 += 256, ++k) {
    const int g = (c0 >> 5) + warp;
    const float ts = wsum(s[k]), tq = wsum(q[k]);
    if (lane == 0) {
      atomicAdd(ws + 2 * g, static_cast<double>(ts));
      atomicAdd(ws + 2 * g + 1, static_cast<double>(tq));
    }
  }
}

__global__ void __launch_bounds__(256)
gn32_gelu_apply_kernel(const float* __restrict__ x, const float* __restrict__ gamma,
                       const float* __restrict__ beta, __nv_bfloat16* __restrict__ out, int L, int C,
                       double* __restrict__ ws) {
  const size_t total4 = static_cast<size_t>(L) * C / 4;
  __shared__ float gmean[32], grstd[32];
  gn_group_stats(ws, C >> 5, 32.0 * L, gmean, grstd);
  for (size_t i = blockIdx.x * static_cast<size_t>(blockDim.x) + threadIdx.x; i < total4;
       i += static_cast<size_t>(gridDim.x) * blockDim.x) {
    const int c = static_cast<int>((i * 4) % C);
    const int g = c >> 5;
    const float mean = gmean[g], rstd = grstd[g];
    const float4 v = *reinterpret_cast<const float4*>(x + i * 4);
    const float4 ga = *reinterpret_cast<const float4*>(gamma + c);
    const float4 be = *reinterpret_cast<const float4*>(beta + c);
    float y[4] = {(v.x - mean) * rstd * ga.x + be.x, (v.y - mean) * rstd * ga.y + be.y,
                  (v.z - mean) * rstd * ga.z + be.z, (v.w - mean) * rstd * ga.w + be.w};
#pragma unroll
    for (int j = 0; j < 4; ++j) y[j] = 0.5f * y[j] * (1.f + erff(y[j] * 0.70710678118654752f));   // exact GELU
    uint2 pk;
    __nv_bfloat162* o = reinterpret_cast<__nv_bfloat162*>(&pk);
    o[0] = __floats2bfloat162_rn(y[0], y[1]);
    o[1] = __floats2bfloat162_rn(y[2], y[3]);
    *reinterpret_cast<uint2*>(out + i * 4) = pk;
  }
  gn_ws_release(ws, 2 * (C >> 5));
}

int groupnorm_gelu_tokens_launch(const float* x, const float* gamma, const float* beta, void* out, int L, int C,
                                 int groups, void* workspace, cudaStream_t stream) {
  if (groups < 1 || C > 1024 || (C % 256) || C != groups * 32 || L < 1)
    return set_error(RMEM_ERR_ARG, "groupnorm_gelu_tokens: needs C = 32 * groups, C a multiple of 256 <= 1024");
  gn32_stats_kernel<<<(L + 7) / 8, 256, 0, stream>>>(x, L, C, static_cast<double*>(workspace));
  if (int r = check_launch("gn32_stats_kernel")) return r;
  gn32_gelu_apply_kernel<<<592, 256, 0, stream>>>(x, gamma, beta, static_cast<__nv_bfloat16*>(out), L, C,
                                                   static_cast<double*>(workspace));
  return check_launch("gn32_gelu_apply_kernel");
}

}  // namespace rmem
