// Normalisation / element-wise kernels of the propagation block (HBM/L2-bound, vectorised,
// coalesced): LayerNorm (dual stream, bf16 out at a column offset), 5x5 depth-wise conv over the
// token grid, token GroupNorm, TPE query preparation, foreground probability for the eviction rule.
#include <cuda_bf16.h>
#include <mutex>
#include "sm100_ptx.cuh"
#include <cuda_fp16.h>
#include "rmem_internal.h"

namespace rmem {

__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}

// ------------------------------------------------------------------------------------------------
// LayerNorm, C = 256: one warp per row, 8 channels per lane. Two streams per launch (blockIdx.y).
struct LnStream {
  const void* x;
  const float* gamma;
  const float* beta;
  __nv_bfloat16* out;
  int ld, col0;
};
struct LnParams {
  LnStream s[2];
  int L, in_bf16;      // 0: f32 input, 1: bf16, 2: IEEE fp16
  float* copy_f32;     // optional: stream 0's input rows as f32 [L][256] (the residual stream the block adds into)
};

__global__ void __launch_bounds__(256) layernorm256_kernel(const LnParams p) {
  const LnStream st = p.s[blockIdx.y];
  const int row = blockIdx.x * 8 + (threadIdx.x >> 5);
  const int lane = threadIdx.x & 31;
  pdl_wait();
  pdl_trigger();
  if (row >= p.L) return;
  float v[8];
  if (p.in_bf16 == 2) {
    const uint4 raw = *reinterpret_cast<const uint4*>(static_cast<const __half*>(st.x) +
                                                      static_cast<size_t>(row) * 256 + lane * 8);
    const __half2* h = reinterpret_cast<const __half2*>(&raw);
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      const float2 f = __half22float2(h[i]);
      v[2 * i] = f.x;
      v[2 * i + 1] = f.y;
    }
  } else if (p.in_bf16) {
    const uint4 raw = *reinterpret_cast<const uint4*>(static_cast<const __nv_bfloat16*>(st.x) +
                                                      static_cast<size_t>(row) * 256 + lane * 8);
    const __nv_bfloat162* h = reinterpret_cast<const __nv_bfloat162*>(&raw);
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      v[2 * i] = __bfloat162float(h[i].x);
      v[2 * i + 1] = __bfloat162float(h[i].y);
    }
  } else {
    const float4* px = reinterpret_cast<const float4*>(static_cast<const float*>(st.x) +
                                                       static_cast<size_t>(row) * 256 + lane * 8);
    const float4 a = px[0], b = px[1];
    v[0] = a.x; v[1] = a.y; v[2] = a.z; v[3] = a.w; v[4] = b.x; v[5] = b.y; v[6] = b.z; v[7] = b.w;
  }
  if (p.copy_f32 && blockIdx.y == 0) {
    float4* pc = reinterpret_cast<float4*>(p.copy_f32 + static_cast<size_t>(row) * 256 + lane * 8);
    pc[0] = make_float4(v[0], v[1], v[2], v[3]);
    pc[1] = make_float4(v[4], v[5], v[6], v[7]);
  }
  float s = 0.f;
#pragma unroll
  for (int i = 0; i < 8; ++i) s += v[i];
  const float mean = warp_sum(s) * (1.f / 256.f);
  float q = 0.f;
#pragma unroll
  for (int i = 0; i < 8; ++i) { const float d = v[i] - mean; q += d * d; }
  const float rstd = rsqrtf(warp_sum(q) * (1.f / 256.f) + 1e-5f);
  const float4 g0 = *reinterpret_cast<const float4*>(st.gamma + lane * 8);
  const float4 g1 = *reinterpret_cast<const float4*>(st.gamma + lane * 8 + 4);
  const float4 b0 = *reinterpret_cast<const float4*>(st.beta + lane * 8);
  const float4 b1 = *reinterpret_cast<const float4*>(st.beta + lane * 8 + 4);
  const float g[8] = {g0.x, g0.y, g0.z, g0.w, g1.x, g1.y, g1.z, g1.w};
  const float b[8] = {b0.x, b0.y, b0.z, b0.w, b1.x, b1.y, b1.z, b1.w};
  uint4 pk;
  __nv_bfloat162* o = reinterpret_cast<__nv_bfloat162*>(&pk);
#pragma unroll
  for (int i = 0; i < 4; ++i)
    o[i] = __floats2bfloat162_rn((v[2 * i] - mean) * rstd * g[2 * i] + b[2 * i],
                                 (v[2 * i + 1] - mean) * rstd * g[2 * i + 1] + b[2 * i + 1]);
  *reinterpret_cast<uint4*>(st.out + static_cast<size_t>(row) * st.ld + st.col0 + lane * 8) = pk;
}

int layernorm_launch(const void* x0, const float* g0, const float* b0, void* o0, int ld0, int c0,
                     const void* x1, const float* g1, const float* b1, void* o1, int ld1, int c1,
                     int L, int C, int in_bf16, cudaStream_t stream) {
  return layernorm_copy_launch(x0, g0, b0, o0, ld0, c0, x1, g1, b1, o1, ld1, c1, L, C, in_bf16, nullptr, stream);
}

int layernorm_copy_launch(const void* x0, const float* g0, const float* b0, void* o0, int ld0, int c0,
                          const void* x1, const float* g1, const float* b1, void* o1, int ld1, int c1,
                          int L, int C, int in_kind, float* copy_f32, cudaStream_t stream) {
  const int in_bf16 = in_kind;
  if (in_kind < 0 || in_kind > 2) return set_error(RMEM_ERR_ARG, "layernorm: input kind must be 0 (f32), 1 (bf16), 2 (fp16)");
  if (C != 256 || L < 1 || (ld0 % 8) || (c0 % 8) || (x1 && ((ld1 % 8) || (c1 % 8))))
    return set_error(RMEM_ERR_ARG, "layernorm: C must be 256 and ld/col offsets multiples of 8");
  LnParams p;
  p.L = L; p.in_bf16 = in_bf16; p.copy_f32 = copy_f32;
  p.s[0] = {x0, g0, b0, static_cast<__nv_bfloat16*>(o0), ld0, c0};
  p.s[1] = {x1, g1, b1, static_cast<__nv_bfloat16*>(o1), ld1, c1};
  dim3 grid((L + 7) / 8, x1 ? 2 : 1);
  launch_k(layernorm256_kernel, grid, 256, 0, stream, p);
  return check_launch("layernorm256_kernel");
}

// ------------------------------------------------------------------------------------------------
// 5x5 depth-wise conv over the (h, w) token grid, channels-last bf16, zero padding 2.
// CTA = (band of kDwRows token rows, 64 channels): one TMA box brings the band plus its 2-pixel halo into
// shared memory -- rows / columns outside the image are zero-filled by the TMA unit, which IS the padding --
// so the global-memory latency is paid once per CTA.  Thread = (channel pair, group of columns): the 25
// weights of its two channels live in registers and a 5x5 window slides along each row, so an output costs
// five 4-byte shared-memory loads (conflict-free: consecutive lanes = consecutive channel pairs) and 25
// packed f32x2 FMAs.  Weights tap-major [25][C].
constexpr int kDwRows = 2;          // output rows per CTA
constexpr int kDwCh = 64;           // channels per CTA (128-byte rows in shared memory)
constexpr int kDwThreads = 256;     // 32 channel pairs x 8 column groups

__device__ __forceinline__ void fma2(float2& d, const float2 a, const float2 b) {
  asm("{\n\t.reg .b64 ra, rb, rc;\n\t"
      "mov.b64 ra, {%2, %3};\n\tmov.b64 rb, {%4, %5};\n\tmov.b64 rc, {%0, %1};\n\t"
      "fma.rn.f32x2 rc, ra, rb, rc;\n\t"
      "mov.b64 {%0, %1}, rc;\n\t}\n"
      : "+f"(d.x), "+f"(d.y)
      : "f"(a.x), "f"(a.y), "f"(b.x), "f"(b.y));
}

__global__ void __launch_bounds__(kDwThreads)
dwconv5_kernel(const __grid_constant__ CUtensorMap tm_x, const float* __restrict__ wt,
               __nv_bfloat16* __restrict__ y, int ldy, int ycol0, int h, int w, int C, int box_w) {
  // blockIdx.z = sequence of a batch of independent token grids stacked as [S][h*w] rows
  y += static_cast<size_t>(blockIdx.z) * h * w * ldy;
  extern __shared__ __align__(128) uint8_t dw_smem[];
  uint64_t* bar = reinterpret_cast<uint64_t*>(dw_smem);
  const uint8_t* tile = dw_smem + 128;                         // [kDwRows + 4][box_w][64 ch] bf16
  const int y0 = blockIdx.x * kDwRows;
  const int c0 = blockIdx.y * kDwCh;
  if (threadIdx.x == 0) {
    sm100::tma_prefetch_desc(&tm_x);
    sm100::mbar_init(bar, 1);
    sm100::mbar_fence_init();
  }
  const int cp = threadIdx.x & 31;              // channel pair inside the CTA's 64 channels
  const int grp = threadIdx.x >> 5;             // column group
  const int c = c0 + 2 * cp;
  float2 wgt[25];
#pragma unroll
  for (int t = 0; t < 25; ++t) wgt[t] = *reinterpret_cast<const float2*>(wt + t * C + c);
  pdl_wait();                                   // weights above; activations from here on
  pdl_trigger();
  if (threadIdx.x == 0) {
    sm100::mbar_expect_tx(bar, static_cast<uint32_t>((kDwRows + 4) * box_w * kDwCh * 2));
    sm100::tma_load_4d(dw_smem + 128, &tm_x, bar, c0, -2, y0 - 2, static_cast<int>(blockIdx.z));
  }
  const int per = (w + 7) / 8;                  // columns per group
  const int xa = grp * per, xb = min(xa + per, w);
  __syncthreads();                              // barrier init visible
  sm100::mbar_wait(bar, 0);
  auto ld = [&](int ry, int sx) -> float2 {     // ry: tile row (0 = y0 - 2), sx: tile column (0 = x = -2)
    const uint32_t raw = *reinterpret_cast<const uint32_t*>(tile + (static_cast<size_t>(ry) * box_w + sx) * 128 + cp * 4);
    return make_float2(__uint_as_float(raw << 16), __uint_as_float(raw & 0xffff0000u));
  };
#pragma unroll 1
  for (int r = 0; r < kDwRows; ++r) {
    const int py = y0 + r;
    if (py >= h || xa >= xb) break;
    // col[slot][dy]: window column; output j of a chunk of five uses slots (j + k) % 5 for tap column k and
    // loads its new rightmost column into slot (j + 4) % 5 -- static indices, so the window never moves.
    float2 col[5][5];
#pragma unroll
    for (int k = 0; k < 4; ++k) {
#pragma unroll
      for (int dy = 0; dy < 5; ++dy) col[k][dy] = ld(r + dy, xa + k);
    }
#pragma unroll 1
    for (int xc = xa; xc < xb; xc += 5) {
#pragma unroll
      for (int j = 0; j < 5; ++j) {
        const int xx = xc + j;
        if (xx < xb) {
#pragma unroll
          for (int dy = 0; dy < 5; ++dy) col[(j + 4) % 5][dy] = ld(r + dy, xx + 4);
          float2 acc = make_float2(0.f, 0.f);
#pragma unroll
          for (int dy = 0; dy < 5; ++dy) {
#pragma unroll
            for (int k = 0; k < 5; ++k) fma2(acc, col[(j + k) % 5][dy], wgt[dy * 5 + k]);
          }
          const __nv_bfloat162 o = __floats2bfloat162_rn(acc.x, acc.y);
          *reinterpret_cast<__nv_bfloat162*>(y + static_cast<size_t>(py * w + xx) * ldy + ycol0 + c) = o;
        }
      }
    }
  }
}

int dwconv5_launch(const void* x, int ldx, int xcol0, const float* wt, void* y, int ldy, int ycol0,
                   int h, int w, int C, cudaStream_t stream) {
  return dwconv5_batched_launch(x, ldx, xcol0, wt, y, ldy, ycol0, h, w, C, 1, stream);
}

int dwconv5_batched_launch(const void* x, int ldx, int xcol0, const float* wt, void* y, int ldy, int ycol0,
                           int h, int w, int C, int S, cudaStream_t stream) {
  if ((C % kDwCh) || (ldx % 8) || (ldy % 8) || (xcol0 % 8) || (ycol0 % 8) || h < 1 || w < 1 || w + 4 > 256 || S < 1 ||
      S > 65535)
    return set_error(RMEM_ERR_ARG, "dwconv5: C must be a multiple of 64, ld / col offsets multiples of 8, w <= 252");
  auto fn = get_tensor_map_encoder();
  if (!fn) return set_error(RMEM_ERR_DRIVER, "cuTensorMapEncodeTiled entry point not found");
  const int box_w = w + 4;
  CUtensorMap tm_x;
  {
    const uint64_t dims[4] = {static_cast<uint64_t>(C), static_cast<uint64_t>(w), static_cast<uint64_t>(h),
                              static_cast<uint64_t>(S)};
    const uint64_t strides[3] = {static_cast<uint64_t>(ldx) * 2, static_cast<uint64_t>(w) * ldx * 2,
                                 static_cast<uint64_t>(h) * w * ldx * 2};
    const uint32_t box[4] = {kDwCh, static_cast<uint32_t>(box_w), kDwRows + 4, 1};
    const uint32_t es[4] = {1, 1, 1, 1};
    const void* base = static_cast<const __nv_bfloat16*>(x) + xcol0;
    CUresult r = fn(&tm_x, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 4, const_cast<void*>(base), dims, strides, box, es,
                    CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                    CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_SUCCESS) return set_error(RMEM_ERR_DRIVER, "cuTensorMapEncodeTiled (dwconv5) failed");
  }
  const size_t smem = 128 + static_cast<size_t>(kDwRows + 4) * box_w * kDwCh * 2;
  static SmemOptIn optin;
  if (int r = ensure_dynamic_smem(reinterpret_cast<const void*>(dwconv5_kernel), &optin, static_cast<int>(smem),
                                  "dwconv5: cudaFuncSetAttribute"))
    return r;
  dim3 grid((h + kDwRows - 1) / kDwRows, C / kDwCh, S);
  launch_k(dwconv5_kernel, grid, kDwThreads, smem, stream, tm_x, wt, static_cast<__nv_bfloat16*>(y), ldy, ycol0, h, w,
           C, box_w);
  return check_launch("dwconv5_kernel");
}

// ------------------------------------------------------------------------------------------------
// Token GroupNorm (GroupNorm1D): pass 1 accumulates per-group sum / sum-of-squares in fp64 atomics,
// pass 2 normalises. x = cat(x_a [L][Ca], x_b [L][C-Ca]).
__global__ void __launch_bounds__(256)
gn_stats_kernel(const float* __restrict__ xa, int Ca, const float* __restrict__ xb, int L, int C,
                int groups, int rows_per_cta, double* __restrict__ ws) {
  pdl_wait();        // programmatic dependent launch: see rmem_internal.h
  pdl_trigger();
  // blockIdx.y = sequence: inputs stacked [S][L] rows, one (2 groups + 1)-double workspace per sequence
  xa += static_cast<size_t>(blockIdx.y) * L * Ca;
  if (xb) xb += static_cast<size_t>(blockIdx.y) * L * (C - Ca);
  ws += static_cast<size_t>(blockIdx.y) * (2 * groups + 1);
  // One CTA per SM, rows_per_cta rows each.  A thread owns four consecutive channels (one float4 per row; Ca, C
  // and C / groups are multiples of 4), so C / 4 threads cover a row and 1024 / C rows are in flight per pass.
  // Every CTA ends in 2 * groups fp64 atomics: ~150 CTAs keep that contention below the cost of the reduction.
  const int cpg = C / groups;
  const int tpr = C >> 2;                         // threads per row (<= 256)
  const int c = (threadIdx.x % tpr) << 2;
  const int r_in_pass = threadIdx.x / tpr, rows_per_pass = 256 / tpr;
  const bool active = r_in_pass < rows_per_pass;  // 256 % tpr threads idle when tpr does not divide 256
  const float* src = c < Ca ? xa + c : xb + (c - Ca);
  const int ld = c < Ca ? Ca : C - Ca;
  const int row0 = blockIdx.x * rows_per_cta, row1 = min(row0 + rows_per_cta, L);
  float s = 0.f, q = 0.f;
  if (active) {
#pragma unroll 4
    for (int row = row0 + r_in_pass; row < row1; row += rows_per_pass) {
      const float4 v = *reinterpret_cast<const float4*>(src + static_cast<size_t>(row) * ld);
      s += (v.x + v.y) + (v.z + v.w);
      q += (v.x * v.x + v.y * v.y) + (v.z * v.z + v.w * v.w);
    }
  }
  __shared__ float sh_s[8], sh_q[8];
  if (threadIdx.x < 8) { sh_s[threadIdx.x] = 0.f; sh_q[threadIdx.x] = 0.f; }
  __syncthreads();
  const int g = c / cpg;
  if ((cpg & 127) == 0 && (tpr & 31) == 0) {      // a warp's 128 channels lie in one group
    const float ws_ = warp_sum(s), wq_ = warp_sum(q);
    if ((threadIdx.x & 31) == 0) { atomicAdd(&sh_s[g], ws_); atomicAdd(&sh_q[g], wq_); }
  } else if (active) {
    atomicAdd(&sh_s[g], s); atomicAdd(&sh_q[g], q);
  }
  __syncthreads();
  if (threadIdx.x < groups) {
    atomicAdd(ws + 2 * threadIdx.x, static_cast<double>(sh_s[threadIdx.x]));
    atomicAdd(ws + 2 * threadIdx.x + 1, static_cast<double>(sh_q[threadIdx.x]));
  }
}

__global__ void __launch_bounds__(256)
gn_apply_kernel(const float* __restrict__ xa, int Ca, const float* __restrict__ xb,
                const float* __restrict__ gamma, const float* __restrict__ beta,
                __nv_bfloat16* __restrict__ out, float* __restrict__ out_f32, int L, int C, int groups,
                double* __restrict__ ws, int out_is_f16) {
  pdl_wait();        // programmatic dependent launch: see rmem_internal.h
  pdl_trigger();
  xa += static_cast<size_t>(blockIdx.y) * L * Ca;
  if (xb) xb += static_cast<size_t>(blockIdx.y) * L * (C - Ca);
  ws += static_cast<size_t>(blockIdx.y) * (2 * groups + 1);
  if (out) out += static_cast<size_t>(blockIdx.y) * L * C;
  if (out_f32) out_f32 += static_cast<size_t>(blockIdx.y) * L * C;
  // a thread owns four consecutive channels (one float4 in, four bf16 or f32 out); Ca, C and C / groups are
  // multiples of 4
  const int cpg = C / groups;
  const size_t total4 = static_cast<size_t>(L) * C / 4;
  __shared__ float gmean[32], grstd[32];
  gn_group_stats(ws, groups, static_cast<double>(cpg) * L, gmean, grstd);
  for (size_t i = blockIdx.x * static_cast<size_t>(blockDim.x) + threadIdx.x; i < total4;
       i += static_cast<size_t>(gridDim.x) * blockDim.x) {
    const int row = static_cast<int>((i * 4) / C), c = static_cast<int>(i * 4 - static_cast<size_t>(row) * C);
    const int g = c / cpg;
    const float mean = gmean[g], rstd = grstd[g];
    const float4 v = c < Ca ? *reinterpret_cast<const float4*>(xa + static_cast<size_t>(row) * Ca + c)
                            : *reinterpret_cast<const float4*>(xb + static_cast<size_t>(row) * (C - Ca) + (c - Ca));
    const float4 ga = *reinterpret_cast<const float4*>(gamma + c);
    const float4 be = *reinterpret_cast<const float4*>(beta + c);
    const float4 y = make_float4((v.x - mean) * rstd * ga.x + be.x, (v.y - mean) * rstd * ga.y + be.y,
                                 (v.z - mean) * rstd * ga.z + be.z, (v.w - mean) * rstd * ga.w + be.w);
    if (out_f32) {
      *reinterpret_cast<float4*>(out_f32 + i * 4) = y;
    } else if (out_is_f16) {                   // `out` holds IEEE fp16 (the fp16 decoder reads it as it is)
      const __half2 h0 = __floats2half2_rn(y.x, y.y), h1 = __floats2half2_rn(y.z, y.w);
      uint2 pk;
      pk.x = *reinterpret_cast<const uint32_t*>(&h0);
      pk.y = *reinterpret_cast<const uint32_t*>(&h1);
      *reinterpret_cast<uint2*>(out + i * 4) = pk;
    } else {
      uint2 pk;
      __nv_bfloat162* o = reinterpret_cast<__nv_bfloat162*>(&pk);
      o[0] = __floats2bfloat162_rn(y.x, y.y);
      o[1] = __floats2bfloat162_rn(y.z, y.w);
      *reinterpret_cast<uint2*>(out + i * 4) = pk;
    }
  }
  gn_ws_release(ws, 2 * groups);
}

int groupnorm_tokens_launch(const float* xa, int Ca, const float* xb, const float* gamma,
                            const float* beta, void* out, int out_f32, int L, int C, int groups,
                            void* workspace, cudaStream_t stream) {
  return groupnorm_tokens_batched_launch(xa, Ca, xb, gamma, beta, out, out_f32, L, C, groups, 1, workspace, stream);
}

int groupnorm_tokens_batched_launch(const float* xa, int Ca, const float* xb, const float* gamma,
                                    const float* beta, void* out, int out_f32, int L, int C, int groups, int S,
                                    void* workspace, cudaStream_t stream) {
  if (groups < 1 || groups > 8 || C > 1024 || (C % groups) || ((C / groups) % 32) || Ca > C ||
      (Ca < C && !xb) || (Ca & 3) || (C & 3) || S < 1 || S > 65535)
    return set_error(RMEM_ERR_ARG, "groupnorm_tokens: unsupported C / groups");
  const int per = (148 + S - 1) / S;                               // CTAs per sequence: about one per SM in total
  const int rows = (L + per - 1) / per < 4 ? 4 : (L + per - 1) / per;
  launch_k(gn_stats_kernel, dim3((L + rows - 1) / rows, S), 256, 0, stream, xa, Ca, xb, L, C, groups, rows,
                                                                      static_cast<double*>(workspace));
  if (int r = check_launch("gn_stats_kernel")) return r;
  launch_k(gn_apply_kernel, dim3((296 + S - 1) / S, S), 256, 0, stream, xa, Ca, xb, gamma, beta,
                                           out_f32 == 1 ? nullptr : static_cast<__nv_bfloat16*>(out),
                                           out_f32 == 1 ? static_cast<float*>(out) : nullptr, L, C, groups,
                                           static_cast<double*>(workspace), out_f32 == 2 ? 1 : 0);
  return check_launch("gn_apply_kernel");
}

// ------------------------------------------------------------------------------------------------
// Token GroupNorm without a caller workspace (rmem_groupnorm_tokens_static): the statistics scratch is a
// file-scope device array instead of a pointer the caller allocated.  Every GPU that loads the library holds
// these kGnStaticBytes bytes until the process exits.
//
// In HOST code the name `g_gn_static_ws` is only a host-side stand-in for the symbol, not a device address:
// handing it to a kernel would fault (or, where the GPU can reach host memory, silently use the wrong memory).
// The device address is looked up once per device with cudaGetSymbolAddress, with that device current, and it
// is that pointer the kernels receive.
constexpr int kGnStaticMaxGroups = RMEM_GN_STATIC_MAX_GROUPS;
constexpr size_t kGnStaticBytes = 256;                    // >= 16 * kGnStaticMaxGroups + 8, one 256-byte line
static_assert(16 * kGnStaticMaxGroups + 8 <= kGnStaticBytes, "static GroupNorm scratch too small");
static __device__ __align__(256) unsigned char g_gn_static_ws[kGnStaticBytes];

static void* g_gn_static_ptr[kMaxDevices];                // device address of g_gn_static_ws, per device
static std::mutex g_gn_static_mu;

static int gn_static_scratch(void** ptr) {
  int dev = -1;
  cudaError_t e = cudaGetDevice(&dev);
  if (e != cudaSuccess) return set_cuda_error(e, "groupnorm_tokens_static: cudaGetDevice");
  if (dev < 0 || dev >= kMaxDevices) return set_error(RMEM_ERR_ARG, "groupnorm_tokens_static: device ordinal >= 64");
  std::lock_guard<std::mutex> lock(g_gn_static_mu);
  if (!g_gn_static_ptr[dev]) {
    void* p = nullptr;
    e = cudaGetSymbolAddress(&p, g_gn_static_ws);         // the current device's copy of the array
    if (e != cudaSuccess) return set_cuda_error(e, "groupnorm_tokens_static: cudaGetSymbolAddress");
    if (!p) return set_error(RMEM_ERR_CUDA, "groupnorm_tokens_static: cudaGetSymbolAddress returned NULL");
    g_gn_static_ptr[dev] = p;
  }
  *ptr = g_gn_static_ptr[dev];
  return RMEM_OK;
}

int groupnorm_tokens_static_launch(const float* xa, int Ca, const float* xb, const float* gamma, const float* beta,
                                   void* out, int out_f32, int L, int C, int groups, cudaStream_t stream) {
  if (groups < 1 || groups > kGnStaticMaxGroups)
    return set_error(RMEM_ERR_ARG, "groupnorm_tokens_static: groups exceeds RMEM_GN_STATIC_MAX_GROUPS (8): the "
                                   "built-in scratch holds 16 * groups + 8 <= 256 bytes");
  if (L < 1) return set_error(RMEM_ERR_ARG, "groupnorm_tokens_static: L must be >= 1");
  void* ws = nullptr;
  if (int r = gn_static_scratch(&ws)) return r;
  // the array is shared by every call on this device and keeps the last call's contents: the accumulators and
  // the arrival counter must start at zero (a memset is not an allocation; it is ordered on `stream`)
  cudaError_t e = cudaMemsetAsync(ws, 0, kGnStaticBytes, stream);
  if (e != cudaSuccess) return set_cuda_error(e, "groupnorm_tokens_static: cudaMemsetAsync");
  return groupnorm_tokens_launch(xa, Ca, xb, gamma, beta, out, out_f32, L, C, groups, ws, stream);
}

// ------------------------------------------------------------------------------------------------
// TPE preparation: q_time = bf16(q + cur), bias[q][r] = log2e/sqrt(C) * ((q + cur) . mem[r]).  C = 128.
__global__ void __launch_bounds__(256)
tpe_prepare_kernel(const __nv_bfloat16* __restrict__ q, const float* __restrict__ cur,
                   const float* __restrict__ mem, __nv_bfloat16* __restrict__ q_time,
                   float* __restrict__ bias, int L, float scale) {
  pdl_wait();        // programmatic dependent launch: see rmem_internal.h
  pdl_trigger();
  const int row = blockIdx.x * 8 + (threadIdx.x >> 5);
  const int lane = threadIdx.x & 31;
  if (row >= L) return;
  const uint2 raw = *reinterpret_cast<const uint2*>(q + static_cast<size_t>(row) * 128 + lane * 4);
  const __nv_bfloat162* hv = reinterpret_cast<const __nv_bfloat162*>(&raw);
  const float4 c4 = *reinterpret_cast<const float4*>(cur + lane * 4);
  float v[4] = {__bfloat162float(hv[0].x) + c4.x, __bfloat162float(hv[0].y) + c4.y,
                __bfloat162float(hv[1].x) + c4.z, __bfloat162float(hv[1].y) + c4.w};
  uint2 pk;
  __nv_bfloat162* o = reinterpret_cast<__nv_bfloat162*>(&pk);
  o[0] = __floats2bfloat162_rn(v[0], v[1]);
  o[1] = __floats2bfloat162_rn(v[2], v[3]);
  *reinterpret_cast<uint2*>(q_time + static_cast<size_t>(row) * 128 + lane * 4) = pk;
  // the bias is a rank-1 term kept in f32: it uses the un-rounded q' (the reference adds the embedding to K in
  // fp32, transformer.py:1142-1172), so it carries only the bf16 rounding q already has
#ifdef RMEM_TPE_BIAS_FROM_ROUNDED_Q
  const float r0 = __bfloat162float(o[0].x), r1 = __bfloat162float(o[0].y);
  const float r2 = __bfloat162float(o[1].x), r3 = __bfloat162float(o[1].y);
#else
  const float r0 = v[0], r1 = v[1], r2 = v[2], r3 = v[3];
#endif
#pragma unroll
  for (int r = 0; r < RMEM_TPE_ROWS; ++r) {
    const float4 m4 = *reinterpret_cast<const float4*>(mem + r * 128 + lane * 4);
    const float d = warp_sum(r0 * m4.x + r1 * m4.y + r2 * m4.z + r3 * m4.w);
    if (lane == 0) bias[row * RMEM_TPE_ROWS + r] = d * scale;
  }
}

int tpe_prepare_launch(const void* q, const float* cur, const float* mem, void* q_time, float* bias,
                       int L, int C, cudaStream_t stream) {
  if (C != 128) return set_error(RMEM_ERR_ARG, "tpe_prepare: C must be 128");
  launch_k(tpe_prepare_kernel, (L + 7) / 8, 256, 0, stream, 
      static_cast<const __nv_bfloat16*>(q), cur, mem, static_cast<__nv_bfloat16*>(q_time), bias, L,
      1.4426950408889634f / sqrtf(128.f));
  return check_launch("tpe_prepare_kernel");
}

// ------------------------------------------------------------------------------------------------
// fg[q] = 1 - softmax_c(bilinear(align_corners=True) logits)[0]  (aot_engine.py:355-362)
__global__ void __launch_bounds__(128)
foreground_prob_kernel(const float* __restrict__ logits, int C, int Hs, int Ws, float* __restrict__ fg,
                       int h, int w) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= h * w) return;
  const int y = i / w, x = i - y * w;
  const float sy = h > 1 ? static_cast<float>(Hs - 1) / static_cast<float>(h - 1) : 0.f;
  const float sx = w > 1 ? static_cast<float>(Ws - 1) / static_cast<float>(w - 1) : 0.f;
  const float fy = sy * y, fx = sx * x;
  const int y0 = static_cast<int>(fy), x0 = static_cast<int>(fx);
  const int y1 = min(y0 + 1, Hs - 1), x1 = min(x0 + 1, Ws - 1);
  const float ly = fy - y0, lx = fx - x0;
  float mx = -INFINITY, v0 = 0.f;
  float vals[16];
  for (int c = 0; c < C; ++c) {
    const float* pl = logits + static_cast<size_t>(c) * Hs * Ws;
    const float v = (1.f - ly) * ((1.f - lx) * pl[y0 * Ws + x0] + lx * pl[y0 * Ws + x1]) +
                    ly * ((1.f - lx) * pl[y1 * Ws + x0] + lx * pl[y1 * Ws + x1]);
    vals[c] = v;
    mx = fmaxf(mx, v);
    if (c == 0) v0 = v;
  }
  float den = 0.f;
  for (int c = 0; c < C; ++c) den += expf(vals[c] - mx);
  fg[i] = 1.f - expf(v0 - mx) / den;
}

int foreground_prob_launch(const float* logits, int C, int Hs, int Ws, float* fg, int h, int w,
                           cudaStream_t stream) {
  if (C < 1 || C > 16) return set_error(RMEM_ERR_ARG, "foreground_prob: C must be in [1,16]");
  foreground_prob_kernel<<<(h * w + 127) / 128, 128, 0, stream>>>(logits, C, Hs, Ws, fg, h, w);
  return check_launch("foreground_prob_kernel");
}

}  // namespace rmem
