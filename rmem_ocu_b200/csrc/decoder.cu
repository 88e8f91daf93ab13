// Normalisation / resampling kernels of the FPN decoding head (reference
// aot_plus/networks/decoders/fpn.py:36-68, ConvGN = Conv2d + GroupNorm(8), layers/basic.py:60-70).
// The convolutions themselves stay library (cuDNN) calls; what PyTorch issues around them for a
// channels-last batch-1 tensor (a GroupNorm reduction on 8 CTAs, a generic bilinear kernel) is replaced
// by two bandwidth-bound kernels over the [H*W][C] view:
//   gn_nhwc_stats / gn_nhwc_apply : GroupNorm over (C/groups channels x H*W pixels) + optional ReLU
//   upsample_add_nhwc             : out = a + bilinear_align_corners(x)   (adapter(shortcut) + upsampled x)
#include <cuda_bf16.h>
#include <cuda_fp16.h>
#include "rmem_internal.h"

// CTAs per SM the grid-stride GroupNorm consumers are capped at (4 measured slower at 120 x 216 x 128: 16.6 vs 13.8 us)
#ifndef RMEM_GN_CTAS_PER_SM
#define RMEM_GN_CTAS_PER_SM 8
#endif
namespace rmem {

// Four consecutive channels of a channels-last activation, fp32 or fp16 storage (the decoder runs in either).
template <typename T> __device__ __forceinline__ float4 ld4(const T* p);
template <> __device__ __forceinline__ float4 ld4<float>(const float* p) { return *reinterpret_cast<const float4*>(p); }
template <> __device__ __forceinline__ float4 ld4<__half>(const __half* p) {
  const uint2 r = *reinterpret_cast<const uint2*>(p);
  const float2 a = __half22float2(*reinterpret_cast<const __half2*>(&r.x));
  const float2 b = __half22float2(*reinterpret_cast<const __half2*>(&r.y));
  return make_float4(a.x, a.y, b.x, b.y);
}
template <typename T> __device__ __forceinline__ void st4(T* p, float4 v);
template <> __device__ __forceinline__ void st4<float>(float* p, float4 v) { *reinterpret_cast<float4*>(p) = v; }
template <> __device__ __forceinline__ void st4<__half>(__half* p, float4 v) {
  const __half2 a = __floats2half2_rn(v.x, v.y), b = __floats2half2_rn(v.z, v.w);
  uint2 r;
  r.x = *reinterpret_cast<const uint32_t*>(&a);
  r.y = *reinterpret_cast<const uint32_t*>(&b);
  *reinterpret_cast<uint2*>(p) = r;
}


// ------------------------------------------------------------------------------------------------
// stats: a thread owns four consecutive channels (one float4, always inside one group since cpg % 4 == 0) of
// every (1024 / C)-th pixel of the CTA's pixel range; partial sums are reduced over the lanes of a group with
// shuffles, over the CTA in shared memory, and added to the fp64 workspace with one atomic pair per
// (CTA, group).  The range per CTA is sized so the grid is ~4 CTAs per SM whatever the resolution.
template <typename T>
__global__ void __launch_bounds__(256)
gn_nhwc_stats_kernel(const T* __restrict__ x, const float* __restrict__ pre_bias, int P, int C, int cpg,
                     int rows_per_cta, double* __restrict__ ws) {
  pdl_wait();        // programmatic dependent launch: see rmem_internal.h
  pdl_trigger();
  // blockIdx.y = image of a batch [S][P][C] (independent sequences): statistics and workspace per image
  x += static_cast<size_t>(blockIdx.y) * P * C;
  ws += static_cast<size_t>(blockIdx.y) * (2 * (C / cpg) + 1);
  const int tpp = C >> 2;                        // threads per pixel (C = 64, 128, 256 -> 16, 32, 64)
  const int ppb = 256 / tpp;                     // pixels per step
  const int c = (threadIdx.x % tpp) * 4;
  const int ph = threadIdx.x / tpp;
  const int p0 = blockIdx.x * rows_per_cta;
  const int p1 = min(p0 + rows_per_cta, P);
  float s = 0.f, q = 0.f;
  const float4 pb = pre_bias ? *reinterpret_cast<const float4*>(pre_bias + c) : make_float4(0.f, 0.f, 0.f, 0.f);
#pragma unroll 4
  for (int pix = p0 + ph; pix < p1; pix += ppb) {
    float4 v = ld4(x + static_cast<size_t>(pix) * C + c);
    v.x += pb.x; v.y += pb.y; v.z += pb.z; v.w += pb.w;
    s += (v.x + v.y) + (v.z + v.w);
    q += (v.x * v.x + v.y * v.y) + (v.z * v.z + v.w * v.w);
  }
  // lanes sharing a group: cpg / 4 consecutive lanes (cpg is a power of two in [4, 32])
  const int span = cpg >> 2;
  for (int o = span >> 1; o > 0; o >>= 1) {
    s += __shfl_xor_sync(0xffffffffu, s, o);
    q += __shfl_xor_sync(0xffffffffu, q, o);
  }
  __shared__ float sh[64];
  if (threadIdx.x < 64) sh[threadIdx.x] = 0.f;
  __syncthreads();
  if ((threadIdx.x & (span - 1)) == 0) {
    const int g = c / cpg;
    atomicAdd(&sh[2 * g], s);
    atomicAdd(&sh[2 * g + 1], q);
  }
  __syncthreads();
  const int groups = C / cpg;
  if (threadIdx.x < 2 * groups) atomicAdd(ws + threadIdx.x, static_cast<double>(sh[threadIdx.x]));
}

template <typename T>
__global__ void __launch_bounds__(256)
gn_nhwc_apply_kernel(const T* __restrict__ x, const float* __restrict__ pre_bias,
                     const float* __restrict__ gamma, const float* __restrict__ beta, T* __restrict__ y,
                     int P, int C, int cpg, int relu, double* __restrict__ ws) {
  pdl_wait();        // programmatic dependent launch: see rmem_internal.h
  pdl_trigger();
  x += static_cast<size_t>(blockIdx.y) * P * C;
  y += static_cast<size_t>(blockIdx.y) * P * C;
  ws += static_cast<size_t>(blockIdx.y) * (2 * (C / cpg) + 1);
  const size_t total4 = static_cast<size_t>(P) * C / 4;
  __shared__ float gmean[32], grstd[32];
  gn_group_stats(ws, C / cpg, static_cast<double>(cpg) * P, gmean, grstd);
  for (size_t i = blockIdx.x * static_cast<size_t>(blockDim.x) + threadIdx.x; i < total4;
       i += static_cast<size_t>(gridDim.x) * blockDim.x) {
    const int c = static_cast<int>((i * 4) % C);      // 4 consecutive channels, same group (cpg % 4 == 0)
    const int g = c / cpg;
    const float mean = gmean[g], rstd = grstd[g];
    float4 v = ld4(x + i * 4);
    if (pre_bias) {
      const float4 pb = *reinterpret_cast<const float4*>(pre_bias + c);
      v.x += pb.x; v.y += pb.y; v.z += pb.z; v.w += pb.w;
    }
    const float4 ga = *reinterpret_cast<const float4*>(gamma + c);
    const float4 be = *reinterpret_cast<const float4*>(beta + c);
    float4 o;
    o.x = (v.x - mean) * rstd * ga.x + be.x;
    o.y = (v.y - mean) * rstd * ga.y + be.y;
    o.z = (v.z - mean) * rstd * ga.z + be.z;
    o.w = (v.w - mean) * rstd * ga.w + be.w;
    if (relu) { o.x = fmaxf(o.x, 0.f); o.y = fmaxf(o.y, 0.f); o.z = fmaxf(o.z, 0.f); o.w = fmaxf(o.w, 0.f); }
    st4(y + i * 4, o);
  }
  gn_ws_release(ws, 2 * (C / cpg));
}

int groupnorm_nhwc_launch(const float* x, const float* pre_bias, const float* gamma, const float* beta, float* y,
                          int P, int C, int groups, int relu, void* workspace, cudaStream_t stream) {
  return groupnorm_nhwc_batched_launch(x, pre_bias, gamma, beta, y, P, C, groups, relu, 1, workspace, stream);
}

int groupnorm_nhwc_batched_launch(const float* x, const float* pre_bias, const float* gamma, const float* beta,
                                  float* y, int P, int C, int groups, int relu, int S, void* workspace,
                                  cudaStream_t stream) {
  return groupnorm_nhwc_ex_launch(x, pre_bias, gamma, beta, y, P, C, groups, relu, S, 0, workspace, stream);
}

int groupnorm_nhwc_ex_launch(const void* x, const float* pre_bias, const float* gamma, const float* beta, void* y, int P,
                             int C, int groups, int relu, int S, int f16, void* workspace, cudaStream_t stream) {
  if (groups < 1 || groups > 32 || (C % groups) || (C != 64 && C != 128 && C != 256) || P < 1 || S < 1 || S > 65535)
    return set_error(RMEM_ERR_ARG, "groupnorm_nhwc: C must be 64 / 128 / 256");
  const int cpg = C / groups;
  if ((cpg & (cpg - 1)) || cpg < 4 || cpg > 32)
    return set_error(RMEM_ERR_ARG, "groupnorm_nhwc: channels per group must be a power of two in [4, 32]");
  const int per = (592 + S - 1) / S;             // ~4 CTAs per SM at large P over the whole batch ...
  int rows = (P + per - 1) / per;
  if (rows < 8) rows = 8;                        // ... but never less than two steps of work per CTA
  double* wsd = static_cast<double*>(workspace);
  const dim3 gs((P + rows - 1) / rows, S);
  if (f16) launch_k(gn_nhwc_stats_kernel<__half>, gs, 256, 0, stream, static_cast<const __half*>(x), pre_bias, P, C, cpg, rows, wsd);
  else launch_k(gn_nhwc_stats_kernel<float>, gs, 256, 0, stream, static_cast<const float*>(x), pre_bias, P, C, cpg, rows, wsd);
  if (int r = check_launch("gn_nhwc_stats_kernel")) return r;
  const size_t total4 = static_cast<size_t>(P) * C / 4;
  int blocks = static_cast<int>((total4 + 255) / 256);
  if (blocks > (148 * RMEM_GN_CTAS_PER_SM + S - 1) / S) blocks = (148 * RMEM_GN_CTAS_PER_SM + S - 1) / S;
  if (f16)
    launch_k(gn_nhwc_apply_kernel<__half>, dim3(blocks, S), 256, 0, stream, static_cast<const __half*>(x), pre_bias, gamma,
             beta, static_cast<__half*>(y), P, C, cpg, relu, wsd);
  else
    launch_k(gn_nhwc_apply_kernel<float>, dim3(blocks, S), 256, 0, stream, static_cast<const float*>(x), pre_bias, gamma,
             beta, static_cast<float*>(y), P, C, cpg, relu, wsd);
  return check_launch("gn_nhwc_apply_kernel");
}

// ------------------------------------------------------------------------------------------------
// out[y][x][c] = a[y][x][c] + bilinear(x_in)[y][x][c], align_corners = True (F.interpolate, fpn.py:49-60)
template <typename T>
__global__ void __launch_bounds__(256)
upsample_add_nhwc_kernel(const T* __restrict__ xin, int hi, int wi, const float* __restrict__ a,
                         const float* __restrict__ a_bias, T* __restrict__ out, int ho, int wo, int C) {
  pdl_wait();        // programmatic dependent launch: see rmem_internal.h
  pdl_trigger();
  // blockIdx.y = image of a batch: xin [S][hi*wi][C], a / out [S][ho*wo][C]
  xin += static_cast<size_t>(blockIdx.y) * hi * wi * C;
  if (a) a += static_cast<size_t>(blockIdx.y) * ho * wo * C;
  out += static_cast<size_t>(blockIdx.y) * ho * wo * C;
  const int c4 = C / 4;
  const size_t total = static_cast<size_t>(ho) * wo * c4;
  const float sy = ho > 1 ? static_cast<float>(hi - 1) / static_cast<float>(ho - 1) : 0.f;
  const float sx = wo > 1 ? static_cast<float>(wi - 1) / static_cast<float>(wo - 1) : 0.f;
  for (size_t i = blockIdx.x * static_cast<size_t>(blockDim.x) + threadIdx.x; i < total;
       i += static_cast<size_t>(gridDim.x) * blockDim.x) {
    const int c = static_cast<int>(i % c4) * 4;
    const int pix = static_cast<int>(i / c4);
    const int y = pix / wo, x = pix - y * wo;
    const float fy = sy * y, fx = sx * x;
    const int y0 = static_cast<int>(fy), x0 = static_cast<int>(fx);
    const int y1 = min(y0 + 1, hi - 1), x1 = min(x0 + 1, wi - 1);
    const float ly = fy - y0, lx = fx - x0, hy = 1.f - ly, hx = 1.f - lx;
    const float4 v00 = ld4(xin + (static_cast<size_t>(y0) * wi + x0) * C + c);
    const float4 v01 = ld4(xin + (static_cast<size_t>(y0) * wi + x1) * C + c);
    const float4 v10 = ld4(xin + (static_cast<size_t>(y1) * wi + x0) * C + c);
    const float4 v11 = ld4(xin + (static_cast<size_t>(y1) * wi + x1) * C + c);
    float4 o = a ? *reinterpret_cast<const float4*>(a + static_cast<size_t>(pix) * C + c)
                 : make_float4(0.f, 0.f, 0.f, 0.f);
    if (a_bias) {
      const float4 ab = *reinterpret_cast<const float4*>(a_bias + c);
      o.x += ab.x; o.y += ab.y; o.z += ab.z; o.w += ab.w;
    }
    o.x += hy * (hx * v00.x + lx * v01.x) + ly * (hx * v10.x + lx * v11.x);
    o.y += hy * (hx * v00.y + lx * v01.y) + ly * (hx * v10.y + lx * v11.y);
    o.z += hy * (hx * v00.z + lx * v01.z) + ly * (hx * v10.z + lx * v11.z);
    o.w += hy * (hx * v00.w + lx * v01.w) + ly * (hx * v10.w + lx * v11.w);
    st4(out + static_cast<size_t>(pix) * C + c, o);
  }
}

int upsample_add_nhwc_launch(const float* xin, int hi, int wi, const float* a, const float* a_bias, float* out,
                             int ho, int wo, int C, cudaStream_t stream) {
  return upsample_add_nhwc_batched_launch(xin, hi, wi, a, a_bias, out, ho, wo, C, 1, stream);
}

int upsample_add_nhwc_batched_launch(const float* xin, int hi, int wi, const float* a, const float* a_bias, float* out,
                                     int ho, int wo, int C, int S, cudaStream_t stream) {
  return upsample_add_nhwc_ex_launch(xin, hi, wi, a, a_bias, out, ho, wo, C, S, 0, stream);
}

int upsample_add_nhwc_ex_launch(const void* xin, int hi, int wi, const float* a, const float* a_bias, void* out, int ho,
                                int wo, int C, int S, int f16, cudaStream_t stream) {
  if ((C % 4) || hi < 1 || wi < 1 || ho < 1 || wo < 1 || S < 1 || S > 65535)
    return set_error(RMEM_ERR_ARG, "upsample_add_nhwc: C must be a multiple of 4");
  const size_t total = static_cast<size_t>(ho) * wo * (C / 4);
  int blocks = static_cast<int>((total + 255) / 256);
  if (blocks > (148 * RMEM_GN_CTAS_PER_SM + S - 1) / S) blocks = (148 * RMEM_GN_CTAS_PER_SM + S - 1) / S;
  if (f16)
    launch_k(upsample_add_nhwc_kernel<__half>, dim3(blocks, S), 256, 0, stream, static_cast<const __half*>(xin), hi, wi, a,
             a_bias, static_cast<__half*>(out), ho, wo, C);
  else
    launch_k(upsample_add_nhwc_kernel<float>, dim3(blocks, S), 256, 0, stream, static_cast<const float*>(xin), hi, wi, a,
             a_bias, static_cast<float*>(out), ho, wo, C);
  return check_launch("upsample_add_nhwc_kernel");
}

// ------------------------------------------------------------------------------------------------
// One decoder stage in one pass (decoders/fpn.py:44-63): GroupNorm(8) + ReLU of a convolution's raw output x, its
// bilinear (align_corners) resize, and the add of the next shortcut's adapter branch --
//   out[y][x][c] = a[y][x][c] + a_bias[c] + bilinear(relu(gn(x + pre_bias)))[y][x][c].
// The four neighbours are normalised on the fly (statistics from gn_nhwc_stats_kernel), so the normalised map is
// never stored and the stage is two launches (statistics, this) instead of three.
template <typename T>
__global__ void __launch_bounds__(256)
gn_upsample_add_nhwc_kernel(const T* __restrict__ xin, const float* __restrict__ pre_bias,
                            const float* __restrict__ gamma, const float* __restrict__ beta, int hi, int wi, int cpg,
                            const float* __restrict__ a, const float* __restrict__ a_bias, T* __restrict__ out, int ho,
                            int wo, int C, double* __restrict__ ws) {
  pdl_wait();        // programmatic dependent launch: see rmem_internal.h
  pdl_trigger();
  xin += static_cast<size_t>(blockIdx.y) * hi * wi * C;
  if (a) a += static_cast<size_t>(blockIdx.y) * ho * wo * C;
  out += static_cast<size_t>(blockIdx.y) * ho * wo * C;
  ws += static_cast<size_t>(blockIdx.y) * (2 * (C / cpg) + 1);
  const int c4 = C / 4;
  const size_t total = static_cast<size_t>(ho) * wo * c4;
  const float sy = ho > 1 ? static_cast<float>(hi - 1) / static_cast<float>(ho - 1) : 0.f;
  const float sx = wo > 1 ? static_cast<float>(wi - 1) / static_cast<float>(wo - 1) : 0.f;
  __shared__ float gmean[32], grstd[32];
  gn_group_stats(ws, C / cpg, static_cast<double>(cpg) * hi * wi, gmean, grstd);
  for (size_t i = blockIdx.x * static_cast<size_t>(blockDim.x) + threadIdx.x; i < total;
       i += static_cast<size_t>(gridDim.x) * blockDim.x) {
    const int c = static_cast<int>(i % c4) * 4;
    const int pix = static_cast<int>(i / c4);
    const int g = c / cpg;
    const float rstd = grstd[g];
    float4 sc = *reinterpret_cast<const float4*>(gamma + c);          // y = relu(v * sc + sh)
    sc.x *= rstd; sc.y *= rstd; sc.z *= rstd; sc.w *= rstd;
    float4 sh = *reinterpret_cast<const float4*>(beta + c);
    float4 pb = make_float4(0.f, 0.f, 0.f, 0.f);
    if (pre_bias) pb = *reinterpret_cast<const float4*>(pre_bias + c);
    const float mean = gmean[g];
    sh.x += (pb.x - mean) * sc.x; sh.y += (pb.y - mean) * sc.y; sh.z += (pb.z - mean) * sc.z; sh.w += (pb.w - mean) * sc.w;
    const int y = pix / wo, x = pix - y * wo;
    const float fy = sy * y, fx = sx * x;
    const int y0 = static_cast<int>(fy), x0 = static_cast<int>(fx);
    const int y1 = min(y0 + 1, hi - 1), x1 = min(x0 + 1, wi - 1);
    const float ly = fy - y0, lx = fx - x0, hy = 1.f - ly, hx = 1.f - lx;
    auto nrm = [&](int yy, int xx) {
      float4 v = ld4(xin + (static_cast<size_t>(yy) * wi + xx) * C + c);
      v.x = fmaxf(fmaf(v.x, sc.x, sh.x), 0.f); v.y = fmaxf(fmaf(v.y, sc.y, sh.y), 0.f);
      v.z = fmaxf(fmaf(v.z, sc.z, sh.z), 0.f); v.w = fmaxf(fmaf(v.w, sc.w, sh.w), 0.f);
      return v;
    };
    const float4 v00 = nrm(y0, x0), v01 = nrm(y0, x1), v10 = nrm(y1, x0), v11 = nrm(y1, x1);
    float4 o = a ? *reinterpret_cast<const float4*>(a + static_cast<size_t>(pix) * C + c)
                 : make_float4(0.f, 0.f, 0.f, 0.f);
    if (a_bias) {
      const float4 ab = *reinterpret_cast<const float4*>(a_bias + c);
      o.x += ab.x; o.y += ab.y; o.z += ab.z; o.w += ab.w;
    }
    o.x += hy * (hx * v00.x + lx * v01.x) + ly * (hx * v10.x + lx * v11.x);
    o.y += hy * (hx * v00.y + lx * v01.y) + ly * (hx * v10.y + lx * v11.y);
    o.z += hy * (hx * v00.z + lx * v01.z) + ly * (hx * v10.z + lx * v11.z);
    o.w += hy * (hx * v00.w + lx * v01.w) + ly * (hx * v10.w + lx * v11.w);
    st4(out + static_cast<size_t>(pix) * C + c, o);
  }
  gn_ws_release(ws, 2 * (C / cpg));
}

int gn_upsample_add_nhwc_launch(const void* x, const float* pre_bias, const float* gamma, const float* beta, int hi, int wi,
                                int C, int groups, const float* a, const float* a_bias, void* out, int ho, int wo, int S,
                                int f16, void* workspace, cudaStream_t stream) {
  if (groups < 1 || (C % groups) || (C != 64 && C != 128 && C != 256) || hi < 1 || wi < 1 || ho < 1 || wo < 1 || S < 1 ||
      S > 65535)
    return set_error(RMEM_ERR_ARG, "gn_upsample_add_nhwc: C must be 64 / 128 / 256");
  const int cpg = C / groups, P = hi * wi;
  if ((cpg & (cpg - 1)) || cpg < 4 || cpg > 32)
    return set_error(RMEM_ERR_ARG, "gn_upsample_add_nhwc: channels per group must be a power of two in [4, 32]");
  const int per = (592 + S - 1) / S;
  int rows = (P + per - 1) / per;
  if (rows < 8) rows = 8;
  double* wsd = static_cast<double*>(workspace);
  const dim3 gs((P + rows - 1) / rows, S);
  if (f16) launch_k(gn_nhwc_stats_kernel<__half>, gs, 256, 0, stream, static_cast<const __half*>(x), pre_bias, P, C, cpg, rows, wsd);
  else launch_k(gn_nhwc_stats_kernel<float>, gs, 256, 0, stream, static_cast<const float*>(x), pre_bias, P, C, cpg, rows, wsd);
  if (int r = check_launch("gn_nhwc_stats_kernel")) return r;
  const size_t total = static_cast<size_t>(ho) * wo * (C / 4);
  int blocks = static_cast<int>((total + 255) / 256);
  if (blocks > (148 * RMEM_GN_CTAS_PER_SM + S - 1) / S) blocks = (148 * RMEM_GN_CTAS_PER_SM + S - 1) / S;
  if (f16)
    launch_k(gn_upsample_add_nhwc_kernel<__half>, dim3(blocks, S), 256, 0, stream, static_cast<const __half*>(x), pre_bias,
             gamma, beta, hi, wi, cpg, a, a_bias, static_cast<__half*>(out), ho, wo, C, wsd);
  else
    launch_k(gn_upsample_add_nhwc_kernel<float>, dim3(blocks, S), 256, 0, stream, static_cast<const float*>(x), pre_bias,
             gamma, beta, hi, wi, cpg, a, a_bias, static_cast<float*>(out), ho, wo, C, wsd);
  return check_launch("gn_upsample_add_nhwc_kernel");
}

// ------------------------------------------------------------------------------------------------
// decode_current_logits tail (aot_engine.py:451-463): logits = conv_out + bias, channels beyond the
// sequence's objects masked to -1e10, kept as pred_id_logits [C][Hs][Ws]; then the bilinear
// (align_corners = True) resize to the output size, [C][Ho][Wo].
struct ValidCounts {
  int n[64];                 // valid channels (objects + 1) per image of the batch
};

__global__ void __launch_bounds__(256)
finish_logits_kernel(const float* __restrict__ conv, long long sn, long long sc, long long sy, long long sx,
                     const float* __restrict__ bias, float* __restrict__ pred, int C, int Hs, int Ws,
                     const ValidCounts valid) {
  pdl_wait();        // programmatic dependent launch: see rmem_internal.h
  pdl_trigger();
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  const int total = C * Hs * Ws;
  if (i >= total) return;
  // blockIdx.y = image of a batch: conv strided by sn, pred [S][C][Hs][Ws]
  conv += blockIdx.y * sn;
  pred += static_cast<size_t>(blockIdx.y) * total;
  const int n_valid = valid.n[blockIdx.y];
  const int c = i / (Hs * Ws), r = i - c * (Hs * Ws);
  const int y = r / Ws, x = r - y * Ws;
  const float v = conv[c * sc + y * sy + x * sx] + (bias ? bias[c] : 0.f);
  pred[i] = c < n_valid ? v : -1e10f;
}

__global__ void __launch_bounds__(256)
upsample_logits_kernel(const float* __restrict__ pred, int C, int Hs, int Ws, float* __restrict__ out, int Ho,
                       int Wo, float* __restrict__ label) {
  pdl_wait();        // programmatic dependent launch: see rmem_internal.h
  pdl_trigger();
  pred += static_cast<size_t>(blockIdx.y) * C * Hs * Ws;
  out += static_cast<size_t>(blockIdx.y) * C * Ho * Wo;
  if (label) label += static_cast<size_t>(blockIdx.y) * Ho * Wo;   // argmax over the channels (first maximum), f32
  // thread = 4 consecutive output pixels of one row, all C channels (weights computed once per pixel)
  const int wq = (Wo + 3) / 4;
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= Ho * wq) return;
  const int y = i / wq, x4 = (i - y * wq) * 4;
  const float sy = Ho > 1 ? static_cast<float>(Hs - 1) / static_cast<float>(Ho - 1) : 0.f;
  const float sx = Wo > 1 ? static_cast<float>(Ws - 1) / static_cast<float>(Wo - 1) : 0.f;
  const float fy = sy * y;
  const int y0 = static_cast<int>(fy), y1 = min(y0 + 1, Hs - 1);
  const float ly = fy - y0, hy = 1.f - ly;
  int x0[4], x1[4];
  float lx[4];
#pragma unroll
  for (int k = 0; k < 4; ++k) {
    const float fx = sx * (x4 + k);
    x0[k] = min(static_cast<int>(fx), Ws - 1);
    x1[k] = min(x0[k] + 1, Ws - 1);
    lx[k] = fx - x0[k];
  }
  const bool full = x4 + 3 < Wo && (Wo & 3) == 0;
  float best[4] = {-INFINITY, -INFINITY, -INFINITY, -INFINITY};
  float arg[4] = {0.f, 0.f, 0.f, 0.f};
  for (int c = 0; c < C; ++c) {
    const float* r0 = pred + (static_cast<size_t>(c) * Hs + y0) * Ws;
    const float* r1 = pred + (static_cast<size_t>(c) * Hs + y1) * Ws;
    float o[4];
#pragma unroll
    for (int k = 0; k < 4; ++k) {
      o[k] = hy * ((1.f - lx[k]) * r0[x0[k]] + lx[k] * r0[x1[k]]) + ly * ((1.f - lx[k]) * r1[x0[k]] + lx[k] * r1[x1[k]]);
      if (c == 0 || o[k] > best[k]) { best[k] = o[k]; arg[k] = static_cast<float>(c); }
    }
    float* d = out + (static_cast<size_t>(c) * Ho + y) * Wo + x4;
    if (full) {
      *reinterpret_cast<float4*>(d) = make_float4(o[0], o[1], o[2], o[3]);
    } else {
      for (int k = 0; k < 4 && x4 + k < Wo; ++k) d[k] = o[k];
    }
  }
  if (label) {
    float* d = label + static_cast<size_t>(y) * Wo + x4;
    if (full) {
      *reinterpret_cast<float4*>(d) = make_float4(arg[0], arg[1], arg[2], arg[3]);
    } else {
      for (int k = 0; k < 4 && x4 + k < Wo; ++k) d[k] = arg[k];
    }
  }
}

int finish_logits_launch(const float* conv, long long sc, long long sy, long long sx, const float* bias, float* pred,
                         int C, int Hs, int Ws, int n_valid, float* out, int Ho, int Wo, cudaStream_t stream) {
  return finish_logits_batched_launch(conv, 0, sc, sy, sx, bias, pred, C, Hs, Ws, &n_valid, 1, out, Ho, Wo, stream);
}

int finish_logits_batched_launch(const float* conv, long long sn, long long sc, long long sy, long long sx,
                                 const float* bias, float* pred, int C, int Hs, int Ws, const int* n_valid_host, int S,
                                 float* out, int Ho, int Wo, cudaStream_t stream) {
  if (C < 1 || Hs < 1 || Ws < 1 || S < 1 || S > 64 || !n_valid_host)
    return set_error(RMEM_ERR_ARG, "finish_logits: bad shape (at most 64 images per call)");
  ValidCounts valid;
  for (int s = 0; s < 64; ++s) valid.n[s] = s < S ? n_valid_host[s] : 0;
  const int total = C * Hs * Ws;
  launch_k(finish_logits_kernel, dim3((total + 255) / 256, S), 256, 0, stream, conv, sn, sc, sy, sx, bias, pred, C, Hs, Ws,
                                                                        valid);
  if (int r = check_launch("finish_logits_kernel")) return r;
  if (out) {
    const int n = Ho * ((Wo + 3) / 4);
    launch_k(upsample_logits_kernel, dim3((n + 255) / 256, S), 256, 0, stream, pred, C, Hs, Ws, out, Ho, Wo,
             static_cast<float*>(nullptr));
    return check_launch("upsample_logits_kernel");
  }
  return RMEM_OK;
}

// ------------------------------------------------------------------------------------------------
// The decoder's last stage in one pass over the conv_4x output (decoders/fpn.py:63-68 + aot_engine.py:451-463):
// GroupNorm(8) + ReLU of the 3x3 convolution's raw output x [P][C] (statistics from gn_nhwc_stats_kernel), the 1x1
// conv_out (C -> Co <= 16 classes) as fp32 dot products, + bias, classes beyond the sequence's objects -> -1e10,
// written as pred_id_logits [Co][P].  The normalised C-channel map is never stored (13 MB at 480p), the cuDNN 1x1
// convolution and the bias / mask pass disappear.
constexpr int kHeadMaxCo = 16;
struct HeadParams {
  const void* x;             // fp32 or fp16 [S][P][C]
  const float* pre_bias;
  const float* gamma;
  const float* beta;
  const float* w_out;        // [Co][C]
  const float* b_out;        // [Co]
  float* pred;               // [S][Co][P]
  double* ws;
  int P, C, cpg, Co;
  ValidCounts valid;
};

// Four threads per pixel: thread j takes the 8-channel chunks q = j (mod 4), normalises them and folds them into up to
// 16 class sums held in registers (weights, scales and shifts are shared-memory reads: the four threads of a pixel read
// four different rows, laid out 32 bytes apart modulo the bank width so a warp's access is conflict-free); two xor
// shuffles per class finish the sums.  A thread carries four pixels at once, so a weight row read from shared memory
// feeds 64 FMAs instead of 16 (with one pixel per thread the kernel was bound by those reads).  (The first version spread a pixel over a whole warp and reduced 11 partial sums
// over 32 lanes: 99 shuffle / shared-memory wavefronts per pixel made it MIO-bound at 24 us; one thread per pixel left
// 5 warps per SM and was latency-bound at 14 us.)
constexpr int kHeadThreads = 128;
constexpr int kHeadSplit = 4;                           // threads per pixel
template <typename T> struct Vec8;
template <> struct Vec8<float> {
  static __device__ __forceinline__ void load(const float* p, float* v) {
    const float4 a = *reinterpret_cast<const float4*>(p), b = *reinterpret_cast<const float4*>(p + 4);
    v[0] = a.x; v[1] = a.y; v[2] = a.z; v[3] = a.w; v[4] = b.x; v[5] = b.y; v[6] = b.z; v[7] = b.w;
  }
};
template <> struct Vec8<__half> {
  static __device__ __forceinline__ void load(const __half* p, float* v) {
    const uint4 r = *reinterpret_cast<const uint4*>(p);
    const __half2* h = reinterpret_cast<const __half2*>(&r);
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      const float2 f = __half22float2(h[i]);
      v[2 * i] = f.x; v[2 * i + 1] = f.y;
    }
  }
};
// shared-memory float offset of the weight row of channel c: rows of one thread-of-a-pixel are contiguous (16 floats
// each), the four groups of rows start 8 floats (32 bytes) further apart than a multiple of the bank width
__device__ __forceinline__ int head_row(int c, int C) {
  const int q = c >> 3, k = c & 7;                      // chunk, channel inside it
  const int j = q & (kHeadSplit - 1), r = (q >> 2) * 8 + k;
  return j * ((C / kHeadSplit) * kHeadMaxCo + 8) + r * kHeadMaxCo;
}

template <typename T>
__global__ void __launch_bounds__(kHeadThreads) gn_head_kernel(const HeadParams p) {
  pdl_wait();        // programmatic dependent launch: see rmem_internal.h
  pdl_trigger();
  extern __shared__ __align__(16) float head_sh[];      // weight rows (see head_row) | scale [C] | shift [C]
  __shared__ float gmean[32], grstd[32];
  const int C = p.C, Co = p.Co, P = p.P;
  float* w_t = head_sh;
  float* scale = head_sh + C * kHeadMaxCo + kHeadSplit * 8;
  float* shift = scale + C;
  const T* x = static_cast<const T*>(p.x) + static_cast<size_t>(blockIdx.y) * P * C;
  float* pred = p.pred + static_cast<size_t>(blockIdx.y) * Co * P;
  double* ws = p.ws + static_cast<size_t>(blockIdx.y) * (2 * (C / p.cpg) + 1);
  const int n_valid = p.valid.n[blockIdx.y];
  gn_group_stats(ws, C / p.cpg, static_cast<double>(p.cpg) * P, gmean, grstd);
  // (global reads coalesced along c: read the other way round, every CTA issued 2048 single-sector requests and the
  // fill alone cost 10 us over the grid)
  for (int i = threadIdx.x; i < C * kHeadMaxCo; i += kHeadThreads) {
    const int o = i / C, c = i - o * C;
    w_t[head_row(c, C) + o] = o < Co ? p.w_out[o * C + c] : 0.f;
  }
  for (int c = threadIdx.x; c < C; c += kHeadThreads) {
    const int g = c / p.cpg;
    const float sc = p.gamma[c] * grstd[g];
    scale[c] = sc;
    shift[c] = p.beta[c] + ((p.pre_bias ? p.pre_bias[c] : 0.f) - gmean[g]) * sc;
  }
  __syncthreads();
  const int j = threadIdx.x & (kHeadSplit - 1);
  const int rounds = C / (8 * kHeadSplit);              // 8-channel chunks per thread: 2 (C = 64) or 4 (C = 128)
  constexpr int kPix = 4;                               // pixels per thread: every weight row read feeds 4 pixels
  constexpr int kStepPix = kHeadThreads / kHeadSplit * kPix;   // pixels per CTA step (128)
  const int n_steps = (P + kStepPix - 1) / kStepPix;
  for (int step = blockIdx.x; step < n_steps; step += gridDim.x) {   // whole warps stay in the loop (shuffles below)
    // the thread's pixels: pix0 + 32 i (i = 0..3), so that for a fixed i a warp's 8 pixel groups are consecutive pixels
    const int pix0 = step * kStepPix + (threadIdx.x >> 2);
    float acc[kPix][kHeadMaxCo];
#pragma unroll
    for (int i = 0; i < kPix; ++i)
#pragma unroll
      for (int o = 0; o < kHeadMaxCo; ++o) acc[i][o] = 0.f;
#pragma unroll 1
    for (int r = 0; r < rounds; ++r) {
      float v[kPix][8];
#pragma unroll
      for (int i = 0; i < kPix; ++i) {
        const int pix = pix0 + i * (kHeadThreads / kHeadSplit);
        Vec8<T>::load(x + static_cast<size_t>(pix < P ? pix : 0) * C + (r * kHeadSplit + j) * 8, v[i]);
      }
#pragma unroll
      for (int k = 0; k < 8; ++k) {
        const int c = (r * kHeadSplit + j) * 8 + k;
        const float sc = scale[c], sh = shift[c];
        const float4* wr = reinterpret_cast<const float4*>(w_t + j * ((C / kHeadSplit) * kHeadMaxCo + 8) +
                                                           (r * 8 + k) * kHeadMaxCo);
        float y[kPix];
#pragma unroll
        for (int i = 0; i < kPix; ++i) y[i] = fmaxf(fmaf(v[i][k], sc, sh), 0.f);
#pragma unroll
        for (int q = 0; q < kHeadMaxCo / 4; ++q) {
          const float4 w4 = wr[q];
#pragma unroll
          for (int i = 0; i < kPix; ++i) {
            acc[i][4 * q] = fmaf(y[i], w4.x, acc[i][4 * q]);
            acc[i][4 * q + 1] = fmaf(y[i], w4.y, acc[i][4 * q + 1]);
            acc[i][4 * q + 2] = fmaf(y[i], w4.z, acc[i][4 * q + 2]);
            acc[i][4 * q + 3] = fmaf(y[i], w4.w, acc[i][4 * q + 3]);
          }
        }
      }
    }
#pragma unroll
    for (int i = 0; i < kPix; ++i) {
      const int pix = pix0 + i * (kHeadThreads / kHeadSplit);
#pragma unroll
      for (int o = 0; o < kHeadMaxCo; ++o) {
        acc[i][o] += __shfl_xor_sync(0xffffffffu, acc[i][o], 1);
        acc[i][o] += __shfl_xor_sync(0xffffffffu, acc[i][o], 2);
      }
      // thread j of the pixel writes classes j, j + 4, ...: for a fixed class the warp's 8 pixels are one 32-byte run
#pragma unroll
      for (int o = 0; o < kHeadMaxCo; ++o)
        if (pix < P && (o & (kHeadSplit - 1)) == j && o < Co)
          pred[static_cast<size_t>(o) * P + pix] = o < n_valid ? acc[i][o] + p.b_out[o] : -1e10f;
    }
  }
  gn_ws_release(ws, 2 * (C / p.cpg));
}

int logits_head_batched_launch(const void* x, int x_f16, const float* pre_bias, const float* gamma, const float* beta,
                               int Hs, int Ws, int C, int groups, const float* w_out, const float* b_out, int Co,
                               const int* n_valid_host, int S, float* pred, float* out, int Ho, int Wo, float* label,
                               void* workspace, cudaStream_t stream) {
  const int P = Hs * Ws;
  if (groups < 1 || (C % groups) || (C != 64 && C != 128) || Co < 1 || Co > kHeadMaxCo || P < 1 || S < 1 || S > 64 ||
      !n_valid_host)
    return set_error(RMEM_ERR_ARG, "logits_head: C must be 64 / 128, at most 16 classes and 64 images");
  const int cpg = C / groups;
  if ((cpg & (cpg - 1)) || cpg < 4 || cpg > 32)
    return set_error(RMEM_ERR_ARG, "logits_head: channels per group must be a power of two in [4, 32]");
  const int per = (592 + S - 1) / S;
  int rows = (P + per - 1) / per;
  if (rows < 8) rows = 8;
  if (x_f16)
    launch_k(gn_nhwc_stats_kernel<__half>, dim3((P + rows - 1) / rows, S), 256, 0, stream, static_cast<const __half*>(x),
             pre_bias, P, C, cpg, rows, static_cast<double*>(workspace));
  else
    launch_k(gn_nhwc_stats_kernel<float>, dim3((P + rows - 1) / rows, S), 256, 0, stream, static_cast<const float*>(x),
             pre_bias, P, C, cpg, rows, static_cast<double*>(workspace));
  if (int r = check_launch("gn_nhwc_stats_kernel")) return r;
  HeadParams hp;
  hp.x = x; hp.pre_bias = pre_bias; hp.gamma = gamma; hp.beta = beta; hp.w_out = w_out; hp.b_out = b_out;
  hp.pred = pred; hp.ws = static_cast<double*>(workspace); hp.P = P; hp.C = C; hp.cpg = cpg; hp.Co = Co;
  for (int s = 0; s < 64; ++s) hp.valid.n[s] = s < S ? n_valid_host[s] : 0;
  const int ppc = kHeadThreads / kHeadSplit * 4;                        // pixels per CTA step (4 per thread)
  int blocks = (P + ppc - 1) / ppc;                                     // 203 steps at 480p
  if (blocks > (148 * 2 + S - 1) / S) blocks = (148 * 2 + S - 1) / S;   // every CTA ends in one same-address atomic
  const size_t smem = (static_cast<size_t>(C) * (kHeadMaxCo + 2) + kHeadSplit * 8) * sizeof(float);
  if (x_f16) launch_k(gn_head_kernel<__half>, dim3(blocks, S), kHeadThreads, smem, stream, hp);
  else launch_k(gn_head_kernel<float>, dim3(blocks, S), kHeadThreads, smem, stream, hp);
  if (int r = check_launch("gn_head_kernel")) return r;
  if (out) {
    const int n = Ho * ((Wo + 3) / 4);
    launch_k(upsample_logits_kernel, dim3((n + 255) / 256, S), 256, 0, stream, pred, Co, Hs, Ws, out, Ho, Wo, label);
    return check_launch("upsample_logits_kernel");
  }
  return RMEM_OK;
}

// Encoder input: f32 planes [3][H][W] -> bf16 channels-last [H][W][8] with channels 3..7 zero (the padded first
// convolution reads 8-channel pixels so the tensor-core conv kernels apply; the extra weights are zero).
__global__ void __launch_bounds__(256)
pack_image_kernel(const float* __restrict__ img, int HW, __nv_bfloat16* __restrict__ out) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= HW) return;
  const __nv_bfloat162 a = __floats2bfloat162_rn(img[i], img[HW + i]);
  const __nv_bfloat162 b = __floats2bfloat162_rn(img[2 * HW + i], 0.f);
  uint4 pk;
  pk.x = *reinterpret_cast<const uint32_t*>(&a);
  pk.y = *reinterpret_cast<const uint32_t*>(&b);
  pk.z = 0u; pk.w = 0u;
  reinterpret_cast<uint4*>(out)[i] = pk;
}

int pack_image_launch(const float* img, int H, int W, void* out, cudaStream_t stream) {
  if (H < 1 || W < 1) return set_error(RMEM_ERR_ARG, "pack_image: bad shape");
  const int HW = H * W;
  pack_image_kernel<<<(HW + 255) / 256, 256, 0, stream>>>(img, HW, static_cast<__nv_bfloat16*>(out));
  return check_launch("pack_image_kernel");
}

// 16-bit activation types of the cuDNN backbone: bf16 or IEEE fp16 (same storage size, same tensor-core rate; fp16
// carries 3 more mantissa bits, which is what keeps the masks within 99.9 % of the fp32 reference).
template <typename T2> struct Pair16;
template <> struct Pair16<__nv_bfloat162> {
  static __device__ __forceinline__ __nv_bfloat162 from(float a, float b) { return __floats2bfloat162_rn(a, b); }
  static __device__ __forceinline__ float2 to(__nv_bfloat162 v) { return __bfloat1622float2(v); }
};
template <> struct Pair16<__half2> {
  static __device__ __forceinline__ __half2 from(float a, float b) { return __floats2half2_rn(a, b); }
  static __device__ __forceinline__ float2 to(__half2 v) { return __half22float2(v); }
};

// Encoder input for the space-to-depth form of the stem: the 7x7 stride-2 convolution on 3 channels equals a 4x4
// stride-1 convolution on the 2x2 space-to-depth image (12 channels, padded to 16), which the tensor-core
// convolution kernels run ~4x faster than the 8-channel strided form.  f32 planes [3][H][W] -- or u8 interleaved
// [H][W][3] pixels, normalised here as (u8 / 255 - mean[c]) * inv_std[c] so a caller uploads a quarter of the bytes
// -- -> 16-bit channels-last [Hs+3][Ws+3][16], Hs = ceil(H/2): out[y'+2][x'+2][(dy*2+dx)*3+c] = img[c][2y'+dy][2x'+dx];
// the two leading / one trailing border rows and columns (the convolution's padding), out-of-image pixels and
// channels 12..15 are zero.
struct PackNorm {
  float mean[3], inv_std[3];
};

template <typename T2, bool kU8>
__global__ void __launch_bounds__(256)
pack_image_s2d_kernel(const void* __restrict__ img_, PackNorm nrm, int H, int W, int Hs, int Ws, T2* __restrict__ out) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  const int Wp = Ws + 3;
  if (i >= (Hs + 3) * Wp) return;
  // blockIdx.y = image of a batch: img [S][3][H][W] f32 or [S][H][W][3] u8, out [S][Hs+3][Ws+3][16]
  img_ = static_cast<const uint8_t*>(img_) + static_cast<size_t>(blockIdx.y) * 3 * H * W * (kU8 ? 1 : 4);
  out += static_cast<size_t>(blockIdx.y) * (Hs + 3) * Wp * 8;
  const int yp = i / Wp - 2, xp = i % Wp - 2;
  float v[16];
#pragma unroll
  for (int k = 0; k < 16; ++k) v[k] = 0.f;
  if (yp >= 0 && yp < Hs && xp >= 0 && xp < Ws) {
#pragma unroll
    for (int dy = 0; dy < 2; ++dy) {
#pragma unroll
      for (int dx = 0; dx < 2; ++dx) {
        const int y = 2 * yp + dy, x = 2 * xp + dx;
        if (y < H && x < W) {
          if (kU8) {
            const uint8_t* px = static_cast<const uint8_t*>(img_) + (static_cast<size_t>(y) * W + x) * 3;
#pragma unroll
            for (int c = 0; c < 3; ++c)
              v[(dy * 2 + dx) * 3 + c] = (static_cast<float>(px[c]) * (1.f / 255.f) - nrm.mean[c]) * nrm.inv_std[c];
          } else {
            const float* img = static_cast<const float*>(img_);
#pragma unroll
            for (int c = 0; c < 3; ++c) v[(dy * 2 + dx) * 3 + c] = img[(static_cast<size_t>(c) * H + y) * W + x];
          }
        }
      }
    }
  }
  uint4 pk[2];
  T2* h = reinterpret_cast<T2*>(pk);
#pragma unroll
  for (int k = 0; k < 8; ++k) h[k] = Pair16<T2>::from(v[2 * k], v[2 * k + 1]);
  reinterpret_cast<uint4*>(out)[2 * i] = pk[0];
  reinterpret_cast<uint4*>(out)[2 * i + 1] = pk[1];
}

int pack_image_s2d_ex_launch(const void* img, int img_u8_hwc, const float* mean_host, const float* inv_std_host,
                             int H, int W, void* out, int out_f16, cudaStream_t stream) {
  return pack_image_s2d_batched_launch(img, img_u8_hwc, mean_host, inv_std_host, H, W, out, out_f16, 1, stream);
}

int pack_image_s2d_batched_launch(const void* img, int img_u8_hwc, const float* mean_host, const float* inv_std_host,
                                  int H, int W, void* out, int out_f16, int S, cudaStream_t stream) {
  if (H < 1 || W < 1 || S < 1 || S > 65535) return set_error(RMEM_ERR_ARG, "pack_image_s2d: bad shape");
  if (img_u8_hwc && (!mean_host || !inv_std_host))
    return set_error(RMEM_ERR_ARG, "pack_image_s2d: u8 input needs mean / inv_std");
  PackNorm nrm = {{0.f, 0.f, 0.f}, {1.f, 1.f, 1.f}};
  if (img_u8_hwc)
    for (int c = 0; c < 3; ++c) { nrm.mean[c] = mean_host[c]; nrm.inv_std[c] = inv_std_host[c]; }
  const int Hs = (H + 1) / 2, Ws = (W + 1) / 2;
  const int n = (Hs + 3) * (Ws + 3);
  const dim3 grid((n + 255) / 256, S);
  if (out_f16) {
    if (img_u8_hwc) pack_image_s2d_kernel<__half2, true><<<grid, 256, 0, stream>>>(img, nrm, H, W, Hs, Ws, static_cast<__half2*>(out));
    else pack_image_s2d_kernel<__half2, false><<<grid, 256, 0, stream>>>(img, nrm, H, W, Hs, Ws, static_cast<__half2*>(out));
  } else {
    if (img_u8_hwc) pack_image_s2d_kernel<__nv_bfloat162, true><<<grid, 256, 0, stream>>>(img, nrm, H, W, Hs, Ws, static_cast<__nv_bfloat162*>(out));
    else pack_image_s2d_kernel<__nv_bfloat162, false><<<grid, 256, 0, stream>>>(img, nrm, H, W, Hs, Ws, static_cast<__nv_bfloat162*>(out));
  }
  return check_launch("pack_image_s2d_kernel");
}

int pack_image_s2d_launch(const float* img, int H, int W, void* out, cudaStream_t stream) {
  return pack_image_s2d_ex_launch(img, 0, nullptr, nullptr, H, W, out, 0, stream);
}

// Encoder stem max-pool (encoders/resnet.py:154, MaxPool2d(3, stride 2, padding 1)) on 16-bit channels-last
// [H][W][C] -> [Ho][Wo][C]: one thread per output pixel and 8-channel group, 128-bit loads and stores.  With a
// bias and / or relu it computes relu(maxpool(x) + bias) = maxpool(relu(x + bias)) (both maps are monotone per
// channel), so the stem convolution can run without its own bias + ReLU pass over the full-resolution map.
template <typename T2>
__global__ void __launch_bounds__(256)
maxpool3s2_nhwc_kernel(const T2* __restrict__ x, int H, int W, int C8, const float* __restrict__ bias,
                       int relu, T2* __restrict__ y, int Ho, int Wo) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= Ho * Wo * C8) return;
  x += static_cast<size_t>(blockIdx.y) * H * W * C8 * 4;          // blockIdx.y = image of a batch
  y += static_cast<size_t>(blockIdx.y) * Ho * Wo * C8 * 4;
  const int c8 = i % C8, pix = i / C8;
  const int ox = pix % Wo, oy = pix / Wo;
  const T2 ninf = Pair16<T2>::from(-INFINITY, -INFINITY);
  T2 m[4] = {ninf, ninf, ninf, ninf};
#pragma unroll
  for (int dy = 0; dy < 3; ++dy) {
    const int iy = 2 * oy - 1 + dy;
    if (iy < 0 || iy >= H) continue;
#pragma unroll
    for (int dx = 0; dx < 3; ++dx) {
      const int ix = 2 * ox - 1 + dx;
      if (ix < 0 || ix >= W) continue;
      const uint4 v = *reinterpret_cast<const uint4*>(x + ((static_cast<size_t>(iy) * W + ix) * C8 + c8) * 4);
      const T2* h = reinterpret_cast<const T2*>(&v);
#pragma unroll
      for (int e = 0; e < 4; ++e) m[e] = __hmax2(m[e], h[e]);
    }
  }
  uint4 o;
  T2* oh = reinterpret_cast<T2*>(&o);
#pragma unroll
  for (int e = 0; e < 4; ++e) {
    if (bias || relu) {
      float2 f = Pair16<T2>::to(m[e]);
      if (bias) { f.x += bias[c8 * 8 + 2 * e]; f.y += bias[c8 * 8 + 2 * e + 1]; }
      if (relu) { f.x = fmaxf(f.x, 0.f); f.y = fmaxf(f.y, 0.f); }
      oh[e] = Pair16<T2>::from(f.x, f.y);
    } else {
      oh[e] = m[e];
    }
  }
  *reinterpret_cast<uint4*>(y + (static_cast<size_t>(pix) * C8 + c8) * 4) = o;
}

int maxpool3s2_nhwc_ex_launch(const void* x, int H, int W, int C, const float* bias, int relu, void* y, int f16,
                              cudaStream_t stream) {
  return maxpool3s2_nhwc_batched_launch(x, H, W, C, bias, relu, y, f16, 1, stream);
}

int maxpool3s2_nhwc_batched_launch(const void* x, int H, int W, int C, const float* bias, int relu, void* y, int f16,
                                   int S, cudaStream_t stream) {
  if (H < 1 || W < 1 || C < 8 || (C & 7) || S < 1 || S > 65535)
    return set_error(RMEM_ERR_ARG, "maxpool3s2: bad shape (C % 8 != 0)");
  const int Ho = (H - 1) / 2 + 1, Wo = (W - 1) / 2 + 1;
  const int n = Ho * Wo * (C / 8);
  const dim3 grid((n + 255) / 256, S);
  if (f16)
    maxpool3s2_nhwc_kernel<__half2><<<grid, 256, 0, stream>>>(static_cast<const __half2*>(x), H, W, C / 8, bias,
                                                             relu, static_cast<__half2*>(y), Ho, Wo);
  else
    maxpool3s2_nhwc_kernel<__nv_bfloat162><<<grid, 256, 0, stream>>>(
        static_cast<const __nv_bfloat162*>(x), H, W, C / 8, bias, relu, static_cast<__nv_bfloat162*>(y), Ho, Wo);
  return check_launch("maxpool3s2_nhwc_kernel");
}

int maxpool3s2_nhwc_launch(const void* x, int H, int W, int C, const float* bias, int relu, void* y,
                           cudaStream_t stream) {
  return maxpool3s2_nhwc_ex_launch(x, H, W, C, bias, relu, y, 0, stream);
}

// label[y][x] = argmax_c logits[c][y][x] (first maximum, like torch.argmax) as f32 -- the evaluator's
// softmax + argmax (networks/managers/evaluator.py:430-441; softmax is monotone, so it is skipped).
__global__ void __launch_bounds__(256)
argmax_label_kernel(const float* __restrict__ logits, int C, int HW, float* __restrict__ label) {
  pdl_wait();        // programmatic dependent launch: see rmem_internal.h
  pdl_trigger();
  const int i = (blockIdx.x * blockDim.x + threadIdx.x) * 4;
  if (i >= HW) return;
  logits += static_cast<size_t>(blockIdx.y) * C * HW;              // blockIdx.y = image of a batch
  label += static_cast<size_t>(blockIdx.y) * HW;
  if (i + 3 < HW && (HW & 3) == 0) {
    float4 best = *reinterpret_cast<const float4*>(logits + i);
    float4 arg = make_float4(0.f, 0.f, 0.f, 0.f);
    for (int c = 1; c < C; ++c) {
      const float4 v = *reinterpret_cast<const float4*>(logits + static_cast<size_t>(c) * HW + i);
      if (v.x > best.x) { best.x = v.x; arg.x = c; }
      if (v.y > best.y) { best.y = v.y; arg.y = c; }
      if (v.z > best.z) { best.z = v.z; arg.z = c; }
      if (v.w > best.w) { best.w = v.w; arg.w = c; }
    }
    *reinterpret_cast<float4*>(label + i) = arg;
  } else {
    for (int k = i; k < HW && k < i + 4; ++k) {
      float best = logits[k];
      int arg = 0;
      for (int c = 1; c < C; ++c) {
        const float v = logits[static_cast<size_t>(c) * HW + k];
        if (v > best) { best = v; arg = c; }
      }
      label[k] = static_cast<float>(arg);
    }
  }
}

int argmax_label_launch(const float* logits, int C, int HW, float* label, cudaStream_t stream) {
  return argmax_label_batched_launch(logits, C, HW, label, 1, stream);
}

int argmax_label_batched_launch(const float* logits, int C, int HW, float* label, int S, cudaStream_t stream) {
  if (C < 1 || HW < 1 || S < 1 || S > 65535) return set_error(RMEM_ERR_ARG, "argmax_label: bad shape");
  const int n = (HW + 3) / 4;
  launch_k(argmax_label_kernel, dim3((n + 255) / 256, S), 256, 0, stream, logits, C, HW, label);
  return check_launch("argmax_label_kernel");
}

}  // namespace rmem
