// Normalisation / resampling kernels of the FPN decoding head (reference
// aot_plus/networks/decoders/fpn.py:36-68, ConvGN = Conv2d + GroupNorm(8), layers/basic.py:60-70).
// The convolutions themselves stay library (cuDNN) calls; what PyTorch issues around them for a
// channels-last batch-1 tensor (a GroupNorm reduction on 8 CTAs, a generic bilinear kernel) is replaced
// by two bandwidth-bound kernels over the [H*W][C] view:
//   gn_nhwc_stats / gn_nhwc_apply : GroupNorm over (C/groups channels x H*W pixels) + optional ReLU
//   upsample_add_nhwc             : out = a + bilinear_align_corners(x)   (adapter(shortcut) + upsampled x)
#include <cuda_bf16.h>
#include "rmem_internal.h"

namespace rmem {

// ------------------------------------------------------------------------------------------------
// stats: a thread owns four consecutive channels (one float4, always inside one group since cpg % 4 == 0) of
// every (1024 / C)-th pixel of the CTA's pixel range; partial sums are reduced over the lanes of a group with
// shuffles, over the CTA in shared memory, and added to the fp64 workspace with one atomic pair per
// (CTA, group).  The range per CTA is sized so the grid is ~4 CTAs per SM whatever the resolution.
__global__ void __launch_bounds__(256)
gn_nhwc_stats_kernel(const float* __restrict__ x, int P, int C, int cpg, int rows_per_cta,
                     double* __restrict__ ws) {
  const int tpp = C >> 2;                        // threads per pixel (C = 64, 128, 256 -> 16, 32, 64)
  const int ppb = 256 / tpp;                     // pixels per step
  const int c = (threadIdx.x % tpp) * 4;
  const int ph = threadIdx.x / tpp;
  const int p0 = blockIdx.x * rows_per_cta;
  const int p1 = min(p0 + rows_per_cta, P);
  float s = 0.f, q = 0.f;
#pragma unroll 4
  for (int pix = p0 + ph; pix < p1; pix += ppb) {
    const float4 v = *reinterpret_cast<const float4*>(x + static_cast<size_t>(pix) * C + c);
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

__global__ void __launch_bounds__(256)
gn_nhwc_apply_kernel(const float* __restrict__ x, const float* __restrict__ gamma,
                     const float* __restrict__ beta, float* __restrict__ y, int P, int C, int cpg,
                     int relu, const double* __restrict__ ws) {
  const size_t total4 = static_cast<size_t>(P) * C / 4;
  const double cnt = static_cast<double>(cpg) * P;
  for (size_t i = blockIdx.x * static_cast<size_t>(blockDim.x) + threadIdx.x; i < total4;
       i += static_cast<size_t>(gridDim.x) * blockDim.x) {
    const int c = static_cast<int>((i * 4) % C);      // 4 consecutive channels, same group (cpg % 4 == 0)
    const int g = c / cpg;
    const double mean_d = ws[2 * g] / cnt;
    const double var_d = ws[2 * g + 1] / cnt - mean_d * mean_d;
    const float mean = static_cast<float>(mean_d);
    const float rstd = rsqrtf(static_cast<float>(var_d) + 1e-5f);
    const float4 v = reinterpret_cast<const float4*>(x)[i];
    const float4 ga = *reinterpret_cast<const float4*>(gamma + c);
    const float4 be = *reinterpret_cast<const float4*>(beta + c);
    float4 o;
    o.x = (v.x - mean) * rstd * ga.x + be.x;
    o.y = (v.y - mean) * rstd * ga.y + be.y;
    o.z = (v.z - mean) * rstd * ga.z + be.z;
    o.w = (v.w - mean) * rstd * ga.w + be.w;
    if (relu) { o.x = fmaxf(o.x, 0.f); o.y = fmaxf(o.y, 0.f); o.z = fmaxf(o.z, 0.f); o.w = fmaxf(o.w, 0.f); }
    reinterpret_cast<float4*>(y)[i] = o;
  }
}

int groupnorm_nhwc_launch(const float* x, const float* gamma, const float* beta, float* y, int P, int C,
                          int groups, int relu, void* workspace, cudaStream_t stream) {
  if (groups < 1 || groups > 32 || (C % groups) || (C != 64 && C != 128 && C != 256) || P < 1)
    return set_error(RMEM_ERR_ARG, "groupnorm_nhwc: C must be 64 / 128 / 256");
  const int cpg = C / groups;
  if ((cpg & (cpg - 1)) || cpg < 4 || cpg > 32)
    return set_error(RMEM_ERR_ARG, "groupnorm_nhwc: channels per group must be a power of two in [4, 32]");
  cudaError_t e = cudaMemsetAsync(workspace, 0, sizeof(double) * 2 * groups, stream);
  if (e != cudaSuccess) return set_cuda_error(e, "groupnorm_nhwc: memset");
  int rows = (P + 591) / 592;                    // ~4 CTAs per SM at large P ...
  if (rows < 8) rows = 8;                        // ... but never less than two steps of work per CTA
  gn_nhwc_stats_kernel<<<(P + rows - 1) / rows, 256, 0, stream>>>(x, P, C, cpg, rows,
                                                                 static_cast<double*>(workspace));
  if (int r = check_launch("gn_nhwc_stats_kernel")) return r;
  const size_t total4 = static_cast<size_t>(P) * C / 4;
  int blocks = static_cast<int>((total4 + 255) / 256);
  if (blocks > 148 * 8) blocks = 148 * 8;
  gn_nhwc_apply_kernel<<<blocks, 256, 0, stream>>>(x, gamma, beta, y, P, C, cpg, relu,
                                                   static_cast<const double*>(workspace));
  return check_launch("gn_nhwc_apply_kernel");
}

// ------------------------------------------------------------------------------------------------
// out[y][x][c] = a[y][x][c] + bilinear(x_in)[y][x][c], align_corners = True (F.interpolate, fpn.py:49-60)
__global__ void __launch_bounds__(256)
upsample_add_nhwc_kernel(const float* __restrict__ xin, int hi, int wi, const float* __restrict__ a,
                         float* __restrict__ out, int ho, int wo, int C) {
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
    const float4 v00 = *reinterpret_cast<const float4*>(xin + (static_cast<size_t>(y0) * wi + x0) * C + c);
    const float4 v01 = *reinterpret_cast<const float4*>(xin + (static_cast<size_t>(y0) * wi + x1) * C + c);
    const float4 v10 = *reinterpret_cast<const float4*>(xin + (static_cast<size_t>(y1) * wi + x0) * C + c);
    const float4 v11 = *reinterpret_cast<const float4*>(xin + (static_cast<size_t>(y1) * wi + x1) * C + c);
    float4 o = a ? *reinterpret_cast<const float4*>(a + static_cast<size_t>(pix) * C + c)
                 : make_float4(0.f, 0.f, 0.f, 0.f);
    o.x += hy * (hx * v00.x + lx * v01.x) + ly * (hx * v10.x + lx * v11.x);
    o.y += hy * (hx * v00.y + lx * v01.y) + ly * (hx * v10.y + lx * v11.y);
    o.z += hy * (hx * v00.z + lx * v01.z) + ly * (hx * v10.z + lx * v11.z);
    o.w += hy * (hx * v00.w + lx * v01.w) + ly * (hx * v10.w + lx * v11.w);
    *reinterpret_cast<float4*>(out + static_cast<size_t>(pix) * C + c) = o;
  }
}

int upsample_add_nhwc_launch(const float* xin, int hi, int wi, const float* a, float* out, int ho, int wo,
                             int C, cudaStream_t stream) {
  if ((C % 4) || hi < 1 || wi < 1 || ho < 1 || wo < 1)
    return set_error(RMEM_ERR_ARG, "upsample_add_nhwc: C must be a multiple of 4");
  const size_t total = static_cast<size_t>(ho) * wo * (C / 4);
  int blocks = static_cast<int>((total + 255) / 256);
  if (blocks > 148 * 8) blocks = 148 * 8;
  upsample_add_nhwc_kernel<<<blocks, 256, 0, stream>>>(xin, hi, wi, a, out, ho, wo, C);
  return check_launch("upsample_add_nhwc_kernel");
}

}  // namespace rmem
