// Identity embedding of a label map: one_hot_mask + ignore channel + 17x17 stride-16 conv
// (patch_wise_id_bank) + DeAOT's id_norm LayerNorm, as one gather kernel -- reference
// AOTEngine.assign_identity (networks/engines/aot_engine.py:208-232), utils/image.py:69-74,
// AOT.get_id_emb / DeAOT.get_id_emb (networks/models/aot.py:111-114, deaot.py:65-69).
//
// A one-hot input makes the convolution a table lookup: out[token][c] = bias[c] +
// sum over the token's 17x17 window of W[c][class(y,x)][ky][kx].  The weights are repacked as
// row-prefix sums P[class][ky][kx'][c] = sum_{kx < kx'} W[c][class][ky][kx], so a run of equal labels
// in a window row costs two loads instead of one per tap (label maps are piecewise constant).  Behind them sit
// the 2-D prefix sums P2[class][ky'][kx'][c] = sum_{ky < ky', kx < kx'} W[c][class][ky][kx]: rows with the same
// run structure collapse into rectangles.
#include <cuda_bf16.h>
#include "rmem_internal.h"

namespace rmem {

constexpr int kIdK = 17, kIdStride = 16, kIdPad = 8, kIdC = 256;

struct IdEmbedParams {
  const float* label;      // [H][W] label map: 0..max_obj = object id, 255 = ignore, others = nothing
  const float* prefix;     // [n_cls][17][18][256] row prefixes, then [n_cls][18][18][256] 2-D prefixes
  const float* bias;       // [256]
  const float* gamma;      // LayerNorm weight [256] or nullptr (AOT: no norm)
  const float* beta;
  __nv_bfloat16* dst[4];   // up to 4 destinations
  int ld[4], col0[4];
  int n_dst;
  int H, W, h, w, n_cls;   // n_cls = max_obj + 2 (last class = ignore)
};

__global__ void __launch_bounds__(kIdC) id_embed_kernel(const IdEmbedParams p) {
  __shared__ int cls_sh[kIdK][kIdK + 1];
  __shared__ float red[2][kIdC / 32];
  const int tok = blockIdx.x;
  const int ty = tok / p.w, tx = tok - ty * p.w;
  const int c = threadIdx.x;
  for (int i = threadIdx.x; i < kIdK * kIdK; i += kIdC) {
    const int ky = i / kIdK, kx = i - ky * kIdK;
    const int y = ty * kIdStride - kIdPad + ky, x = tx * kIdStride - kIdPad + kx;
    int cls = -1;
    if (y >= 0 && y < p.H && x >= 0 && x < p.W) {
      const int l = __float2int_rn(p.label[static_cast<size_t>(y) * p.W + x]);
      if (l >= 0 && l < p.n_cls - 1) cls = l;
      else if (l == 255) cls = p.n_cls - 1;
    }
    cls_sh[ky][kx] = cls;
  }
  __syncthreads();
  // runs of equal labels per window row, found once per CTA (one thread per row) instead of by every channel
  __shared__ short run_cls[kIdK][kIdK], run_lo[kIdK][kIdK], run_hi[kIdK][kIdK];
  __shared__ int n_runs[kIdK];
  if (threadIdx.x < kIdK) {
    const int ky = threadIdx.x;
    int n = 0, run_start = 0, rc = cls_sh[ky][0];
    for (int kx = 1; kx <= kIdK; ++kx) {
      const int cur = (kx < kIdK) ? cls_sh[ky][kx] : -2;
      if (cur != rc) {
        if (rc >= 0) {
          run_cls[ky][n] = static_cast<short>(rc);
          run_lo[ky][n] = static_cast<short>(run_start);
          run_hi[ky][n] = static_cast<short>(kx);
          ++n;
        }
        run_start = kx;
        rc = cur;
      }
    }
    n_runs[ky] = n;
  }
  __syncthreads();
  // consecutive window rows with identical run lists form a group: its contribution is a rectangle sum, four
  // loads from the 2-D prefix table whatever the group's height (an interior window of a single object is one
  // group with one run); a lone row uses the row-prefix table (two loads per run)
  __shared__ short grp_lo[kIdK], grp_hi[kIdK];
  __shared__ int n_grp;
  if (threadIdx.x == 0) {
    int n = 0, r0 = 0;
    for (int ky = 1; ky <= kIdK; ++ky) {
      bool same = ky < kIdK && n_runs[ky] == n_runs[r0];
      if (same) {
        for (int r = 0; r < n_runs[r0]; ++r)
          same = same && run_cls[ky][r] == run_cls[r0][r] && run_lo[ky][r] == run_lo[r0][r] &&
                 run_hi[ky][r] == run_hi[r0][r];
      }
      if (!same) {
        if (n_runs[r0] > 0) { grp_lo[n] = static_cast<short>(r0); grp_hi[n] = static_cast<short>(ky); ++n; }
        r0 = ky;
      }
    }
    n_grp = n;
  }
  __syncthreads();
  const float* prefix2 = p.prefix + static_cast<size_t>(p.n_cls) * kIdK * (kIdK + 1) * kIdC;
  float acc = p.bias[c];
  const int ng = n_grp;
  for (int g = 0; g < ng; ++g) {
    const int r0 = grp_lo[g], r1 = grp_hi[g];
    const int n = n_runs[r0];
    if (r1 - r0 == 1) {
      for (int r = 0; r < n; ++r) {
        const float* base = p.prefix + (static_cast<size_t>(run_cls[r0][r]) * kIdK + r0) * (kIdK + 1) * kIdC + c;
        acc += base[run_hi[r0][r] * kIdC] - base[run_lo[r0][r] * kIdC];
      }
    } else {
      for (int r = 0; r < n; ++r) {
        const float* base = prefix2 + static_cast<size_t>(run_cls[r0][r]) * (kIdK + 1) * (kIdK + 1) * kIdC + c;
        const int lo = run_lo[r0][r], hi = run_hi[r0][r];
        acc += (base[(r1 * (kIdK + 1) + hi) * kIdC] - base[(r0 * (kIdK + 1) + hi) * kIdC]) -
               (base[(r1 * (kIdK + 1) + lo) * kIdC] - base[(r0 * (kIdK + 1) + lo) * kIdC]);
      }
    }
  }
  float out = acc;
  if (p.gamma) {
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    float s = acc;
    for (int o = 16; o > 0; o >>= 1) s += __shfl_xor_sync(0xffffffffu, s, o);
    if (lane == 0) red[0][warp] = s;
    __syncthreads();
    float mean = 0.f;
    for (int i = 0; i < kIdC / 32; ++i) mean += red[0][i];
    mean *= (1.f / kIdC);
    const float d = acc - mean;
    float q = d * d;
    for (int o = 16; o > 0; o >>= 1) q += __shfl_xor_sync(0xffffffffu, q, o);
    if (lane == 0) red[1][warp] = q;
    __syncthreads();
    float var = 0.f;
    for (int i = 0; i < kIdC / 32; ++i) var += red[1][i];
    out = d * rsqrtf(var * (1.f / kIdC) + 1e-5f) * p.gamma[c] + p.beta[c];
  }
  const __nv_bfloat16 ob = __float2bfloat16(out);
  for (int i = 0; i < p.n_dst; ++i) p.dst[i][static_cast<size_t>(tok) * p.ld[i] + p.col0[i] + c] = ob;
}

int id_embed_launch(const float* label, int H, int W, const float* prefix, const float* bias,
                    const float* gamma, const float* beta, int n_cls, void* const* dst, const int* ld,
                    const int* col0, int n_dst, int h, int w, cudaStream_t stream) {
  if (n_dst < 1 || n_dst > 4 || n_cls < 2 || h != (H + 2 * kIdPad - kIdK) / kIdStride + 1 ||
      w != (W + 2 * kIdPad - kIdK) / kIdStride + 1)
    return set_error(RMEM_ERR_ARG, "id_embed: bad destination count / class count / token grid");
  IdEmbedParams p;
  p.label = label; p.prefix = prefix; p.bias = bias; p.gamma = gamma; p.beta = beta;
  p.n_dst = n_dst; p.H = H; p.W = W; p.h = h; p.w = w; p.n_cls = n_cls;
  for (int i = 0; i < 4; ++i) {
    p.dst[i] = i < n_dst ? static_cast<__nv_bfloat16*>(dst[i]) : nullptr;
    p.ld[i] = i < n_dst ? ld[i] : 0;
    p.col0[i] = i < n_dst ? col0[i] : 0;
  }
  id_embed_kernel<<<h * w, kIdC, 0, stream>>>(p);
  return check_launch("id_embed_kernel");
}

}  // namespace rmem
