// Identity embedding of a label map: one_hot_mask + ignore channel + 17x17 stride-16 conv
// (patch_wise_id_bank) + DeAOT's id_norm LayerNorm, as one gather kernel -- reference
// AOTEngine.assign_identity (networks/engines/aot_engine.py:208-232), utils/image.py:69-74,
// AOT.get_id_emb / DeAOT.get_id_emb (networks/models/aot.py:111-114, deaot.py:65-69).
//
// A one-hot input makes the convolution a table lookup: out[token][c] = bias[c] +
// sum over the token's 17x17 window of W[c][class(y,x)][ky][kx].  The weights are repacked as
// row-prefix sums P[class][ky][kx'][c] = sum_{kx < kx'} W[c][class][ky][kx], so a run of equal labels
// in a window row costs two loads instead of one per tap (label maps are piecewise constant).  Behind them sit
// the 2-D prefix sums P2[class][ky'][kx'][c] = sum_{ky < ky', kx < kx'} W[c][class][ky][kx]: identical rows
// collapse into rectangles.
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

// id_norm LayerNorm over the 256 channels of one token (one channel per thread) + the bf16 stores
__device__ __forceinline__ void finish_token(float acc, const float* gamma, const float* beta, float (*red)[kIdC / 32],
                                             __nv_bfloat16* const* dst, const int* ld, const int* col0, int n_dst) {
  const int c = threadIdx.x, warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  float out = acc;
  if (gamma) {
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
    out = d * rsqrtf(var * (1.f / kIdC) + 1e-5f) * gamma[c] + beta[c];
  }
  const __nv_bfloat16 ob = __float2bfloat16(out);
  for (int i = 0; i < n_dst; ++i) dst[i][static_cast<size_t>(blockIdx.x) * ld[i] + col0[i] + c] = ob;
}

__global__ void __launch_bounds__(kIdC) id_embed_kernel(const IdEmbedParams p) {
  pdl_wait();        // programmatic dependent launch: see rmem_internal.h
  pdl_trigger();
  __shared__ int cls_sh[kIdK][kIdK + 1];
  __shared__ float red[2][kIdC / 32];
  // blockIdx.x = (sequence, token): label maps stacked [S][H][W], destination rows stacked [S][h*w]
  const int seq = blockIdx.x / (p.h * p.w);
  const int tok = blockIdx.x - seq * (p.h * p.w);
  const int ty = tok / p.w, tx = tok - ty * p.w;
  const int c = threadIdx.x;
  const float* label = p.label + static_cast<size_t>(seq) * p.H * p.W;
  for (int i = threadIdx.x; i < kIdK * kIdK; i += kIdC) {
    const int ky = i / kIdK, kx = i - ky * kIdK;
    const int y = ty * kIdStride - kIdPad + ky, x = tx * kIdStride - kIdPad + kx;
    int cls = -1;
    if (y >= 0 && y < p.H && x >= 0 && x < p.W) {
      const int l = __float2int_rn(label[static_cast<size_t>(y) * p.W + x]);
      if (l >= 0 && l < p.n_cls - 1) cls = l;
      else if (l == 255) cls = p.n_cls - 1;
    }
    cls_sh[ky][kx] = cls;
  }
  __syncthreads();
  // The window is cut into rectangles of one class: maximal runs of equal labels in a row, extended downwards
  // over the following rows that are identical to it.  A rectangle costs four loads from the 2-D prefix table
  // whatever its size (an interior window of a single object is one rectangle), a one-row rectangle two loads
  // from the row-prefix table.  The cutting is warp-parallel (ballots over the 17 columns of a row), the list is
  // built in (row, column) order so the summation order -- and the fp32 result -- does not depend on timing.
  __shared__ unsigned run_mask[kIdK];       // bit kx: a run starts at column kx of row ky
  __shared__ int row_new[kIdK];             // row ky differs from row ky - 1 (or ky == 0): it opens a row group
  __shared__ int rect_cnt[kIdK];            // rectangles opened by row ky
  __shared__ int rect_cls[kIdK * kIdK];
  __shared__ uchar4 rect_box[kIdK * kIdK];  // ky0, ky1, kx0, kx1
  __shared__ int n_rect_sh;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  for (int ky = warp; ky < kIdK; ky += kIdC / 32) {
    const bool in = lane < kIdK;
    const int cur = in ? cls_sh[ky][lane] : -3;
    const int left = (in && lane > 0) ? cls_sh[ky][lane - 1] : -2;
    const bool differs_up = in && (ky == 0 || cur != cls_sh[ky - 1][lane]);
    const bool start = in && cur != left;
    const unsigned starts = __ballot_sync(0xffffffffu, start);
    const unsigned opens = __ballot_sync(0xffffffffu, start && cur >= 0);
    const bool is_new = __any_sync(0xffffffffu, differs_up);
    if (lane == 0) {
      run_mask[ky] = starts;
      row_new[ky] = is_new ? 1 : 0;
      rect_cnt[ky] = is_new ? __popc(opens) : 0;
    }
  }
  __syncthreads();
  for (int ky = warp; ky < kIdK; ky += kIdC / 32) {
    const unsigned rows = __ballot_sync(0xffffffffu, lane < kIdK && row_new[lane] != 0);
    int before = (lane < ky) ? rect_cnt[lane] : 0;      // ky <= 16 < 32: one lane per earlier row
    for (int o = 16; o > 0; o >>= 1) before += __shfl_xor_sync(0xffffffffu, before, o);
    if (!((rows >> ky) & 1u)) continue;                  // warp-uniform
    const unsigned later_rows = rows >> (ky + 1);
    const int ky1 = later_rows ? ky + __ffs(later_rows) : kIdK;
    const unsigned starts = run_mask[ky];
    const int cur = lane < kIdK ? cls_sh[ky][lane] : -3;
    const bool opens = ((starts >> lane) & 1u) && cur >= 0;
    const unsigned open_mask = __ballot_sync(0xffffffffu, opens);
    if (opens) {
      const unsigned later = starts >> (lane + 1);
      const int kx1 = later ? lane + __ffs(later) : kIdK;
      const int slot = before + __popc(open_mask & ((1u << lane) - 1u));
      rect_cls[slot] = cur;
      rect_box[slot] = make_uchar4(ky, ky1, lane, kx1);
    }
  }
  if (threadIdx.x == 0) {
    int n = 0;
    for (int ky = 0; ky < kIdK; ++ky) n += rect_cnt[ky];
    n_rect_sh = n;
  }
  __syncthreads();
  const float* prefix2 = p.prefix + static_cast<size_t>(p.n_cls) * kIdK * (kIdK + 1) * kIdC;
  float acc = p.bias[c];
  const int n_rect = n_rect_sh;
  for (int r = 0; r < n_rect; ++r) {
    const uchar4 bx = rect_box[r];
    const int r0 = bx.x, r1 = bx.y, lo = bx.z, hi = bx.w;
    if (r1 - r0 == 1) {
      const float* base = p.prefix + (static_cast<size_t>(rect_cls[r]) * kIdK + r0) * (kIdK + 1) * kIdC + c;
      acc += base[hi * kIdC] - base[lo * kIdC];
    } else {
      const float* base = prefix2 + static_cast<size_t>(rect_cls[r]) * (kIdK + 1) * (kIdK + 1) * kIdC + c;
      acc += (base[(r1 * (kIdK + 1) + hi) * kIdC] - base[(r0 * (kIdK + 1) + hi) * kIdC]) -
             (base[(r1 * (kIdK + 1) + lo) * kIdC] - base[(r0 * (kIdK + 1) + lo) * kIdC]);
    }
  }
  finish_token(acc, p.gamma, p.beta, red, p.dst, p.ld, p.col0, p.n_dst);
}

int id_embed_launch(const float* label, int H, int W, const float* prefix, const float* bias,
                    const float* gamma, const float* beta, int n_cls, void* const* dst, const int* ld,
                    const int* col0, int n_dst, int h, int w, cudaStream_t stream) {
  return id_embed_batched_launch(label, H, W, prefix, bias, gamma, beta, n_cls, dst, ld, col0, n_dst, h, w, 1, stream);
}

int id_embed_batched_launch(const float* label, int H, int W, const float* prefix, const float* bias,
                            const float* gamma, const float* beta, int n_cls, void* const* dst, const int* ld,
                            const int* col0, int n_dst, int h, int w, int S, cudaStream_t stream) {
  if (n_dst < 1 || n_dst > 4 || n_cls < 2 || h != (H + 2 * kIdPad - kIdK) / kIdStride + 1 ||
      w != (W + 2 * kIdPad - kIdK) / kIdStride + 1 || S < 1)
    return set_error(RMEM_ERR_ARG, "id_embed: bad destination count / class count / token grid");
  IdEmbedParams p;
  p.label = label; p.prefix = prefix; p.bias = bias; p.gamma = gamma; p.beta = beta;
  p.n_dst = n_dst; p.H = H; p.W = W; p.h = h; p.w = w; p.n_cls = n_cls;
  for (int i = 0; i < 4; ++i) {
    p.dst[i] = i < n_dst ? static_cast<__nv_bfloat16*>(dst[i]) : nullptr;
    p.ld[i] = i < n_dst ? ld[i] : 0;
    p.col0[i] = i < n_dst ? col0[i] : 0;
  }
  launch_k(id_embed_kernel, S * h * w, kIdC, 0, stream, p);
  return check_launch("id_embed_kernel");
}

// ---------------------------------------------------------------------------------------------
// The same embedding from a one-hot or PROBABILITY mask (aot_engine.py:330-336: a mask that already is
// [max_obj + 1] channels goes to assign_identity as it is): the convolution is then a real weighted sum,
// out[token][c] = bias[c] + sum_{k, ky, kx} m[k][y][x] W[c][k][ky][kx].  One CTA per token stages the token's window of
// every channel in shared memory and walks the taps in (channel, ky, kx) order; taps whose mask value is zero are
// skipped (CTA-uniform), so a one-hot mask costs 289 weight rows and a saturated probability mask little more.
struct IdSoftParams {
  const float* prob;       // [S][n_ch][H][W]
  const float* ignore;     // [S][H][W] or nullptr: channel 0 is multiplied by (ignore == 0), the ignore map is class n_cls - 1
  const float* taps;       // [n_cls][17][17][256] = W[c][k][ky][kx] transposed
  const float* bias;
  const float* gamma;
  const float* beta;
  __nv_bfloat16* dst[4];
  int ld[4], col0[4];
  int n_dst;
  int H, W, h, w, n_ch, n_cls;
};
constexpr int kIdMaxCls = 16;

__global__ void __launch_bounds__(kIdC) id_embed_soft_kernel(const IdSoftParams p) {
  extern __shared__ float win[];                      // [n_cls][17 * 17]
  __shared__ float red[2][kIdC / 32];
  const int seq = blockIdx.x / (p.h * p.w);
  const int tok = blockIdx.x - seq * (p.h * p.w);
  const int ty = tok / p.w, tx = tok - ty * p.w;
  const int c = threadIdx.x;
  const size_t HW = static_cast<size_t>(p.H) * p.W;
  const float* prob = p.prob + static_cast<size_t>(seq) * p.n_ch * HW;
  const float* ign = p.ignore ? p.ignore + static_cast<size_t>(seq) * HW : nullptr;
  constexpr int kWin = kIdK * kIdK;
  for (int i = threadIdx.x; i < p.n_cls * kWin; i += kIdC) {
    const int k = i / kWin, t = i - k * kWin;
    const int ky = t / kIdK, kx = t - ky * kIdK;
    const int y = ty * kIdStride - kIdPad + ky, x = tx * kIdStride - kIdPad + kx;
    float v = 0.f;
    if (y >= 0 && y < p.H && x >= 0 && x < p.W) {
      const size_t px = static_cast<size_t>(y) * p.W + x;
      if (k < p.n_ch) {
        v = prob[static_cast<size_t>(k) * HW + px];
        if (k == 0 && ign && ign[px] != 0.f) v = 0.f;
      } else if (k == p.n_cls - 1 && ign) {
        v = ign[px];
      }
    }
    win[i] = v;
  }
  __syncthreads();
  float acc = p.bias[c];
  const int n = p.n_cls * kWin;
  for (int i = 0; i < n; ++i) {
    const float m = win[i];                           // broadcast read: the branch is CTA-uniform
    if (m != 0.f) acc = fmaf(m, p.taps[static_cast<size_t>(i) * kIdC + c], acc);
  }
  finish_token(acc, p.gamma, p.beta, red, p.dst, p.ld, p.col0, p.n_dst);
}

int id_embed_soft_launch(const float* prob, int n_ch, const float* ignore, int H, int W, const float* taps,
                         const float* bias, const float* gamma, const float* beta, int n_cls, void* const* dst,
                         const int* ld, const int* col0, int n_dst, int h, int w, int S, cudaStream_t stream) {
  if (n_dst < 1 || n_dst > 4 || n_cls < 2 || n_cls > kIdMaxCls || n_ch < 1 || n_ch > n_cls ||
      h != (H + 2 * kIdPad - kIdK) / kIdStride + 1 || w != (W + 2 * kIdPad - kIdK) / kIdStride + 1 || S < 1)
    return set_error(RMEM_ERR_ARG, "id_embed_soft: bad destination count / channel count / token grid");
  IdSoftParams p;
  p.prob = prob; p.ignore = ignore; p.taps = taps; p.bias = bias; p.gamma = gamma; p.beta = beta;
  p.n_dst = n_dst; p.H = H; p.W = W; p.h = h; p.w = w; p.n_ch = n_ch; p.n_cls = n_cls;
  for (int i = 0; i < 4; ++i) {
    p.dst[i] = i < n_dst ? static_cast<__nv_bfloat16*>(dst[i]) : nullptr;
    p.ld[i] = i < n_dst ? ld[i] : 0;
    p.col0[i] = i < n_dst ? col0[i] : 0;
  }
  const size_t smem = static_cast<size_t>(n_cls) * kIdK * kIdK * sizeof(float);     // <= 18.5 KB
  id_embed_soft_kernel<<<S * h * w, kIdC, smem, stream>>>(p);
  return check_launch("id_embed_soft_kernel");
}

}  // namespace rmem
