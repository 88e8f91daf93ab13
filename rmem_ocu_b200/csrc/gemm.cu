// C[M,N] = A[M,K] * W[N,K]^T (+ bias) with a fused, column-segmented epilogue -- the Linear layers
// of the propagation blocks (reference nn.Linear calls in GatedPropagationModule.forward,
// layers/transformer.py:1104-1123, 1212-1232; GatedPropagation projections, attention.py:152-172,
// 209; fuse_key_value_id, transformer.py:1238-1244) with their surrounding element-wise ops
// (SiLU, residual add, output split / concat) folded into the epilogue.
//
// bf16 operands (both K-major), fp32 accumulation in TMEM.  128x128 output tile per CTA,
// 64-wide K blocks through a 4-stage TMA ring, tcgen05.mma issued by one thread, four epilogue
// warps read the accumulator with tcgen05.ld (one row per thread).
#include "sm100_ptx.cuh"
#include "rmem_internal.h"

namespace rmem {
using namespace sm100;

constexpr int kGM = 128, kGN = 128, kGK = 64, kGStages = 4;
constexpr uint32_t kGABytes = kGM * kGK * 2;   // 16 KB
constexpr uint32_t kGBBytes = kGN * kGK * 2;   // 16 KB
constexpr uint32_t kGStageBytes = kGABytes + kGBBytes;
constexpr uint32_t kGSmem = kGStages * kGStageBytes + 1024 + 128;
constexpr int kGThreads = 192;

struct GemmParams {
  int M, N, K;
  const float* bias;   // [N] or nullptr
  int n_seg;
  rmem_gemm_segment seg[RMEM_GEMM_MAX_SEGMENTS];
};

__device__ __forceinline__ float silu_f(float x) { return x / (1.f + __expf(-x)); }

__global__ void __launch_bounds__(kGThreads, 1)
gemm_bf16_tn_kernel(const __grid_constant__ CUtensorMap tm_a, const __grid_constant__ CUtensorMap tm_w,
                    const GemmParams p) {
  extern __shared__ uint8_t smem_raw[];
  const uint32_t raw_addr = smem_u32(smem_raw);
  uint8_t* smem = smem_raw + ((1024u - (raw_addr & 1023u)) & 1023u);
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem + kGStages * kGStageBytes);
  uint64_t* full = bars;                  // [4]
  uint64_t* empty = bars + kGStages;      // [4]
  uint64_t* acc_full = bars + 2 * kGStages;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + 2 * kGStages + 1);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int m0 = blockIdx.y * kGM, n0 = blockIdx.x * kGN;
  const int kb = (p.K + kGK - 1) / kGK;

  if (warp == 0 && lane == 0) {
    tma_prefetch_desc(&tm_a);
    tma_prefetch_desc(&tm_w);
    for (int i = 0; i < kGStages; ++i) { mbar_init(full + i, 1); mbar_init(empty + i, 1); }
    mbar_init(acc_full, 1);
    mbar_fence_init();
  }
  if (warp == 1) tmem_alloc<128>(tmem_slot);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;

  if (warp == 0) {
    if (lane == 0) {
      for (int k = 0; k < kb; ++k) {
        const int st = k % kGStages;
        if (k >= kGStages) mbar_wait(empty + st, ((k / kGStages) - 1) & 1);
        uint8_t* a = smem + st * kGStageBytes;
        mbar_expect_tx(full + st, kGStageBytes);
        tma_load_2d(a, &tm_a, full + st, k * kGK, m0);
        tma_load_2d(a + kGABytes, &tm_w, full + st, k * kGK, n0);
      }
    }
  } else if (warp == 1) {
    if (lane == 0) {
      constexpr uint32_t idesc = make_idesc_bf16(kGM, kGN, 0, 0);
      for (int k = 0; k < kb; ++k) {
        const int st = k % kGStages;
        mbar_wait(full + st, (k / kGStages) & 1);
        tc_fence_after();
        const uint32_t a_addr = smem_u32(smem + st * kGStageBytes);
        const uint32_t b_addr = a_addr + kGABytes;
#pragma unroll
        for (int ks = 0; ks < kGK / 16; ++ks) {
          mma_bf16_ss(tmem_base, make_desc_sw128(a_addr + ks * 32, 16, 1024),
                      make_desc_sw128(b_addr + ks * 32, 16, 1024), idesc, (k > 0) || (ks > 0));
        }
        tc_commit(empty + st);
      }
      tc_commit(acc_full);
    }
  } else {
    // ---------------------------------------------------------------- epilogue (4 warps, 32 rows each)
    // Phase 1: the thread that owns TMEM lane r (= output row) reads 32 accumulator columns, adds the
    // bias and applies the activation in registers, and parks them in a per-warp shared-memory tile.
    // Phase 2: the warp re-reads the tile transposed -- half a warp per row, two adjacent columns per
    // lane -- so every global access (store, or load + add + store for the accumulate mode) is a
    // contiguous 64 B / 128 B run instead of 32 scattered rows.
    const int quarter = warp & 3;
    const int row_l = quarter * 32 + lane;
    const uint32_t lane_addr = static_cast<uint32_t>(quarter * 32) << 16;
    float* tile = reinterpret_cast<float*>(smem) + quarter * (32 * 33);      // [32 rows][33] (stage 0 is idle)
    float* bias_s = reinterpret_cast<float*>(smem + kGStageBytes);           // [128]        (stage 1 is idle)
    mbar_wait(acc_full, 0);
    tc_fence_after();
    // every MMA has completed, so every TMA load has landed and the stage buffers are free
    if (p.bias) {
      const int cg = n0 + row_l;
      bias_s[row_l] = cg < p.N ? __ldg(p.bias + cg) : 0.f;
    } else {
      bias_s[row_l] = 0.f;
    }
    named_bar_sync(1, 128);
    const int c2 = 2 * (lane & 15);        // column pair of this lane in phase 2
    const int rpar = lane >> 4;            // row parity of this half-warp
#pragma unroll 1
    for (int cb = 0; cb < kGN / 32; ++cb) {
      const int c0 = n0 + cb * 32;
      if (c0 >= p.N) break;                       // warp-uniform
      // segment lookup (warp-uniform: segments start on multiples of 32 columns)
      int si = -1;
      for (int sgi = 0; sgi < p.n_seg; ++sgi)
        if (c0 >= p.seg[sgi].col_begin && c0 < p.seg[sgi].col_end) si = sgi;
      if (si < 0) continue;
      const rmem_gemm_segment sg = p.seg[si];
      int ncol = sg.col_end - c0;
      if (ncol > 32) ncol = 32;
      uint32_t r[32];
      tmem_ld32(tmem_base + lane_addr + cb * 32, r);
      tmem_ld_wait();
      __syncwarp();                               // phase 2 of the previous block has finished with `tile`
#pragma unroll
      for (int i = 0; i < 32; ++i) {
        float x = __uint_as_float(r[i]) + bias_s[cb * 32 + i];
        if (sg.act == RMEM_ACT_SILU) x = __fdividef(x, 1.f + __expf(-x));
        tile[lane * 33 + i] = x;
      }
      __syncwarp();
      const bool ok0 = c2 < ncol, ok1 = c2 + 1 < ncol;
      const size_t col_off = static_cast<size_t>(sg.dst_col0 + (c0 - sg.col_begin) + c2);
#pragma unroll
      for (int it = 0; it < 16; ++it) {
        const int rl = 2 * it + rpar;
        const int row = m0 + quarter * 32 + rl;
        const float v0 = tile[rl * 33 + c2], v1 = tile[rl * 33 + c2 + 1];
        if (row >= p.M || !ok0) continue;
        const size_t off = static_cast<size_t>(row) * sg.ld + col_off;
        if (sg.mode == RMEM_OUT_BF16) {
          __nv_bfloat16* d = static_cast<__nv_bfloat16*>(sg.dst) + off;
          if (ok1) *reinterpret_cast<uint32_t*>(d) = pack_bf16x2(v0, v1);
          else d[0] = __float2bfloat16(v0);
        } else {
          float* d = static_cast<float*>(sg.dst) + off;
          if (ok1) {
            float2 o = make_float2(v0, v1);
            if (sg.mode == RMEM_OUT_F32_ACCUM) {
              const float2 old = *reinterpret_cast<const float2*>(d);
              o.x += old.x; o.y += old.y;
            }
            *reinterpret_cast<float2*>(d) = o;
          } else {
            d[0] = (sg.mode == RMEM_OUT_F32_ACCUM) ? d[0] + v0 : v0;
          }
        }
      }
    }
    tc_fence_before();
  }
  __syncthreads();
  if (warp == 1) {
    tc_fence_after();
    tmem_dealloc<128>(tmem_base);
  }
}

static int encode_2d(CUtensorMap* map, const void* base, uint64_t inner, uint64_t outer,
                     uint64_t ld_bytes, uint32_t box_outer) {
  auto fn = get_tensor_map_encoder();
  if (!fn) return set_error(RMEM_ERR_DRIVER, "cuTensorMapEncodeTiled entry point not found");
  const uint64_t dims[2] = {inner, outer};
  const uint64_t strides[1] = {ld_bytes};
  const uint32_t box[2] = {64, box_outer};
  const uint32_t es[2] = {1, 1};
  CUresult r = fn(map, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, const_cast<void*>(base), dims, strides,
                  box, es, CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B,
                  CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) return set_error(RMEM_ERR_DRIVER, "cuTensorMapEncodeTiled (gemm) failed");
  return RMEM_OK;
}

int gemm_launch(const void* a, int lda, const void* w, int ldw, const float* bias, int M, int N, int K,
                const rmem_gemm_segment* segs, int n_seg, cudaStream_t stream) {
  if (M < 1 || N < 1 || K < 1 || (lda % 8) || (ldw % 8) || n_seg < 1 || n_seg > RMEM_GEMM_MAX_SEGMENTS)
    return set_error(RMEM_ERR_ARG, "gemm: bad shape / leading dimension / segment count");
  GemmParams p;
  p.M = M; p.N = N; p.K = K; p.bias = bias; p.n_seg = n_seg;
  for (int s = 0; s < n_seg; ++s) {
    p.seg[s] = segs[s];
    if (segs[s].col_begin % 32 || segs[s].col_end <= segs[s].col_begin || segs[s].col_end > N ||
        !segs[s].dst)
      return set_error(RMEM_ERR_ARG, "gemm: segment must start on a multiple of 32 columns");
    if ((segs[s].ld % 2) || (segs[s].dst_col0 % 2) ||
        (reinterpret_cast<uintptr_t>(segs[s].dst) % (segs[s].mode == RMEM_OUT_BF16 ? 4 : 8)))
      return set_error(RMEM_ERR_ARG, "gemm: destination ld / column offset must be even and the base 8-byte aligned");
  }
  CUtensorMap tm_a, tm_w;
  if (int e = encode_2d(&tm_a, a, K, M, static_cast<uint64_t>(lda) * 2, kGM)) return e;
  if (int e = encode_2d(&tm_w, w, K, N, static_cast<uint64_t>(ldw) * 2, kGN)) return e;
  static bool attr_set = false;
  if (!attr_set) {
    cudaError_t e = cudaFuncSetAttribute(gemm_bf16_tn_kernel,
                                         cudaFuncAttributeMaxDynamicSharedMemorySize, kGSmem);
    if (e != cudaSuccess) return set_cuda_error(e, "gemm: cudaFuncSetAttribute");
    attr_set = true;
  }
  dim3 grid((N + kGN - 1) / kGN, (M + kGM - 1) / kGM);
  gemm_bf16_tn_kernel<<<grid, kGThreads, kGSmem, stream>>>(tm_a, tm_w, p);
  return check_launch("gemm_bf16_tn_kernel");
}

}  // namespace rmem
