// C[M,N] = A[M,K] * W[N,K]^T (+ bias) with a fused, column-segmented epilogue -- the Linear layers
// of the propagation blocks (reference nn.Linear calls in GatedPropagationModule.forward,
// layers/transformer.py:1104-1123, 1212-1232; GatedPropagation projections, attention.py:152-172,
// 209; fuse_key_value_id, transformer.py:1238-1244) with their surrounding element-wise ops
// (SiLU, residual add, output split / concat) folded into the epilogue.
//
// bf16 operands (both K-major), fp32 accumulation in TMEM.  128 x {128, 64} output tile per CTA,
// 64-wide K blocks through a 4-stage TMA ring, tcgen05.mma issued by one thread, eight epilogue
// warps read the accumulator with tcgen05.ld (one row per thread) and store through a transposing
// shared-memory tile so global accesses are contiguous per row.
#include "sm100_ptx.cuh"
#include "rmem_internal.h"

namespace rmem {
using namespace sm100;

constexpr int kGM = 128, kGK = 64;
constexpr uint32_t kGABytes = kGM * kGK * 2;   // 16 KB
constexpr int kGEpiWarps = 8;                  // two per TMEM lane quarter, each owning half of the tile's columns
constexpr int kGThreads = 64 + 32 * kGEpiWarps;
constexpr int kGTileLd = 36;                   // floats per staged row: 16-byte aligned, conflict-free for 128-bit accesses

template <int BN, int kGStages>
struct GemmCfg {
  static constexpr uint32_t kBBytes = BN * kGK * 2;
  static constexpr uint32_t kStageBytes = kGABytes + kBBytes;
  static constexpr uint32_t kEpiBytes = kGEpiWarps * 32 * kGTileLd * 4;   // staging tiles: reuse the (idle) ring
  static_assert(kEpiBytes <= kGStages * kStageBytes, "epilogue staging must fit in the pipeline ring");
  static constexpr uint32_t kSmem = kGStages * kStageBytes + BN * 4 + 1024 + 128;
};

struct GemmParams {
  int M, N, K;
  const float* bias;   // [N] or nullptr
  int n_seg;
  int ab_f16;          // A and W are IEEE fp16 instead of bf16 (the backbone's feature maps feeding the adapters)
  rmem_gemm_segment seg[RMEM_GEMM_MAX_SEGMENTS];
};

// BN = 128 for wide outputs, 64 when that would leave most SMs idle (M is only ~13 row tiles on this path).
// Three stages keep two (BN 128) or three (BN 64) CTAs on an SM so main loops and epilogues overlap; the
// long-K narrow GEMMs (fewer CTAs than SMs, 16-32 K blocks each) instead get a six-deep ring to cover TMA latency.
// (Split-K over two CTAs per tile with atomic fp32 reductions was tried for the K = 2048 projections that accumulate
// into the residual streams: 7.5 / 11.2 us became 10.0 / 12.5 us -- the K loop is not what these launches wait for.)
template <int BN, int kGStages>
__global__ void __launch_bounds__(kGThreads, 1)
gemm_bf16_tn_kernel(const __grid_constant__ CUtensorMap tm_a, const __grid_constant__ CUtensorMap tm_w,
                    const GemmParams p) {
  using Cfg = GemmCfg<BN, kGStages>;
  extern __shared__ uint8_t smem_raw[];
  const uint32_t raw_addr = smem_u32(smem_raw);
  uint8_t* smem = smem_raw + ((1024u - (raw_addr & 1023u)) & 1023u);
  float* epi = reinterpret_cast<float*>(smem);          // valid once acc_full fired: every stage is idle then
  float* bias_s = reinterpret_cast<float*>(smem + kGStages * Cfg::kStageBytes);
  uint64_t* bars = reinterpret_cast<uint64_t*>(bias_s + BN);
  uint64_t* full = bars;                  // [kGStages]
  uint64_t* empty = bars + kGStages;      // [kGStages]
  uint64_t* acc_full = bars + 2 * kGStages;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + 2 * kGStages + 1);

  // warp index made warp-uniform (shuffle from lane 0): the MMA warp below then keeps its descriptors in uniform
  // registers and each tcgen05.mma is ONE instruction -- inside an `if (lane == 0)` region the compiler wraps every
  // MMA in an ELECT / R2UR.BROADCAST loop (~90 cycles per MMA, measured on the attention kernel)
  const int warp = __shfl_sync(0xffffffffu, static_cast<int>(threadIdx.x >> 5), 0), lane = threadIdx.x & 31;
  const int m0 = blockIdx.y * kGM, n0 = blockIdx.x * BN;
  const int kb = (p.K + kGK - 1) / kGK;

  if (warp == 0 && lane == 0) {
    tma_prefetch_desc(&tm_a);
    tma_prefetch_desc(&tm_w);
    for (int i = 0; i < kGStages; ++i) { mbar_init(full + i, 1); mbar_init(empty + i, 1); }
    mbar_init(acc_full, 1);
    mbar_fence_init();
  }
  if (warp == 1) tmem_alloc<BN>(tmem_slot);
  if (warp >= 2) {
    // bias of this column tile, staged once (the main loop is running meanwhile)
    const int t = threadIdx.x - 64;
    if (t < BN) {
      const int cg = n0 + t;
      bias_s[t] = (p.bias && cg < p.N) ? __ldg(p.bias + cg) : 0.f;
    }
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = __shfl_sync(0xffffffffu, *tmem_slot, 0);
  pdl_wait();       // everything above touched parameters and the bias (a weight) only
  pdl_trigger();

  if (warp == 0) {
    if (lane == 0) {
      for (int k = 0; k < kb; ++k) {
        const int st = k % kGStages;
        if (k >= kGStages) mbar_wait(empty + st, ((k / kGStages) - 1) & 1);
        uint8_t* a = smem + st * Cfg::kStageBytes;
        mbar_expect_tx(full + st, Cfg::kStageBytes);
        tma_load_2d(a, &tm_a, full + st, k * kGK, m0);
        tma_load_2d(a + kGABytes, &tm_w, full + st, k * kGK, n0);
      }
    }
  } else if (warp == 1) {
    // the whole warp runs the loop (and the waits); one elected lane issues
    const uint32_t idesc = make_idesc_16(kGM, BN, 0, 0, p.ab_f16);
    const uint64_t a_desc0 = make_desc_sw128(smem_u32(smem), 16, 1024);
    for (int k = 0; k < kb; ++k) {
      const int st = k % kGStages;
      mbar_wait(full + st, (k / kGStages) & 1);
      tc_fence_after();
      const uint64_t a_desc = a_desc0 + static_cast<uint64_t>((st * Cfg::kStageBytes) >> 4);
      const uint64_t b_desc = a_desc + static_cast<uint64_t>(kGABytes >> 4);
      if (elect_one()) {
#pragma unroll
        for (int ks = 0; ks < kGK / 16; ++ks)
          mma_bf16_ss(tmem_base, a_desc + static_cast<uint64_t>((ks * 32) >> 4),
                      b_desc + static_cast<uint64_t>((ks * 32) >> 4), idesc, (k > 0) || (ks > 0));
        tc_commit(empty + st);
      }
      __syncwarp();
    }
    if (elect_one()) tc_commit(acc_full);
    __syncwarp();
  } else {
    // ---------------------------------------------------------------- epilogue (8 warps)
    // Warp (quarter, half) owns TMEM lanes [32 quarter, +32) = 32 output rows and half of the tile's columns.
    // Per 32-column block: phase 1, the thread that owns row r reads its accumulator columns, adds the bias,
    // applies the activation and parks them in the warp's staging tile with 128-bit stores; phase 2, the warp
    // re-reads the tile with eight lanes per row (four columns each), so every global access is a contiguous
    // 64 B / 128 B run per row; the accumulate mode issues all its loads before the first add.
    const int ew = warp - 2;
    const int quarter = warp & 3;
    const int half = ew >> 2;
    const uint32_t lane_addr = static_cast<uint32_t>(quarter * 32) << 16;
    float* tile = epi + ew * (32 * kGTileLd);
    constexpr int kCbPerWarp = BN / 64;            // 32-column blocks per warp
    const int r_sub = lane >> 3;                   // phase 2: row within a group of four
    const int c4 = (lane & 7) * 4;                 // phase 2: first of this lane's four columns
    mbar_wait(acc_full, 0);
    tc_fence_after();
#pragma unroll 1
    for (int cbi = 0; cbi < kCbPerWarp; ++cbi) {
      const int cb = half * kCbPerWarp + cbi;
      const int c0 = n0 + cb * 32;
      if (c0 >= p.N) break;                       // warp-uniform
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
      for (int i = 0; i < 32; i += 4) {
        float4 x;
        x.x = __uint_as_float(r[i]) + bias_s[cb * 32 + i];
        x.y = __uint_as_float(r[i + 1]) + bias_s[cb * 32 + i + 1];
        x.z = __uint_as_float(r[i + 2]) + bias_s[cb * 32 + i + 2];
        x.w = __uint_as_float(r[i + 3]) + bias_s[cb * 32 + i + 3];
        if (sg.act == RMEM_ACT_SILU) {
          x.x = __fdividef(x.x, 1.f + __expf(-x.x));
          x.y = __fdividef(x.y, 1.f + __expf(-x.y));
          x.z = __fdividef(x.z, 1.f + __expf(-x.z));
          x.w = __fdividef(x.w, 1.f + __expf(-x.w));
        }
        *reinterpret_cast<float4*>(tile + lane * kGTileLd + i) = x;
      }
      __syncwarp();
      const size_t col_off = static_cast<size_t>(sg.dst_col0 + (c0 - sg.col_begin) + c4);
      const int nc = ncol - c4;                    // valid columns of this lane's four (<= 0: none)
      const int row_base = m0 + quarter * 32 + r_sub;
      const bool vec_ok = nc >= 4 && ((sg.ld | sg.dst_col0) & 3) == 0 &&
                          (reinterpret_cast<uintptr_t>(sg.dst) & 15) == 0;     // 16 B (f32) / 8 B (bf16) accesses
      if (sg.mode == RMEM_OUT_BF16) {
        __nv_bfloat16* base = static_cast<__nv_bfloat16*>(sg.dst) + col_off;
#pragma unroll
        for (int it = 0; it < 8; ++it) {
          const int row = row_base + it * 4;
          const float4 v = *reinterpret_cast<const float4*>(tile + (it * 4 + r_sub) * kGTileLd + c4);
          if (row >= p.M || nc <= 0) continue;
          __nv_bfloat16* d = base + static_cast<size_t>(row) * sg.ld;
          if (vec_ok) {
            *reinterpret_cast<uint2*>(d) = make_uint2(pack_bf16x2(v.x, v.y), pack_bf16x2(v.z, v.w));
          } else {
            const float vv[4] = {v.x, v.y, v.z, v.w};
            for (int e = 0; e < 4 && e < nc; ++e) d[e] = __float2bfloat16(vv[e]);
          }
        }
      } else {
        float* base = static_cast<float*>(sg.dst) + col_off;
        const bool accum = sg.mode == RMEM_OUT_F32_ACCUM;
        float4 old[8];
        if (accum && vec_ok) {
#pragma unroll
          for (int it = 0; it < 8; ++it) {
            const int row = row_base + it * 4;
            old[it] = row < p.M ? *reinterpret_cast<const float4*>(base + static_cast<size_t>(row) * sg.ld)
                                : make_float4(0.f, 0.f, 0.f, 0.f);
          }
        }
#pragma unroll
        for (int it = 0; it < 8; ++it) {
          const int row = row_base + it * 4;
          float4 v = *reinterpret_cast<const float4*>(tile + (it * 4 + r_sub) * kGTileLd + c4);
          if (row >= p.M || nc <= 0) continue;
          float* d = base + static_cast<size_t>(row) * sg.ld;
          if (vec_ok) {
            if (accum) { v.x += old[it].x; v.y += old[it].y; v.z += old[it].z; v.w += old[it].w; }
            *reinterpret_cast<float4*>(d) = v;
          } else {
            const float vv[4] = {v.x, v.y, v.z, v.w};
            for (int e = 0; e < 4 && e < nc; ++e) d[e] = accum ? d[e] + vv[e] : vv[e];
          }
        }
      }
    }
    tc_fence_before();
  }
  __syncthreads();
  if (warp == 1) {
    tc_fence_after();
    tmem_dealloc<BN>(tmem_base);
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
  return gemm_ex_launch(a, lda, w, ldw, bias, M, N, K, segs, n_seg, 0, stream);
}

int gemm_ex_launch(const void* a, int lda, const void* w, int ldw, const float* bias, int M, int N, int K,
                   const rmem_gemm_segment* segs, int n_seg, int ab_f16, cudaStream_t stream) {
  if (M < 1 || N < 1 || K < 1 || (lda % 8) || (ldw % 8) || n_seg < 1 || n_seg > RMEM_GEMM_MAX_SEGMENTS)
    return set_error(RMEM_ERR_ARG, "gemm: bad shape / leading dimension / segment count");
  GemmParams p;
  p.M = M; p.N = N; p.K = K; p.bias = bias; p.n_seg = n_seg; p.ab_f16 = ab_f16 ? 1 : 0;
  for (int s = 0; s < n_seg; ++s) {
    p.seg[s] = segs[s];
    if (segs[s].col_begin % 32 || segs[s].col_end <= segs[s].col_begin || segs[s].col_end > N ||
        !segs[s].dst)
      return set_error(RMEM_ERR_ARG, "gemm: segment must start on a multiple of 32 columns");
    if ((segs[s].ld % 2) || (segs[s].dst_col0 % 2) ||
        (reinterpret_cast<uintptr_t>(segs[s].dst) % (segs[s].mode == RMEM_OUT_BF16 ? 4 : 8)))
      return set_error(RMEM_ERR_ARG, "gemm: destination ld / column offset must be even and the base 8-byte aligned");
  }
  // column-tile width: 128 unless that leaves fewer CTAs than SMs (the row dimension gives only ceil(M/128) tiles)
  const int m_tiles = (M + kGM - 1) / kGM;
  const bool narrow = m_tiles * ((N + 127) / 128) < 120;
  const int bn = narrow ? 64 : 128;
  CUtensorMap tm_a, tm_w;
  if (int e = encode_2d(&tm_a, a, K, M, static_cast<uint64_t>(lda) * 2, kGM)) return e;
  if (int e = encode_2d(&tm_w, w, K, N, static_cast<uint64_t>(ldw) * 2, bn)) return e;
  const bool deep = narrow && K >= 1024;
  static SmemOptIn optin_128_3, optin_64_3, optin_64_6;
  {
    int r = deep ? ensure_dynamic_smem(reinterpret_cast<const void*>(gemm_bf16_tn_kernel<64, 6>), &optin_64_6,
                                       GemmCfg<64, 6>::kSmem, "gemm: cudaFuncSetAttribute")
          : narrow ? ensure_dynamic_smem(reinterpret_cast<const void*>(gemm_bf16_tn_kernel<64, 3>), &optin_64_3,
                                         GemmCfg<64, 3>::kSmem, "gemm: cudaFuncSetAttribute")
                   : ensure_dynamic_smem(reinterpret_cast<const void*>(gemm_bf16_tn_kernel<128, 3>), &optin_128_3,
                                         GemmCfg<128, 3>::kSmem, "gemm: cudaFuncSetAttribute");
    if (r) return r;
  }
  dim3 grid((N + bn - 1) / bn, m_tiles);
  if (deep)
    launch_k(gemm_bf16_tn_kernel<64, 6>, grid, kGThreads, GemmCfg<64, 6>::kSmem, stream, tm_a, tm_w, p);
  else if (narrow)
    launch_k(gemm_bf16_tn_kernel<64, 3>, grid, kGThreads, GemmCfg<64, 3>::kSmem, stream, tm_a, tm_w, p);
  else
    launch_k(gemm_bf16_tn_kernel<128, 3>, grid, kGThreads, GemmCfg<128, 3>::kSmem, stream, tm_a, tm_w, p);
  return check_launch("gemm_bf16_tn_kernel");
}

}  // namespace rmem
