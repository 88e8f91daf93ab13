// Short-term local attention (DeAOT), one kernel: 15x15 window over the previous frame, 1 head, d_att 128,
// value width 1024, relative-position bias, softmax over the 225 taps, aggregation of V and the gate --
// LocalGatedPropagation.forward, reference layers/attention.py:289-361 (scores :314-346, aggregation :350-353,
// gate :355) -- in the flash-attention form of csrc/lt_attn.cu, so the probabilities never leave the SM
// (the two-kernel form of csrc/local_attn.cu writes them to a banded buffer and reads them back 8 times).
//
// A CTA owns an 8 x 16 query tile (128 queries = the 128 TMEM lanes) and 128 value channels.  Its 22 x 30 key
// halo is cut into 6 blocks of four halo rows (4 x 32 key slots, 128 keys); per block
//   TMA     : K block [128 keys][128 ch] and V block [128 keys][128 ch] as 4-D boxes straight out of the
//             token-major maps; out-of-image keys are zero-filled (= the reference's F.pad) and masked below
//   tcgen05 : S_b = Q K_b^T (128 x 128, TMEM, double buffered), then O += P_b V_b with P_b read from TMEM
//   softmax : 16 warps, four threads per query; a thread owns ONE halo row of the block (32 key slots), adds the
//             relative-position bias rel[q][dy*15 + dx], masks slots outside the query's window or the image
//             (-1e8 in the reference: probability exactly 0 in fp32), keeps a running max / sum with lazy
//             rescaling of O, and writes bf16 P over the S tile.
// Epilogue: O / l * gate -> bf16.  Every value slice recomputes S (tensor work is not what bounds this op).
#include <cuda_fp16.h>
#include "sm100_ptx.cuh"
#include "rmem_internal.h"

namespace rmem {
using namespace sm100;

namespace lf {
constexpr int kWin = 15, kMaxDis = 7;
constexpr int kTileH = 8, kTileW = 16;                  // 128 queries
constexpr int kHaloH = kTileH + 2 * kMaxDis;            // 22 halo rows
constexpr int kBoxW = 32;                               // 30 halo columns + 2
constexpr int kBlockRows = 4;                           // halo rows per key block
constexpr int kBlockKeys = kBlockRows * kBoxW;          // 128 key slots per block
constexpr int kBlocks = (kHaloH + kBlockRows - 1) / kBlockRows;   // 6 (the last block's rows 22, 23 are masked)
constexpr int kDV = 128;                                // value channels per CTA
constexpr int kStages = 2;
constexpr uint32_t kQBytes = 128 * 128 * 2;             // [2 cblk][128 q][64 ch]
constexpr uint32_t kKBytes = kBlockKeys * 128 * 2;      // [2 cblk][128 keys][64 ch]
constexpr uint32_t kVBytes = kBlockKeys * kDV * 2;      // [2 cblk][128 keys][64 ch]
constexpr int kSoftmaxWarps = 16;
constexpr int kParts = kSoftmaxWarps / 4;               // threads per query
constexpr uint32_t kXchBytes = 2 * kParts * 128 * 4;    // row-max exchange, one buffer per S buffer
constexpr int kRelLd = 232;                             // halves per query row of the staged bias tile (225 + pad)
constexpr uint32_t kRelBytes = 128 * kRelLd * 2;        // 58 KB: rel[q][tap] * log2(e) as fp16
constexpr uint32_t kSmem = kQBytes + kStages * (kKBytes + kVBytes) + kRelBytes + kXchBytes + 256 + 1024;
static_assert(kSmem <= 232448, "exceeds the 227 KB shared memory of an sm_100 CTA");
constexpr int kThreads = 64 + 32 * kSoftmaxWarps;
constexpr uint32_t kTmemCols = 512;                     // S: 2 x 128, O: 128
constexpr uint32_t kTmemS = 0, kTmemO = 256;
constexpr float kRescale = 8.0f;

struct Params {
  const float* rel;            // [h*w][ld_rel] relative-position bias per query and tap (natural units)
  int ld_rel;
  const __nv_bfloat16* gate;   // [h*w][1024] or nullptr
  __nv_bfloat16* out;          // [h*w][1024]
  int h, w;
  float scale_log2;            // log2(e) / sqrt(128)
};
}  // namespace lf

__global__ void __launch_bounds__(lf::kThreads, 1)
local_attn_kernel(const __grid_constant__ CUtensorMap tm_q, const __grid_constant__ CUtensorMap tm_k,
                  const __grid_constant__ CUtensorMap tm_v, const lf::Params p) {
  using namespace lf;
  extern __shared__ uint8_t smem_raw[];
  const uint32_t raw_addr = smem_u32(smem_raw);
  uint8_t* smem = smem_raw + ((1024u - (raw_addr & 1023u)) & 1023u);
  uint8_t* sQ = smem;
  uint8_t* sK = sQ + kQBytes;
  uint8_t* sV = sK + kStages * kKBytes;
  __half* srel = reinterpret_cast<__half*>(sV + kStages * kVBytes);   // [128][kRelLd]
  float* xch = reinterpret_cast<float*>(sV + kStages * kVBytes + kRelBytes);     // [2][kParts][128]
  uint64_t* bars = reinterpret_cast<uint64_t*>(xch + 2 * kParts * 128);
  uint64_t* q_full = bars;
  uint64_t* k_full = bars + 1;                   // [kStages]
  uint64_t* k_empty = k_full + kStages;
  uint64_t* v_full = k_empty + kStages;
  uint64_t* v_empty = v_full + kStages;
  uint64_t* s_full = v_empty + kStages;          // [2]
  uint64_t* p_full = s_full + 2;                 // [2]
  uint64_t* pv_done = p_full + 2;                // [2]
  uint64_t* o_full = pv_done + 2;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(o_full + 1);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int x0 = blockIdx.x * kTileW, y0 = blockIdx.y * kTileH, dv = blockIdx.z;

  if (warp == 0 && lane == 0) {
    tma_prefetch_desc(&tm_q);
    tma_prefetch_desc(&tm_k);
    tma_prefetch_desc(&tm_v);
    mbar_init(q_full, 1);
    for (int i = 0; i < kStages; ++i) {
      mbar_init(k_full + i, 1); mbar_init(k_empty + i, 1);
      mbar_init(v_full + i, 1); mbar_init(v_empty + i, 1);
    }
    for (int i = 0; i < 2; ++i) {
      mbar_init(s_full + i, 1);
      mbar_init(p_full + i, 32 * kSoftmaxWarps);
      mbar_init(pv_done + i, 1);
    }
    mbar_init(o_full, 1);
    mbar_fence_init();
  }
  if (warp == 1) tmem_alloc<kTmemCols>(tmem_slot);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;

  if (warp == 0) {
    // ------------------------------------------------------------------ TMA producer
    if (lane == 0) {
      mbar_expect_tx(q_full, kQBytes);
      tma_load_4d(sQ, &tm_q, q_full, 0, x0, y0, 0);
      for (int b = 0; b < kBlocks; ++b) {
        const int st = b % kStages;
        const uint32_t ph = ((b / kStages) - 1) & 1;
        if (b >= kStages) mbar_wait(k_empty + st, ph);
        mbar_expect_tx(k_full + st, kKBytes);
        tma_load_4d(sK + st * kKBytes, &tm_k, k_full + st, 0, x0 - kMaxDis, y0 - kMaxDis + kBlockRows * b, 0);
        if (b >= kStages) mbar_wait(v_empty + st, ph);
        mbar_expect_tx(v_full + st, kVBytes);
        tma_load_4d(sV + st * kVBytes, &tm_v, v_full + st, 0, x0 - kMaxDis, y0 - kMaxDis + kBlockRows * b,
                    dv * (kDV / 64));
      }
    }
  } else if (warp == 1) {
    // ------------------------------------------------------------------ MMA issuer
    if (lane == 0) {
      constexpr uint32_t idesc_qk = make_idesc_bf16(128, kBlockKeys, 0, 0);
      constexpr uint32_t idesc_pv = make_idesc_bf16(128, kDV, 0, 1);
      const uint32_t q_addr = smem_u32(sQ);
      auto issue_qk = [&](int b) {
        const int st = b % kStages;
        mbar_wait(k_full + st, (b / kStages) & 1);
        tc_fence_after();
        const uint32_t k_addr = smem_u32(sK + st * kKBytes);
        const uint32_t d = tmem_base + kTmemS + (b & 1) * kBlockKeys;
#pragma unroll
        for (int ks = 0; ks < 8; ++ks) {
          const uint32_t blk = ks >> 2, within = ks & 3;
          mma_bf16_ss(d, make_desc_sw128(q_addr + blk * (128 * 128) + within * 32, 16, 1024),
                      make_desc_sw128(k_addr + blk * (kBlockKeys * 128) + within * 32, 16, 1024), idesc_qk, ks > 0);
        }
        tc_commit(k_empty + st);
        tc_commit(s_full + (b & 1));
      };
      mbar_wait(q_full, 0);
      issue_qk(0);
      for (int b = 0; b < kBlocks; ++b) {
        // S(b+1) goes into the buffer P(b-1) lived in: PV(b-1) was issued before it, and the tcgen05.mma
        // instructions of one thread execute in issue order
        if (b + 1 < kBlocks) issue_qk(b + 1);
        mbar_wait(p_full + (b & 1), (b >> 1) & 1);
        const int st = b % kStages;
        mbar_wait(v_full + st, (b / kStages) & 1);
        tc_fence_after();
        const uint32_t v_addr = smem_u32(sV + st * kVBytes);
        const uint32_t p_tmem = tmem_base + kTmemS + (b & 1) * kBlockKeys;       // bf16 P: 64 columns
#pragma unroll
        for (int ks = 0; ks < kBlockKeys / 16; ++ks)
          mma_bf16_ts(tmem_base + kTmemO, p_tmem + ks * 8,
                      make_desc_sw128(v_addr + ks * (16 * 128), kBlockKeys * 128, 1024), idesc_pv, (b > 0) || (ks > 0));
        tc_commit(v_empty + st);
        tc_commit(pv_done + (b & 1));
      }
      tc_commit(o_full);
    }
  } else {
    // ------------------------------------------------------------------ softmax + epilogue
    // 16 warps, four threads per query: warp w serves TMEM lane quarter w % 4; part = (w - 2) / 4 selects the halo
    // row of the block (32 key slots) the thread owns.  The per-element work (bias
    // fetch, mask, exponential) is instruction- and latency-bound, so it is spread over many warps.
    const int quarter = warp & 3;
    const int part = (warp - 2) >> 2;
    const int hr = part;                          // halo row of the block this thread owns (32 key slots)
    const int row = quarter * 32 + lane;          // query inside the tile = TMEM lane
    const int qy_l = row >> 4, qx_l = row & 15;
    const int gy = y0 + qy_l, gx = x0 + qx_l;
    const bool q_ok = gy < p.h && gx < p.w;
    const uint32_t lane_addr = static_cast<uint32_t>(quarter * 32) << 16;
    const uint32_t row_bar = 1 + quarter;         // named barrier of the four warps sharing these 32 queries
    const float c = p.scale_log2;
    constexpr float kLog2e = 1.4426950408889634f;
    float m_used = -INFINITY, l_sum = 0.f;

    // Stage the tile's relative-position bias rel[q][0..224] (x log2 e, fp16) in shared memory with coalesced
    // row reads: the per-tap gather below is lane = query, i.e. 32 different rows per load instruction -- from
    // global memory that is one L1 sector request per lane and instruction, which bounded the whole kernel.
#pragma unroll 2
    for (int r = warp - 2; r < 128; r += kSoftmaxWarps) {
      const int ry = y0 + (r >> 4), rx = x0 + (r & 15);
      if (ry < p.h && rx < p.w) {                                      // warp-uniform
        const float* src = p.rel + static_cast<size_t>(ry * p.w + rx) * p.ld_rel;
        float t8[8];
#pragma unroll
        for (int k = 0; k < 8; ++k) t8[k] = (lane + 32 * k < kWin * kWin) ? __ldg(src + lane + 32 * k) : 0.f;
#pragma unroll
        for (int k = 0; k < 8; ++k)
          if (lane + 32 * k < kWin * kWin) srel[r * kRelLd + lane + 32 * k] = __float2half_rn(t8[k] * kLog2e);
      }
    }
    named_bar_sync(5, 32 * kSoftmaxWarps);
    const __half* rel_q = srel + row * kRelLd;
    // key slots of a halo row the query may attend to, whatever the row: inside its 15-wide window and the image
    uint32_t colmask = 0;
#pragma unroll
    for (int i = 0; i < 32; ++i) {
      const int dx = i - qx_l, kx = x0 - kMaxDis + i;
      if (q_ok && dx >= 0 && dx < kWin && kx >= 0 && kx < p.w) colmask |= 1u << i;
    }
    const uint32_t o_addr = tmem_base + lane_addr + kTmemO + part * (kDV / kParts);

    for (int b = 0; b < kBlocks; ++b) {
      const int sb = b & 1;
      // this thread's halo row: window row dy of the query, image row ky; per slot: window column slot - qx_l
      const int hy = kBlockRows * b + hr;
      const int dy = hy - qy_l;
      const int ky = y0 - kMaxDis + hy;
      const bool row_in = dy >= 0 && dy < kWin && ky >= 0 && ky < p.h;
      const uint32_t mask = row_in ? colmask : 0u;
      const bool live = __any_sync(0xffffffffu, mask != 0u);   // does any query of this warp see these slots?
      float x[32];
      float mx = -INFINITY;
      if (live) {
        // relative-position bias of the taps (shared memory), fetched while S is still being computed
        float bias[32];
        const __half* rel_row = rel_q + dy * kWin - qx_l;
#pragma unroll
        for (int i = 0; i < 32; ++i) bias[i] = (mask >> i) & 1u ? __half2float(rel_row[i]) : -INFINITY;
        mbar_wait(s_full + sb, (b >> 1) & 1);
        tc_fence_after();
        uint32_t sr[32];
        tmem_ld32(tmem_base + lane_addr + kTmemS + sb * kBlockKeys + hr * 32, sr);
        tmem_ld_wait();
        float mx1 = -INFINITY;
#pragma unroll
        for (int i = 0; i < 32; i += 2) {
          x[i] = fmaf(__uint_as_float(sr[i]), c, bias[i]);      // -inf where masked
          x[i + 1] = fmaf(__uint_as_float(sr[i + 1]), c, bias[i + 1]);
          mx = fmaxf(mx, x[i]);
          mx1 = fmaxf(mx1, x[i + 1]);
        }
        mx = fmaxf(mx, mx1);
      } else {
        mbar_wait(s_full + sb, (b >> 1) & 1);
        tc_fence_after();
      }
      xch[(sb * kParts + part) * 128 + row] = mx;
      named_bar_sync(row_bar, 32 * kParts);
#pragma unroll
      for (int k = 0; k < kParts; ++k) mx = fmaxf(mx, xch[(sb * kParts + k) * 128 + row]);
      const bool grow = mx > m_used + kRescale;               // false while mx == -inf
      float scale = 1.f, m_new = m_used;
      if (grow) {
        m_new = mx;
        scale = fast_exp2(m_used - m_new);                    // 0 when m_used == -inf
      }
      if (b > 0 && __any_sync(0xffffffffu, grow)) {            // same decision in the four warps of the quarter
        mbar_wait(pv_done + ((b - 1) & 1), ((b - 1) >> 1) & 1);
        tc_fence_after();
        uint32_t orow[32];
        tmem_ld32(o_addr, orow);
        tmem_ld_wait();
#pragma unroll
        for (int i = 0; i < 32; ++i) orow[i] = __float_as_uint(__uint_as_float(orow[i]) * scale);
        tmem_st32(o_addr, orow);
      }
      l_sum *= scale;
      m_used = m_new;
      uint32_t packed[16];
      if (live) {
        const float neg_m = (m_used == -INFINITY) ? 0.f : -m_used;     // no valid key yet: exp2(-inf + 0) = 0
        float acc0 = 0.f, acc1 = 0.f;
#pragma unroll
        for (int i = 0; i < 32; i += 2) {
          const float p0 = fast_exp2(x[i] + neg_m), p1 = fast_exp2(x[i + 1] + neg_m);
          acc0 += p0;
          acc1 += p1;
          packed[i >> 1] = pack_bf16x2(p0, p1);
        }
        l_sum += acc0 + acc1;
      } else {
#pragma unroll
        for (int i = 0; i < 16; ++i) packed[i] = 0u;
      }
      // P (bf16, 64 columns) overwrites the first half of this S buffer; every thread of the row holds its S
      // values in registers (it passed tcgen05.wait::ld before the row barrier)
      tmem_st16(tmem_base + lane_addr + kTmemS + sb * kBlockKeys + hr * 16, packed);
      tmem_st_wait();
      tc_fence_before();
      mbar_arrive(p_full + sb);
    }

    // epilogue: out = O / l * gate (bf16); this thread writes 32 of the CTA's 128 value channels of its query
    const size_t off = static_cast<size_t>(q_ok ? gy * p.w + gx : 0) * 1024 + dv * kDV + part * (kDV / kParts);
    uint4 g[4];
    if (p.gate && q_ok) {
#pragma unroll
      for (int i = 0; i < 4; ++i) g[i] = __ldg(reinterpret_cast<const uint4*>(p.gate + off) + i);
    }
    mbar_wait(o_full, 0);
    tc_fence_after();
    // the last block used exchange buffer (kBlocks - 1) & 1; its readers may still be there: use the other one
    constexpr int kLBuf = kBlocks & 1;
    xch[(kLBuf * kParts + part) * 128 + row] = l_sum;
    named_bar_sync(row_bar, 32 * kParts);
    float l_tot = 0.f;
#pragma unroll
    for (int k = 0; k < kParts; ++k) l_tot += xch[(kLBuf * kParts + k) * 128 + row];
    const float inv = l_tot > 0.f ? 1.f / l_tot : 0.f;
    {
      uint32_t orow[32];
      tmem_ld32(o_addr, orow);                    // warp-collective: all lanes, stores guarded below
      tmem_ld_wait();
      if (q_ok) {
#pragma unroll
        for (int i = 0; i < 4; ++i) {
          float v[8];
#pragma unroll
          for (int e = 0; e < 8; ++e) v[e] = __uint_as_float(orow[i * 8 + e]) * inv;
          if (p.gate) {
            const __nv_bfloat162* gh = reinterpret_cast<const __nv_bfloat162*>(&g[i]);
#pragma unroll
            for (int e = 0; e < 4; ++e) {
              v[2 * e] *= __bfloat162float(gh[e].x);
              v[2 * e + 1] *= __bfloat162float(gh[e].y);
            }
          }
          uint4 pk;
          __nv_bfloat162* o = reinterpret_cast<__nv_bfloat162*>(&pk);
#pragma unroll
          for (int e = 0; e < 4; ++e) o[e] = __floats2bfloat162_rn(v[2 * e], v[2 * e + 1]);
          reinterpret_cast<uint4*>(p.out + off)[i] = pk;
        }
      }
    }
    tc_fence_before();
  }

  __syncthreads();
  if (warp == 1) {
    tc_fence_after();
    tmem_dealloc<lf::kTmemCols>(tmem_base);
  }
}

int local_attn_fused_launch(const void* q, const void* k, const void* v, const float* rel, int ld_rel,
                            const void* gate, void* out, int h, int w, cudaStream_t stream) {
  using namespace lf;
  if (h < 1 || w < 1 || ld_rel < kWin * kWin || !rel || !out)
    return set_error(RMEM_ERR_ARG, "local_attn: bad h / w / ld_rel, or a null rel / out");
  auto fn = get_tensor_map_encoder();
  if (!fn) return set_error(RMEM_ERR_DRIVER, "cuTensorMapEncodeTiled entry point not found");
  CUtensorMap tm_q, tm_k, tm_v;
  const uint32_t es[4] = {1, 1, 1, 1};
  {
    const uint64_t dims[4] = {64, static_cast<uint64_t>(w), static_cast<uint64_t>(h), 2};
    const uint64_t strides[3] = {128 * 2, static_cast<uint64_t>(w) * 128 * 2, 128};
    const uint32_t box_q[4] = {64, kTileW, kTileH, 2};
    const uint32_t box_k[4] = {64, kBoxW, kBlockRows, 2};
    CUresult r = fn(&tm_q, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 4, const_cast<void*>(q), dims, strides, box_q, es,
                    CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                    CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_SUCCESS) return set_error(RMEM_ERR_DRIVER, "cuTensorMapEncodeTiled (local Q) failed");
    r = fn(&tm_k, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 4, const_cast<void*>(k), dims, strides, box_k, es,
           CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
           CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_SUCCESS) return set_error(RMEM_ERR_DRIVER, "cuTensorMapEncodeTiled (local K) failed");
  }
  {
    const uint64_t dims[4] = {64, static_cast<uint64_t>(w), static_cast<uint64_t>(h), 16};
    const uint64_t strides[3] = {1024 * 2, static_cast<uint64_t>(w) * 1024 * 2, 128};
    const uint32_t box[4] = {64, kBoxW, kBlockRows, kDV / 64};
    CUresult r = fn(&tm_v, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 4, const_cast<void*>(v), dims, strides, box, es,
                    CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                    CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_SUCCESS) return set_error(RMEM_ERR_DRIVER, "cuTensorMapEncodeTiled (local V) failed");
  }
  static SmemOptIn optin;
  if (int r = ensure_dynamic_smem(reinterpret_cast<const void*>(local_attn_kernel), &optin, kSmem,
                                  "local_attn: cudaFuncSetAttribute"))
    return r;
  Params p;
  p.rel = rel; p.ld_rel = ld_rel;
  p.gate = static_cast<const __nv_bfloat16*>(gate);
  p.out = static_cast<__nv_bfloat16*>(out);
  p.h = h; p.w = w;
  p.scale_log2 = 1.4426950408889634f / sqrtf(128.f);
  dim3 grid((w + kTileW - 1) / kTileW, (h + kTileH - 1) / kTileH, 1024 / kDV);
  local_attn_kernel<<<grid, kThreads, kSmem, stream>>>(tm_q, tm_k, tm_v, p);
  return check_launch("local_attn_kernel");
}

}  // namespace rmem
