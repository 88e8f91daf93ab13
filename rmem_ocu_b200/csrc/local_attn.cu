// Short-term local attention (DeAOT): 15x15 window over the previous frame, 1 head, d_att 128,
// value width 1024 -- LocalGatedPropagation.forward, reference layers/attention.py:289-361.
//
// Stage 1  local_scores_kernel   : shared-memory halo tile of the previous frame's keys
//          (vectorised, coalesced loads; zero fill = the reference's F.pad), windowed q.k on the
//          CUDA cores, + relative-position bias, border mask, 16-lane shuffle softmax over the
//          225 taps.  Writes probs[q][225] (out-of-image taps = 0, like exp(-1e8) in fp32) and / or
//          the same probabilities as bf16 in the banded per-tile layout stage 2 multiplies with.
// Stage 2  local_aggregate_kernel: banded P.V on tcgen05.  A CTA owns an 8x16 query tile and 128
//          value channels; the 22x32 halo of V (4-D box, out-of-image = zero fill) and the matching
//          128x64 slices of the banded probability matrix are both streamed by TMA in 2-row blocks
//          of 64 keys, O += P_b V_b accumulates in TMEM; the epilogue applies the gate and writes
//          bf16.  This replaces unfold / local2global / the dense LxL matmul of attention.py:331-335,
//          350-353, 363-402.
#include <cstdlib>
#include "sm100_ptx.cuh"
#include "rmem_internal.h"

namespace rmem {
using namespace sm100;

constexpr int kWin = 15, kTaps = 225, kMaxDis = 7;

// ------------------------------------------------------------------------------------------------
// Geometry shared by the two stages.  The aggregation runs on 8 x 16 query tiles; the 22 x 30 key halo of a
// tile is cut into 11 blocks of two halo rows (2 x 32 key slots, 30 used), and the probabilities of a tile
// live in a "banded" bf16 layout  P[tile][block][query row 0..127][64 key slots]  in which query (qy, qx)
// has its tap (dy, dx) at block (qy + dy) / 2, slot ((qy + dy) % 2) * 32 + qx + dx.  Slots outside a
// query's window are zero and never written: the caller zero-fills the buffer once per sequence.
constexpr int kATileH = 8, kATileW = 16;                 // 128 queries per aggregation tile
constexpr int kAHaloH = kATileH + 2 * kMaxDis;           // 22 halo rows
constexpr int kABoxW = 32;                               // 30 halo columns + 2 (keeps 64 keys / block)
constexpr int kABlocks = kAHaloH / 2;                    // 11 key blocks of 2 halo rows
constexpr int kABlockElems = 128 * 64;                   // bf16 per (tile, block)

// ------------------------------------------------------------------------------------------------
// Stage 1.  CTA = 2 x 8 queries.  The 16 x 22 key halo arrives as one TMA box (two 64-channel blocks of
// 128-byte rows, 128B swizzle, out-of-image keys zero-filled = the reference's F.pad).  All 16 x 352
// query-key products are one small tensor-core job: each warp runs m16n8k16 MMAs over its share of the
// 8-key column blocks with the query fragments held in registers (the swizzle makes the key-fragment
// loads conflict-free) and parks S in shared memory; then thread (query, sub) picks its 15 of the 225
// taps, adds the relative-position bias and joins the 16-lane shuffle softmax.
constexpr int kSTileH = 2, kSTileW = 8;                      // 16 queries per CTA
constexpr int kSHaloH = kSTileH + 2 * kMaxDis;               // 16
constexpr int kSHaloW = kSTileW + 2 * kMaxDis;               // 22
constexpr int kSKeys = kSHaloH * kSHaloW;                    // 352 = 44 blocks of 8 keys
constexpr int kSLd = 360;                                    // floats per S row (352 + pad)
constexpr uint32_t kSSmemK = 2 * kSKeys * 128;               // 90112 B: [2 channel blocks][352 keys][128 B]
constexpr uint32_t kSSmemS = 16 * kSLd * 4;
constexpr uint32_t kSSmem = kSSmemK + kSSmemS + 1024 + 16;

__device__ __forceinline__ void mma_m16n8k16_bf16(float (&c)[4], const uint32_t (&a)[4], uint32_t b0, uint32_t b1) {
  asm volatile(
      "mma.sync.aligned.m16n8k16.row.col.f32.bf16.bf16.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
      : "+f"(c[0]), "+f"(c[1]), "+f"(c[2]), "+f"(c[3])
      : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(b0), "r"(b1));
}

__global__ void __launch_bounds__(256)
local_scores_kernel(const __grid_constant__ CUtensorMap tm_k, const __nv_bfloat16* __restrict__ q,
                    const float* __restrict__ rel, int ld_rel, float* __restrict__ probs,
                    __nv_bfloat16* __restrict__ banded, int h, int w) {
  extern __shared__ uint8_t smem_raw[];
  const uint32_t raw_addr = smem_u32(smem_raw);
  uint8_t* sK = smem_raw + ((1024u - (raw_addr & 1023u)) & 1023u);
  float* sS = reinterpret_cast<float*>(sK + kSSmemK);
  uint64_t* bar = reinterpret_cast<uint64_t*>(sS + 16 * kSLd);
  const int x0 = blockIdx.x * kSTileW, y0 = blockIdx.y * kSTileH;
  const int warp = __shfl_sync(0xffffffffu, static_cast<int>(threadIdx.x >> 5), 0), lane = threadIdx.x & 31;   // warp-uniform
  {
    // blockIdx.z = sequence of a batch of independent frames stacked as [S][h*w] rows ([S][elems] banded buffers)
    const size_t seq = blockIdx.z, L = static_cast<size_t>(h) * w;
    q += seq * L * 128;
    rel += seq * L * ld_rel;
    if (probs) probs += seq * L * kTaps;
    if (banded)
      banded += seq * (static_cast<size_t>((h + kATileH - 1) / kATileH) * ((w + kATileW - 1) / kATileW) *
                       (kABlocks * kABlockElems));
  }
  if (threadIdx.x == 0) {
    tma_prefetch_desc(&tm_k);
    mbar_init(bar, 1);
    mbar_fence_init();
  }
  pdl_wait();        // programmatic dependent launch: see rmem_internal.h
  pdl_trigger();
  if (threadIdx.x == 0) {
    mbar_expect_tx(bar, kSSmemK);
    tma_load_5d(sK, &tm_k, bar, 0, x0 - kMaxDis, y0 - kMaxDis, 0, static_cast<int>(blockIdx.z));
  }
  // query fragments (A operand, row-major 16 x 16 per k-step) straight from global memory while the halo lands
  uint32_t afrag[8][4];
  {
    const int r0 = lane >> 2, r1 = r0 + 8;
    const int c = 2 * (lane & 3);
    const int gy0 = y0 + (r0 >> 3), gx0 = x0 + (r0 & 7), gy1 = y0 + (r1 >> 3), gx1 = x0 + (r1 & 7);
    const bool ok0 = gy0 < h && gx0 < w, ok1 = gy1 < h && gx1 < w;
    const __nv_bfloat16* q0 = q + static_cast<size_t>(ok0 ? gy0 * w + gx0 : 0) * 128 + c;
    const __nv_bfloat16* q1 = q + static_cast<size_t>(ok1 ? gy1 * w + gx1 : 0) * 128 + c;
#pragma unroll
    for (int ks = 0; ks < 8; ++ks) {
      afrag[ks][0] = ok0 ? *reinterpret_cast<const uint32_t*>(q0 + ks * 16) : 0u;
      afrag[ks][1] = ok1 ? *reinterpret_cast<const uint32_t*>(q1 + ks * 16) : 0u;
      afrag[ks][2] = ok0 ? *reinterpret_cast<const uint32_t*>(q0 + ks * 16 + 8) : 0u;
      afrag[ks][3] = ok1 ? *reinterpret_cast<const uint32_t*>(q1 + ks * 16 + 8) : 0u;
    }
  }
  __syncthreads();                 // barrier init visible to every thread
  mbar_wait(bar, 0);
  // S = Q K^T: warp handles key blocks warp, warp + 8, ...
  {
    const int kr = lane >> 2;                        // key inside the block (B column)
    const int cw = 4 * (lane & 3);                   // byte offset of this lane's channel pair inside a 16-byte chunk
#pragma unroll 1
    for (int nb = warp; nb < kSKeys / 8; nb += 8) {
      const int key = nb * 8 + kr;
      const uint8_t* krow = sK + key * 128;
      const uint32_t sw = static_cast<uint32_t>(key & 7);
      float c[4] = {0.f, 0.f, 0.f, 0.f};
#pragma unroll
      for (int ks = 0; ks < 8; ++ks) {
        // channels ks*16 + 2(lane%4) (+8): channel block ks / 4, 16-byte chunks (ks % 4) * 2 and + 1
        const uint8_t* blk = krow + (ks >> 2) * (kSKeys * 128);
        const uint32_t ch0 = static_cast<uint32_t>((ks & 3) * 2);
        const uint32_t b0 = *reinterpret_cast<const uint32_t*>(blk + ((ch0 ^ sw) << 4) + cw);
        const uint32_t b1 = *reinterpret_cast<const uint32_t*>(blk + (((ch0 + 1) ^ sw) << 4) + cw);
        mma_m16n8k16_bf16(c, afrag[ks], b0, b1);
      }
      float* d = sS + (lane >> 2) * kSLd + nb * 8 + 2 * (lane & 3);
      *reinterpret_cast<float2*>(d) = make_float2(c[0], c[1]);
      *reinterpret_cast<float2*>(d + 8 * kSLd) = make_float2(c[2], c[3]);
    }
  }
  __syncthreads();

  // thread = (query, sub): taps sub, sub + 16, ...
  const int qi = threadIdx.x >> 4, sub = threadIdx.x & 15;
  const int qy_l = qi >> 3, qx_l = qi & 7;
  const int gy = y0 + qy_l, gx = x0 + qx_l;
  const bool q_ok = gy < h && gx < w;
  const float scale = rsqrtf(128.f);
  bool ok[15];
  float s[15];
#pragma unroll
  for (int i = 0; i < 15; ++i) {
    const int t = sub + 16 * i;
    const int dy = t / kWin, dx = t - dy * kWin;
    const int ky = gy + dy - kMaxDis, kx = gx + dx - kMaxDis;
    ok[i] = q_ok && t < kTaps && ky >= 0 && ky < h && kx >= 0 && kx < w;
    s[i] = ok[i] ? sS[qi * kSLd + (qy_l + dy) * kSHaloW + qx_l + dx] : 0.f;
  }
  float mx = -INFINITY;
  const float* relq = rel + static_cast<size_t>(q_ok ? gy * w + gx : 0) * ld_rel;
#pragma unroll
  for (int i = 0; i < 15; ++i) {
    s[i] = ok[i] ? s[i] * scale + relq[sub + 16 * i] : -INFINITY;
    mx = fmaxf(mx, s[i]);
  }
#pragma unroll
  for (int o = 8; o > 0; o >>= 1) mx = fmaxf(mx, __shfl_xor_sync(0xffffffffu, mx, o));
  float sum = 0.f;
#pragma unroll
  for (int i = 0; i < 15; ++i) {
    s[i] = (s[i] == -INFINITY) ? 0.f : __expf(s[i] - mx);
    sum += s[i];
  }
#pragma unroll
  for (int o = 8; o > 0; o >>= 1) sum += __shfl_xor_sync(0xffffffffu, sum, o);
  if (!q_ok) return;
  const float inv = 1.f / sum;
  if (probs) {
    float* out = probs + static_cast<size_t>(gy * w + gx) * kTaps;
#pragma unroll
    for (int i = 0; i < 15; ++i) {
      const int t = sub + 16 * i;
      if (t < kTaps) out[t] = s[i] * inv;
    }
  }
  if (banded) {
    const int tiles_x = (w + kATileW - 1) / kATileW;
    const int ty = gy / kATileH, tx = gx / kATileW;
    const int ay = gy - ty * kATileH, ax = gx - tx * kATileW;
    __nv_bfloat16* tb = banded + static_cast<size_t>(ty * tiles_x + tx) * (kABlocks * kABlockElems) +
                        (ay * kATileW + ax) * 64;
#pragma unroll
    for (int i = 0; i < 15; ++i) {
      const int t = sub + 16 * i;
      if (t < kTaps) {
        const int dy = t / kWin, dx = t - dy * kWin;
        const int hy = ay + dy;
        tb[(hy >> 1) * kABlockElems + (hy & 1) * kABoxW + ax + dx] = __float2bfloat16(s[i] * inv);
      }
    }
  }
}

int local_scores_launch(const void* q, const void* k, const float* rel, int ld_rel, float* probs, void* banded,
                        int h, int w, cudaStream_t stream) {
  return local_scores_batched_launch(q, k, rel, ld_rel, probs, banded, h, w, 1, stream);
}

int local_scores_batched_launch(const void* q, const void* k, const float* rel, int ld_rel, float* probs, void* banded,
                                int h, int w, int S, cudaStream_t stream) {
  if (h < 1 || w < 1 || ld_rel < kTaps || (!probs && !banded) || S < 1 || S > 65535)
    return set_error(RMEM_ERR_ARG, "local_scores: bad h / w / ld_rel / batch, or no output");
  auto fn = get_tensor_map_encoder();
  if (!fn) return set_error(RMEM_ERR_DRIVER, "cuTensorMapEncodeTiled entry point not found");
  CUtensorMap tm_k;
  {
    const uint64_t dims[5] = {64, static_cast<uint64_t>(w), static_cast<uint64_t>(h), 2, static_cast<uint64_t>(S)};
    const uint64_t strides[4] = {128 * 2, static_cast<uint64_t>(w) * 128 * 2, 128,
                                 static_cast<uint64_t>(h) * w * 128 * 2};
    const uint32_t box[5] = {64, kSHaloW, kSHaloH, 2, 1};
    const uint32_t es[5] = {1, 1, 1, 1, 1};
    CUresult r = fn(&tm_k, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 5, const_cast<void*>(k), dims, strides, box, es,
                    CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B,
                    CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_SUCCESS) return set_error(RMEM_ERR_DRIVER, "cuTensorMapEncodeTiled (local K) failed");
  }
  static SmemOptIn optin;
  if (int r = ensure_dynamic_smem(reinterpret_cast<const void*>(local_scores_kernel), &optin, kSSmem,
                                  "local_scores: cudaFuncSetAttribute"))
    return r;
  dim3 grid((w + kSTileW - 1) / kSTileW, (h + kSTileH - 1) / kSTileH, S);
  launch_k(local_scores_kernel, grid, 256, kSSmem, stream, tm_k, static_cast<const __nv_bfloat16*>(q), rel, ld_rel,
                                                     probs, static_cast<__nv_bfloat16*>(banded), h, w);
  return check_launch("local_scores_kernel");
}

// ------------------------------------------------------------------------------------------------
// Stage 2: O[tile, dv slice] = sum_b P_b V_b with both operands streamed by TMA (P blocks from the banded
// buffer, V halo blocks as 4-D boxes with out-of-image zero fill), accumulation in TMEM, gate in the epilogue.
// Variants (value channels per CTA, pipeline stages -> CTAs per SM), timed at 30 x 54 tokens with tools/dev_local.py
// (aggregate kernel alone, us, S = 1 / 2 / 4 / 8 stacked sequences):
//   1: 128 ch, 4 stages, 1 CTA/SM   12.2 / 20.0 / 35.8 / 61.0     <- one sequence
//   2: 256 ch, 4 stages, 1 CTA/SM   16.0 / 16.8 / 29.6 / 54.7
//   3: 256 ch, 2 stages, 2 CTA/SM   18.7 / 19.0 / 27.4 / 48.2
//   4: 128 ch, 3 stages, 2 CTA/SM   12.6 / 16.9 / 27.7 / 49.3     <- batches (a second CTA hides the epilogue)
// At S = 8 the kernel moves 270-360 MB from L2 to shared memory (P blocks re-read per value slice, V halo rows re-read
// by neighbouring tiles: 5.5x) in ~50 us = 5.5-6 TB/s, about half the measured L2 ceiling; what would cut the traffic
// itself -- one CTA sweeping a column strip so V rows are fetched once, P multicast across value slices -- is
// described in DESIGN.md.  RMEM_LOCAL_AGG_VARIANT=1..4 forces a variant.
constexpr uint32_t kAPBytes = 128 * 64 * 2;              // 16 KB  [128 q][64 keys]
template <int kADV, int kAStages> struct AggCfg {
  static constexpr uint32_t kVBytes = 64 * kADV * 2;     // 16 / 32 KB  [kADV / 64 cblk][64 keys][64 ch]
  static constexpr uint32_t kStageBytes = kVBytes + kAPBytes;
  static constexpr uint32_t kSmem = kAStages * kStageBytes + 1024 + 128;
  static constexpr int kMinCtas = kSmem <= 110 * 1024 ? 2 : 1;
};
constexpr int kAThreads = 64 + 128;

struct AggParams {
  const __nv_bfloat16* gate;   // [h*w][1024] or nullptr
  __nv_bfloat16* out;          // [h*w][1024]
  int h, w;
};

template <int kADV, int kAStages>
__global__ void __launch_bounds__(kAThreads, AggCfg<kADV, kAStages>::kMinCtas)
local_aggregate_kernel(const __grid_constant__ CUtensorMap tm_p, const __grid_constant__ CUtensorMap tm_v,
                       const AggParams p) {
  extern __shared__ uint8_t smem_raw[];
  const uint32_t raw_addr = smem_u32(smem_raw);
  uint8_t* smem = smem_raw + ((1024u - (raw_addr & 1023u)) & 1023u);
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem + kAStages * AggCfg<kADV, kAStages>::kStageBytes);
  uint64_t* full = bars;                   // [kAStages]
  uint64_t* empty = bars + kAStages;       // [kAStages]
  uint64_t* o_full = bars + 2 * kAStages;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + 2 * kAStages + 1);

  const int warp = __shfl_sync(0xffffffffu, static_cast<int>(threadIdx.x >> 5), 0), lane = threadIdx.x & 31;   // warp-uniform
  constexpr int kDvSlices = 1024 / kADV;
  // blockIdx.z = (sequence, value slice): a batch of independent frames stacked as [S][h*w] rows
  const int x0 = blockIdx.x * kATileW, y0 = blockIdx.y * kATileH, dv = blockIdx.z % kDvSlices;
  const int seq = blockIdx.z / kDvSlices;
  const int tile = (seq * gridDim.y + blockIdx.y) * gridDim.x + blockIdx.x;

  if (warp == 0 && lane == 0) {
    tma_prefetch_desc(&tm_p);
    tma_prefetch_desc(&tm_v);
    for (int i = 0; i < kAStages; ++i) { mbar_init(full + i, 1); mbar_init(empty + i, 1); }
    mbar_init(o_full, 1);
    mbar_fence_init();
  }
  if (warp == 1) tmem_alloc<kADV>(tmem_slot);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = __shfl_sync(0xffffffffu, *tmem_slot, 0);
  pdl_wait();        // programmatic dependent launch: see rmem_internal.h
  pdl_trigger();

  if (warp == 0) {
    if (lane == 0) {
      for (int b = 0; b < kABlocks; ++b) {
        const int st = b % kAStages;
        if (b >= kAStages) mbar_wait(empty + st, ((b / kAStages) - 1) & 1);
        uint8_t* dst = smem + st * AggCfg<kADV, kAStages>::kStageBytes;
        mbar_expect_tx(full + st, AggCfg<kADV, kAStages>::kStageBytes);
        tma_load_2d(dst, &tm_p, full + st, 0, (tile * kABlocks + b) * 128);
        tma_load_5d(dst + kAPBytes, &tm_v, full + st, 0, x0 - kMaxDis, y0 - kMaxDis + 2 * b, dv * (kADV / 64), seq);
      }
    }
  } else if (warp == 1) {
    // whole warp in the loop, one elected lane issues (warp-uniform operands stay in uniform registers: one
    // instruction per tcgen05.mma instead of an ELECT / R2UR.BROADCAST loop)
    constexpr uint32_t idesc = make_idesc_bf16(128, kADV, 0, 1);
    for (int b = 0; b < kABlocks; ++b) {
      const int st = b % kAStages;
      mbar_wait(full + st, (b / kAStages) & 1);
      tc_fence_after();
      const uint32_t p_addr = smem_u32(smem + st * AggCfg<kADV, kAStages>::kStageBytes);
      const uint32_t v_addr = p_addr + kAPBytes;
      if (elect_one()) {
#pragma unroll
        for (int ks = 0; ks < 4; ++ks) {
          mma_bf16_ss(tmem_base, make_desc_sw128(p_addr + ks * 32, 16, 1024),
                      make_desc_sw128(v_addr + ks * (16 * 128), 64 * 128, 1024), idesc, (b > 0) || (ks > 0));
        }
        tc_commit(empty + st);
      }
      __syncwarp();
    }
    if (elect_one()) tc_commit(o_full);
    __syncwarp();
  } else {
    const int quarter = warp & 3;
    const int row = quarter * 32 + lane;
    const int qy_l = row >> 4, qx_l = row & 15;
    const int gy = y0 + qy_l, gx = x0 + qx_l;
    const bool q_ok = gy < p.h && gx < p.w;
    mbar_wait(o_full, 0);
    tc_fence_after();
    const uint32_t lane_addr = static_cast<uint32_t>(quarter * 32) << 16;
#pragma unroll 1
    for (int cb = 0; cb < kADV / 32; ++cb) {
      uint32_t r[32];
      tmem_ld32(tmem_base + lane_addr + cb * 32, r);      // warp-collective: all lanes, stores guarded below
      tmem_ld_wait();
      if (q_ok) {
        const size_t off = (static_cast<size_t>(seq) * p.h * p.w + gy * p.w + gx) * 1024 + dv * kADV + cb * 32;
        uint4 g[4];
        if (p.gate) {
#pragma unroll
          for (int i = 0; i < 4; ++i) g[i] = *reinterpret_cast<const uint4*>(p.gate + off + i * 8);
        }
#pragma unroll
        for (int i = 0; i < 4; ++i) {
          float v[8];
#pragma unroll
          for (int e = 0; e < 8; ++e) v[e] = __uint_as_float(r[i * 8 + e]);
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
          *reinterpret_cast<uint4*>(p.out + off + i * 8) = pk;
        }
      }
    }
    tc_fence_before();
  }
  __syncthreads();
  if (warp == 1) {
    tc_fence_after();
    tmem_dealloc<kADV>(tmem_base);
  }
}

size_t local_banded_elems(int h, int w) {
  const size_t tiles = static_cast<size_t>((h + kATileH - 1) / kATileH) * ((w + kATileW - 1) / kATileW);
  return tiles * kABlocks * kABlockElems;
}

int local_aggregate_launch(const void* banded, const void* v, const void* gate, void* out, int h, int w,
                           cudaStream_t stream) {
  return local_aggregate_batched_launch(banded, v, gate, out, h, w, 1, stream);
}

int local_aggregate_batched_launch(const void* banded, const void* v, const void* gate, void* out, int h, int w, int S,
                                   cudaStream_t stream) {
  if (h < 1 || w < 1 || !banded || S < 1 || S * 8 > 65535)
    return set_error(RMEM_ERR_ARG, "local_aggregate: bad h / w / batch / banded buffer");
  auto fn = get_tensor_map_encoder();
  if (!fn) return set_error(RMEM_ERR_DRIVER, "cuTensorMapEncodeTiled entry point not found");
  const int tiles_x = (w + kATileW - 1) / kATileW, tiles_y = (h + kATileH - 1) / kATileH;
  static const int forced = [] { const char* e = getenv("RMEM_LOCAL_AGG_VARIANT"); return e ? atoi(e) : 0; }();
  const int variant = forced ? forced : (S >= 2 ? 4 : 1);
  const int dv_per_cta = (variant == 2 || variant == 3) ? 256 : 128;
  CUtensorMap tm_v, tm_p;
  {
    const uint64_t dims[5] = {64, static_cast<uint64_t>(w), static_cast<uint64_t>(h), 16, static_cast<uint64_t>(S)};
    const uint64_t strides[4] = {1024 * 2, static_cast<uint64_t>(w) * 1024 * 2, 128,
                                 static_cast<uint64_t>(h) * w * 1024 * 2};
    const uint32_t box[5] = {64, kABoxW, 2, static_cast<uint32_t>(dv_per_cta / 64), 1};
    const uint32_t es[5] = {1, 1, 1, 1, 1};
    CUresult r = fn(&tm_v, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 5, const_cast<void*>(v), dims, strides, box, es,
                    CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B,
                    CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_SUCCESS) return set_error(RMEM_ERR_DRIVER, "cuTensorMapEncodeTiled (local V) failed");
  }
  {
    const uint64_t dims[2] = {64, static_cast<uint64_t>(S) * tiles_x * tiles_y * kABlocks * 128};
    const uint64_t strides[1] = {128};
    const uint32_t box[2] = {64, 128};
    const uint32_t es[2] = {1, 1};
    CUresult r = fn(&tm_p, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, const_cast<void*>(banded), dims, strides, box, es,
                    CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B,
                    CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_SUCCESS) return set_error(RMEM_ERR_DRIVER, "cuTensorMapEncodeTiled (banded P) failed");
  }
  AggParams p;
  p.gate = static_cast<const __nv_bfloat16*>(gate);
  p.out = static_cast<__nv_bfloat16*>(out); p.h = h; p.w = w;
  dim3 grid(tiles_x, tiles_y, (1024 / dv_per_cta) * S);
#define RMEM_AGG_LAUNCH(DV, ST)                                                                                    \
  {                                                                                                                \
    static SmemOptIn optin;                                                                                        \
    if (int r = ensure_dynamic_smem(reinterpret_cast<const void*>(local_aggregate_kernel<DV, ST>), &optin,         \
                                    AggCfg<DV, ST>::kSmem, "local_aggregate: cudaFuncSetAttribute"))               \
      return r;                                                                                                    \
    launch_k(local_aggregate_kernel<DV, ST>, grid, kAThreads, AggCfg<DV, ST>::kSmem, stream, tm_p, tm_v, p);       \
  }
  if (variant == 1) RMEM_AGG_LAUNCH(128, 4)
  else if (variant == 2) RMEM_AGG_LAUNCH(256, 4)
  else if (variant == 3) RMEM_AGG_LAUNCH(256, 2)
  else RMEM_AGG_LAUNCH(128, 3)
#undef RMEM_AGG_LAUNCH
  return check_launch("local_aggregate_kernel");
}

}  // namespace rmem
