// Short-term local attention (DeAOT): 15x15 window over the previous frame, 1 head, d_att 128,
// value width 1024 -- LocalGatedPropagation.forward, reference layers/attention.py:289-361.
//
// Stage 1  local_scores_kernel   : shared-memory halo tile of the previous frame's keys
//          (vectorised, coalesced loads; zero fill = the reference's F.pad), windowed q.k on the
//          CUDA cores, + relative-position bias, border mask, 16-lane shuffle softmax over the
//          225 taps.  Writes probs[q][225] (out-of-image taps = 0, like exp(-1e8) in fp32).
// Stage 2  local_aggregate_kernel: banded P.V on tcgen05.  A CTA owns an 8x16 query tile and 128
//          value channels; the 22x32 halo of V is streamed by TMA (4-D box, out-of-image = zero
//          fill) in 2-row blocks of 64 keys, the matching 128x64 slice of the banded probability
//          matrix is scattered into a 128B-swizzled smem tile, O += P_b V_b accumulates in TMEM;
//          the epilogue applies the gate and writes bf16.  This replaces unfold / local2global /
//          the dense LxL matmul of attention.py:331-335, 350-353, 363-402.
#include "sm100_ptx.cuh"
#include "rmem_internal.h"

namespace rmem {
using namespace sm100;

constexpr int kWin = 15, kTaps = 225, kMaxDis = 7;

// ------------------------------------------------------------------------------------------------
// Stage 1
constexpr int kSTileH = 2, kSTileW = 8;                      // 16 queries per CTA
constexpr int kSHaloH = kSTileH + 2 * kMaxDis;               // 16
constexpr int kSHaloW = kSTileW + 2 * kMaxDis;               // 22
constexpr int kSKeyStride = 136;                             // bf16 per key row (128 + 8 pad)
constexpr uint32_t kSSmemK = kSHaloH * kSHaloW * kSKeyStride * 2;
constexpr uint32_t kSSmemQ = 16 * 128 * 4;
constexpr uint32_t kSSmem = kSSmemK + kSSmemQ;

__global__ void __launch_bounds__(256)
local_scores_kernel(const __nv_bfloat16* __restrict__ q, const __nv_bfloat16* __restrict__ k,
                    const float* __restrict__ rel, int ld_rel, float* __restrict__ probs, int h, int w) {
  extern __shared__ __align__(16) uint8_t smem[];
  __nv_bfloat16* sK = reinterpret_cast<__nv_bfloat16*>(smem);
  float* sQ = reinterpret_cast<float*>(smem + kSSmemK);
  const int x0 = blockIdx.x * kSTileW, y0 = blockIdx.y * kSTileH;
  // halo of K: 16 uint4 per key row
  for (int i = threadIdx.x; i < kSHaloH * kSHaloW * 16; i += 256) {
    const int key = i >> 4, part = i & 15;
    const int hy = key / kSHaloW, hx = key - hy * kSHaloW;
    const int gy = y0 + hy - kMaxDis, gx = x0 + hx - kMaxDis;
    uint4 v = make_uint4(0u, 0u, 0u, 0u);
    if (gy >= 0 && gy < h && gx >= 0 && gx < w)
      v = *reinterpret_cast<const uint4*>(k + static_cast<size_t>(gy * w + gx) * 128 + part * 8);
    *reinterpret_cast<uint4*>(sK + key * kSKeyStride + part * 8) = v;
  }
  for (int i = threadIdx.x; i < 16 * 128; i += 256) {
    const int qi = i >> 7, c = i & 127;
    const int gy = y0 + (qi >> 3), gx = x0 + (qi & 7);
    sQ[i] = (gy < h && gx < w) ? __bfloat162float(q[static_cast<size_t>(gy * w + gx) * 128 + c]) : 0.f;
  }
  __syncthreads();

  const int qi = threadIdx.x >> 4, sub = threadIdx.x & 15;
  const int qy_l = qi >> 3, qx_l = qi & 7;
  const int gy = y0 + qy_l, gx = x0 + qx_l;
  const bool q_ok = gy < h && gx < w;
  const float scale = rsqrtf(128.f);
  const float4* q4 = reinterpret_cast<const float4*>(sQ + qi * 128);
  float s[15];
  float mx = -INFINITY;
#pragma unroll
  for (int i = 0; i < 15; ++i) {
    const int t = sub + 16 * i;
    s[i] = -INFINITY;
    if (t < kTaps) {
      const int dy = t / kWin, dx = t - dy * kWin;
      const int ky = gy + dy - kMaxDis, kx = gx + dx - kMaxDis;
      if (q_ok && ky >= 0 && ky < h && kx >= 0 && kx < w) {
        const uint4* k4 = reinterpret_cast<const uint4*>(
            sK + ((qy_l + dy) * kSHaloW + qx_l + dx) * kSKeyStride);
        float acc = 0.f;
#pragma unroll 4
        for (int c = 0; c < 16; ++c) {
          const uint4 kv = k4[c];
          const float4 qa = q4[2 * c], qb = q4[2 * c + 1];
          acc += qa.x * __uint_as_float(kv.x << 16) + qa.y * __uint_as_float(kv.x & 0xffff0000u);
          acc += qa.z * __uint_as_float(kv.y << 16) + qa.w * __uint_as_float(kv.y & 0xffff0000u);
          acc += qb.x * __uint_as_float(kv.z << 16) + qb.y * __uint_as_float(kv.z & 0xffff0000u);
          acc += qb.z * __uint_as_float(kv.w << 16) + qb.w * __uint_as_float(kv.w & 0xffff0000u);
        }
        s[i] = acc * scale + rel[static_cast<size_t>(gy * w + gx) * ld_rel + t];
      }
    }
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
  float* out = probs + static_cast<size_t>(gy * w + gx) * kTaps;
#pragma unroll
  for (int i = 0; i < 15; ++i) {
    const int t = sub + 16 * i;
    if (t < kTaps) out[t] = s[i] * inv;
  }
}

int local_scores_launch(const void* q, const void* k, const float* rel, int ld_rel, float* probs, int h,
                        int w, cudaStream_t stream) {
  if (h < 1 || w < 1 || ld_rel < kTaps) return set_error(RMEM_ERR_ARG, "local_scores: bad h / w / ld_rel");
  static bool attr_set = false;
  if (!attr_set) {
    cudaError_t e = cudaFuncSetAttribute(local_scores_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, kSSmem);
    if (e != cudaSuccess) return set_cuda_error(e, "local_scores: cudaFuncSetAttribute");
    attr_set = true;
  }
  dim3 grid((w + kSTileW - 1) / kSTileW, (h + kSTileH - 1) / kSTileH);
  local_scores_kernel<<<grid, 256, kSSmem, stream>>>(static_cast<const __nv_bfloat16*>(q),
                                                     static_cast<const __nv_bfloat16*>(k), rel, ld_rel,
                                                     probs, h, w);
  return check_launch("local_scores_kernel");
}

// ------------------------------------------------------------------------------------------------
// Stage 2
constexpr int kATileH = 8, kATileW = 16;                 // 128 queries per CTA
constexpr int kAHaloH = kATileH + 2 * kMaxDis;           // 22 halo rows
constexpr int kABoxW = 32;                               // 30 halo columns + 2 (keeps 64 keys / block)
constexpr int kABlocks = kAHaloH / 2;                    // 11 key blocks of 2 halo rows
constexpr int kADV = 128;                                // value channels per CTA
constexpr int kAStages = 4;
constexpr uint32_t kAVBytes = 64 * kADV * 2;             // 16 KB  [2 cblk][64 keys][64 ch]
constexpr uint32_t kAPBytes = 128 * 64 * 2;              // 16 KB  [128 q][64 keys]
constexpr uint32_t kASmem = kAStages * kAVBytes + 2 * kAPBytes + 1024 + 128;
constexpr int kAThreads = 192;

struct AggParams {
  const float* probs;          // [h*w][225]
  const __nv_bfloat16* gate;   // [h*w][1024] or nullptr
  __nv_bfloat16* out;          // [h*w][1024]
  int h, w;
};

__global__ void __launch_bounds__(kAThreads, 1)
local_aggregate_kernel(const __grid_constant__ CUtensorMap tm_v, const AggParams p) {
  extern __shared__ uint8_t smem_raw[];
  const uint32_t raw_addr = smem_u32(smem_raw);
  uint8_t* smem = smem_raw + ((1024u - (raw_addr & 1023u)) & 1023u);
  uint8_t* sV = smem;
  uint8_t* sP = sV + kAStages * kAVBytes;
  uint64_t* bars = reinterpret_cast<uint64_t*>(sP + 2 * kAPBytes);
  uint64_t* v_full = bars;                 // [4]
  uint64_t* v_empty = bars + 4;            // [4]
  uint64_t* p_full = bars + 8;             // [2]
  uint64_t* pv_done = bars + 10;           // [2]
  uint64_t* o_full = bars + 12;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + 13);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int x0 = blockIdx.x * kATileW, y0 = blockIdx.y * kATileH, dv = blockIdx.z;

  if (warp == 0 && lane == 0) {
    tma_prefetch_desc(&tm_v);
    for (int i = 0; i < kAStages; ++i) { mbar_init(v_full + i, 1); mbar_init(v_empty + i, 1); }
    for (int i = 0; i < 2; ++i) { mbar_init(p_full + i, 128); mbar_init(pv_done + i, 1); }
    mbar_init(o_full, 1);
    mbar_fence_init();
  }
  if (warp == 1) tmem_alloc<128>(tmem_slot);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;

  if (warp == 0) {
    if (lane == 0) {
      for (int b = 0; b < kABlocks; ++b) {
        const int st = b % kAStages;
        if (b >= kAStages) mbar_wait(v_empty + st, ((b / kAStages) - 1) & 1);
        mbar_expect_tx(v_full + st, kAVBytes);
        tma_load_4d(sV + st * kAVBytes, &tm_v, v_full + st, 0, x0 - kMaxDis, y0 - kMaxDis + 2 * b, dv * 2);
      }
    }
  } else if (warp == 1) {
    if (lane == 0) {
      constexpr uint32_t idesc = make_idesc_bf16(128, kADV, 0, 1);
      for (int b = 0; b < kABlocks; ++b) {
        const int st = b % kAStages;
        mbar_wait(v_full + st, (b / kAStages) & 1);
        mbar_wait(p_full + (b & 1), (b >> 1) & 1);
        tc_fence_after();
        const uint32_t v_addr = smem_u32(sV + st * kAVBytes);
        const uint32_t p_addr = smem_u32(sP + (b & 1) * kAPBytes);
#pragma unroll
        for (int ks = 0; ks < 4; ++ks) {
          mma_bf16_ss(tmem_base, make_desc_sw128(p_addr + ks * 32, 16, 1024),
                      make_desc_sw128(v_addr + ks * (16 * 128), 64 * 128, 1024), idesc, (b > 0) || (ks > 0));
        }
        tc_commit(v_empty + st);
        tc_commit(pv_done + (b & 1));
      }
      tc_commit(o_full);
    }
  } else {
    const int quarter = warp & 3;
    const int row = quarter * 32 + lane;
    const int qy_l = row >> 4, qx_l = row & 15;
    const int gy = y0 + qy_l, gx = x0 + qx_l;
    const bool q_ok = gy < p.h && gx < p.w;
    const float* prow = p.probs + static_cast<size_t>(q_ok ? gy * p.w + gx : 0) * kTaps;
    for (int b = 0; b < kABlocks; ++b) {
      if (b >= 2) mbar_wait(pv_done + (b & 1), ((b - 2) >> 1) & 1);
      uint8_t* pbuf = sP + (b & 1) * kAPBytes;
#pragma unroll
      for (int ch = 0; ch < 8; ++ch)
        *reinterpret_cast<uint4*>(pbuf + sw128_offset(row, ch)) = make_uint4(0u, 0u, 0u, 0u);
      if (q_ok) {
#pragma unroll
        for (int rr = 0; rr < 2; ++rr) {
          const int dy = 2 * b + rr - qy_l;
          if (dy >= 0 && dy < kWin) {
            const float* pr = prow + dy * kWin;
#pragma unroll
            for (int dx = 0; dx < kWin; ++dx) {
              const int key = rr * kABoxW + qx_l + dx;
              const __nv_bfloat16 v = __float2bfloat16(pr[dx]);
              *reinterpret_cast<__nv_bfloat16*>(pbuf + sw128_offset(row, key >> 3) + (key & 7) * 2) = v;
            }
          }
        }
      }
      fence_proxy_async_smem();
      mbar_arrive(p_full + (b & 1));
    }
    mbar_wait(o_full, 0);
    tc_fence_after();
    const uint32_t lane_addr = static_cast<uint32_t>(quarter * 32) << 16;
#pragma unroll 1
    for (int cb = 0; cb < kADV / 32; ++cb) {
      uint32_t r[32];
      tmem_ld32(tmem_base + lane_addr + cb * 32, r);
      tmem_ld_wait();
      if (q_ok) {
        const size_t off = static_cast<size_t>(gy * p.w + gx) * 1024 + dv * kADV + cb * 32;
#pragma unroll
        for (int i = 0; i < 32; i += 8) {
          float v[8];
#pragma unroll
          for (int e = 0; e < 8; ++e) v[e] = __uint_as_float(r[i + e]);
          if (p.gate) {
            const uint4 g = *reinterpret_cast<const uint4*>(p.gate + off + i);
            const __nv_bfloat162* gh = reinterpret_cast<const __nv_bfloat162*>(&g);
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
          *reinterpret_cast<uint4*>(p.out + off + i) = pk;
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

int local_aggregate_launch(const float* probs, const void* v, const void* gate, void* out, int h, int w,
                           cudaStream_t stream) {
  if (h < 1 || w < 1) return set_error(RMEM_ERR_ARG, "local_aggregate: bad h / w");
  auto fn = get_tensor_map_encoder();
  if (!fn) return set_error(RMEM_ERR_DRIVER, "cuTensorMapEncodeTiled entry point not found");
  CUtensorMap tm_v;
  const uint64_t dims[4] = {64, static_cast<uint64_t>(w), static_cast<uint64_t>(h), 16};
  const uint64_t strides[3] = {1024 * 2, static_cast<uint64_t>(w) * 1024 * 2, 128};
  const uint32_t box[4] = {64, kABoxW, 2, kADV / 64};
  const uint32_t es[4] = {1, 1, 1, 1};
  CUresult r = fn(&tm_v, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 4, const_cast<void*>(v), dims, strides, box, es,
                  CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B,
                  CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) return set_error(RMEM_ERR_DRIVER, "cuTensorMapEncodeTiled (local V) failed");
  static bool attr_set = false;
  if (!attr_set) {
    cudaError_t e = cudaFuncSetAttribute(local_aggregate_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, kASmem);
    if (e != cudaSuccess) return set_cuda_error(e, "local_aggregate: cudaFuncSetAttribute");
    attr_set = true;
  }
  AggParams p;
  p.probs = probs; p.gate = static_cast<const __nv_bfloat16*>(gate);
  p.out = static_cast<__nv_bfloat16*>(out); p.h = h; p.w = w;
  dim3 grid((w + kATileW - 1) / kATileW, (h + kATileH - 1) / kATileH, 1024 / kADV);
  local_aggregate_kernel<<<grid, kAThreads, kASmem, stream>>>(tm_v, p);
  return check_launch("local_aggregate_kernel");
}

}  // namespace rmem
