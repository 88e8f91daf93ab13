// Multi-head attention of the AOT propagation block (8 heads x 32) on tcgen05 / TMEM / TMA --
// MultiheadAttention.forward, reference aot_plus/networks/layers/attention.py:28-81, in its three
// roles inside SimplifiedTransformerBlock.forward (layers/transformer.py:553-692):
//   long-term attention over the restricted bank (:632-635, explicit softmax path attention.py:45-64)
//     with the temporal positional embedding on the keys (:594-629) and the attention-weight recording
//     (:636-643);  short-term attention over the previous frame (:657-662);  self-attention (:565-571).
//
// Decomposition (split-KV, like csrc/lt_attn.cu):
//   mha_partial_kernel : grid (q_tile, head_pair, chunk).  One CTA owns 128 queries x 2 heads x one key
//       chunk.  A head pair is one 64-channel (128-byte) swizzle block, so the TMA boxes and the UMMA
//       descriptors are the ones lt_attn.cu uses; head 1 of the pair starts 64 bytes into the row.
//       Per 64-key tile and head: S = Q_h K_h^T (K = 32) into a double-buffered 64-column TMEM tile,
//       P (bf16) overwrites it in place, O_h += P V with the whole 64-channel V block as B (only the
//       head's own 32 columns are read back; tensor time is irrelevant here -- at head_dim 32 the
//       exponentials outnumber the MMA cycles 4:1, so this kernel is MUFU-bound by construction and a
//       third of the exponentials run as FMA-pipe polynomials).  Eight softmax warps: four per head,
//       one query row per thread, no cross-thread exchange.
//   mha_combine_kernel : merges the chunk partials per head, adds the per-(head, frame) TPE bias,
//       emits the head-averaged per-frame attention mass W[q,t] and writes bf16 [L][256].
#include "sm100_ptx.cuh"
#include "rmem_internal.h"

namespace rmem {
using namespace sm100;

namespace mha {
constexpr int kBlockM = 128;
constexpr int kBlockN = 64;
constexpr int kHeads = 8, kHeadDim = 32, kC = 256;
constexpr int kStages = 7;                            // 128 KB of shared memory: one CTA per SM (it owns all 512 TMEM columns)
constexpr int kThreads = 320;                       // warp0 TMA, warp1 MMA, warps 2-5 head 0, warps 6-9 head 1
constexpr uint32_t kQBytes = kBlockM * 64 * 2;      // 16 KB  [128][64]
constexpr uint32_t kKBytes = kBlockN * 64 * 2;      //  8 KB  [64][64]
constexpr uint32_t kVBytes = kBlockN * 64 * 2;      //  8 KB
constexpr uint32_t kStageBytes = kKBytes + kVBytes;
constexpr uint32_t kSmemData = kQBytes + kStages * kStageBytes;
constexpr uint32_t kSmemTotal = kSmemData + 256;
constexpr uint32_t kTmemCols = 512;                 // S: [2 heads][2 buffers] x 64 cols = 256; O: 2 x 64 at 256
constexpr uint32_t kTmemO = 256;
constexpr float kRescaleThreshold = 8.0f;
constexpr int kPolyEvery = 3;

struct Params {
  int L, n_chunks, chunks_per_frame;
  const int32_t* order;   // [T] logical position -> slot
  float* part_o;          // [n_chunks][L][256]
  float* part_m;          // [n_chunks][8][L]
  float* part_l;          // [n_chunks][8][L]
  float scale_log2;       // log2(e) / sqrt(32)
};
}  // namespace mha

__global__ void __launch_bounds__(mha::kThreads, 1)
mha_partial_kernel(const __grid_constant__ CUtensorMap tm_q, const __grid_constant__ CUtensorMap tm_k,
                   const __grid_constant__ CUtensorMap tm_v, const mha::Params p) {
  using namespace mha;
  extern __shared__ __align__(1024) uint8_t smem[];
  uint8_t* sQ = smem;
  uint8_t* sKV = sQ + kQBytes;
  uint64_t* bars = reinterpret_cast<uint64_t*>(sKV + kStages * kStageBytes);
  uint64_t* q_full = bars + 0;
  uint64_t* kv_full = bars + 1;                    // [kStages]
  uint64_t* kv_empty = kv_full + kStages;          // [kStages]
  uint64_t* s_full = kv_empty + kStages;           // [head][buf]
  uint64_t* p_full = s_full + 4;                   // [head][buf]
  uint64_t* pv_done = p_full + 4;                  // [head][buf]
  uint64_t* o_full = pv_done + 4;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(o_full + 1);

  const int warp = __shfl_sync(0xffffffffu, static_cast<int>(threadIdx.x >> 5), 0), lane = threadIdx.x & 31;   // warp-uniform
  const int q_tile = blockIdx.x, pair = blockIdx.y, chunk = blockIdx.z;

  const int frame_pos = chunk / p.chunks_per_frame;
  const int sub = chunk - frame_pos * p.chunks_per_frame;
  const int per = (p.L + p.chunks_per_frame - 1) / p.chunks_per_frame;
  const int per_aligned = ((per + kBlockN - 1) / kBlockN) * kBlockN;
  const int key_begin = sub * per_aligned;
  int key_len = p.L - key_begin;
  if (key_len > per_aligned) key_len = per_aligned;
  if (key_len < 0) key_len = 0;
  const int n_tiles = (key_len + kBlockN - 1) / kBlockN;

  if (warp == 0 && lane == 0) {
    if (smem_u32(smem) & 1023u) asm volatile("trap;");
    tma_prefetch_desc(&tm_q);
    tma_prefetch_desc(&tm_k);
    tma_prefetch_desc(&tm_v);
    mbar_init(q_full, 1);
    for (int i = 0; i < kStages; ++i) { mbar_init(kv_full + i, 1); mbar_init(kv_empty + i, 1); }
    for (int i = 0; i < 4; ++i) { mbar_init(s_full + i, 1); mbar_init(p_full + i, 128); mbar_init(pv_done + i, 1); }
    mbar_init(o_full, 1);
    mbar_fence_init();
  }
  if (warp == 1) tmem_alloc<kTmemCols>(tmem_slot);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = __shfl_sync(0xffffffffu, *tmem_slot, 0);

  if (warp == 0) {
    // ------------------------------------------------------------------ TMA producer
    if (lane == 0 && n_tiles > 0) {
      const int slot = p.order[frame_pos];
      mbar_expect_tx(q_full, kQBytes);
      tma_load_3d(sQ, &tm_q, q_full, 0, q_tile * kBlockM, pair);
      for (int j = 0; j < n_tiles; ++j) {
        const int st = j % kStages;
        if (j >= kStages) mbar_wait(kv_empty + st, ((j / kStages) - 1) & 1);
        uint8_t* dK = sKV + st * kStageBytes;
        mbar_expect_tx(kv_full + st, kStageBytes);
        tma_load_4d(dK, &tm_k, kv_full + st, 0, key_begin + j * kBlockN, pair, slot);
        tma_load_4d(dK + kKBytes, &tm_v, kv_full + st, 0, key_begin + j * kBlockN, pair, slot);
      }
    }
  } else if (warp == 1) {
    // ------------------------------------------------------------------ MMA issuer
    // whole warp in the loop, one elected lane issues: warp-uniform operands stay in uniform registers (one
    // instruction per tcgen05.mma instead of an ELECT / R2UR.BROADCAST loop)
    if (n_tiles > 0) {
      constexpr uint32_t idesc_qk = make_idesc_bf16(kBlockM, kBlockN, 0, 0);
      constexpr uint32_t idesc_pv = make_idesc_bf16(kBlockM, 64, 0, 1);
      const uint32_t q_addr = smem_u32(sQ);
      auto issue_qk = [&](int j) {
        const int st = j % kStages;
        mbar_wait(kv_full + st, (j / kStages) & 1);
        tc_fence_after();
        const uint32_t k_addr = smem_u32(sKV + st * kStageBytes);
#pragma unroll
        for (int h = 0; h < 2; ++h) {
          const uint32_t d = tmem_base + (h * 2 + (j & 1)) * kBlockN;
          if (elect_one()) {
#pragma unroll
            for (int ks = 0; ks < kHeadDim / 16; ++ks) {
              const uint32_t off = h * 64 + ks * 32;     // head h = bytes [64h, 64h+64) of the 128-byte row
              mma_bf16_ss(d, make_desc_sw128(q_addr + off, 16, 1024), make_desc_sw128(k_addr + off, 16, 1024),
                          idesc_qk, ks > 0);
            }
            tc_commit(s_full + h * 2 + (j & 1));
          }
          __syncwarp();
        }
      };
      mbar_wait(q_full, 0);
      issue_qk(0);
      for (int j = 0; j < n_tiles; ++j) {
        if (j + 1 < n_tiles) issue_qk(j + 1);     // its S buffers held P(j-1); PV(j-1) was issued first
        const int st = j % kStages;
        const uint32_t v_addr = smem_u32(sKV + st * kStageBytes + kKBytes);
#pragma unroll
        for (int h = 0; h < 2; ++h) {
          mbar_wait(p_full + h * 2 + (j & 1), (j >> 1) & 1);
          tc_fence_after();
          const uint32_t p_tmem = tmem_base + (h * 2 + (j & 1)) * kBlockN;   // bf16 P: 32 columns
          const uint32_t d = tmem_base + kTmemO + h * 64;
          if (elect_one()) {
#pragma unroll
            for (int ks = 0; ks < kBlockN / 16; ++ks)
              mma_bf16_ts(d, p_tmem + ks * 8, make_desc_sw128(v_addr + ks * (16 * 128), kBlockN * 128, 1024),
                          idesc_pv, (j > 0) || (ks > 0));
            tc_commit(pv_done + h * 2 + (j & 1));
          }
          __syncwarp();
        }
        if (elect_one()) tc_commit(kv_empty + st);
        __syncwarp();
      }
      if (elect_one()) tc_commit(o_full);
      __syncwarp();
    }
  } else {
    // ------------------------------------------------------------------ softmax + epilogue
    const int quarter = warp & 3;
    const int h = (warp - 2) >> 2;                // head inside the pair
    const int row = quarter * 32 + lane;
    const uint32_t lane_addr = static_cast<uint32_t>(quarter * 32) << 16;
    const int q_global = q_tile * kBlockM + row;
    const int head = pair * 2 + h;
    float m_used = -INFINITY, l_sum = 0.f;
    const float c = p.scale_log2;
    const uint32_t o_addr = tmem_base + lane_addr + kTmemO + h * 64 + h * 32;   // this head's 32 columns

    for (int j = 0; j < n_tiles; ++j) {
      const int b = j & 1;
      mbar_wait(s_full + h * 2 + b, (j >> 1) & 1);
      tc_fence_after();
      uint32_t sr[64];
      {
        uint32_t(&lo)[32] = *reinterpret_cast<uint32_t(*)[32]>(&sr[0]);
        uint32_t(&hi)[32] = *reinterpret_cast<uint32_t(*)[32]>(&sr[32]);
        const uint32_t s_addr = tmem_base + lane_addr + (h * 2 + b) * kBlockN;
        tmem_ld32(s_addr, lo);
        tmem_ld32(s_addr + 32, hi);
        tmem_ld_wait();
      }
      const int valid = key_len - j * kBlockN;
      if (valid < 64) {
#pragma unroll
        for (int i = 0; i < 64; ++i)
          if (i >= valid) sr[i] = 0xff800000u;
      }
      float mx = -INFINITY;
#pragma unroll
      for (int i = 0; i < 64; i += 2)
        mx = fmaxf(mx, fmaxf(__uint_as_float(sr[i]), __uint_as_float(sr[i + 1])));
      mx *= c;
      const bool grow = mx > m_used + kRescaleThreshold;
      float scale = 1.f, m_new = m_used;
      if (grow) {
        m_new = mx;
        scale = fast_exp2(m_used - m_new);
      }
      if (j > 0 && __any_sync(0xffffffffu, grow)) {
        mbar_wait(pv_done + h * 2 + ((j - 1) & 1), ((j - 1) >> 1) & 1);
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
      const float neg_m = -m_used;
      float acc0 = 0.f, acc1 = 0.f;
      uint32_t packed[32];
#pragma unroll
      for (int i = 0; i < 64; i += 2) {
        float x0, x1, p0, p1;
        fma_f32x2(x0, x1, __uint_as_float(sr[i]), __uint_as_float(sr[i + 1]), c, c, neg_m, neg_m);
        if ((i >> 1) % kPolyEvery == kPolyEvery - 1) {
          exp2_poly_x2(p0, p1, x0, x1);
        } else {
          p0 = fast_exp2(x0);
          p1 = fast_exp2(x1);
        }
        add_f32x2(acc0, acc1, acc0, acc1, p0, p1);
        packed[i >> 1] = pack_bf16x2(p0, p1);
      }
      l_sum += acc0 + acc1;
      tmem_st32(tmem_base + lane_addr + (h * 2 + b) * kBlockN, packed);
      tmem_st_wait();
      tc_fence_before();
      mbar_arrive(p_full + h * 2 + b);
    }

    if (n_tiles > 0) {
      mbar_wait(o_full, 0);
      tc_fence_after();
    }
    // tcgen05.ld is warp-collective (.sync.aligned): every lane takes part, only the stores are guarded
    uint32_t orow[32];
    if (n_tiles > 0) {
      tmem_ld32(o_addr, orow);
      tmem_ld_wait();
    } else {
#pragma unroll
      for (int i = 0; i < 32; ++i) orow[i] = 0u;
    }
    if (q_global < p.L) {
      float* dst = p.part_o + (static_cast<size_t>(chunk) * p.L + q_global) * kC + head * kHeadDim;
#pragma unroll
      for (int i = 0; i < 32; i += 4)
        *reinterpret_cast<uint4*>(dst + i) = make_uint4(orow[i], orow[i + 1], orow[i + 2], orow[i + 3]);
      const size_t ml = (static_cast<size_t>(chunk) * kHeads + head) * p.L + q_global;
      p.part_m[ml] = m_used;
      p.part_l[ml] = l_sum;
    }
    tc_fence_before();
  }

  __syncthreads();
  if (warp == 1) {
    tc_fence_after();
    tmem_dealloc<kTmemCols>(tmem_base);
  }
}

// ---------------------------------------------------------------------------------------------
// combine: per head, out[q, head*32 + i] = sum_c w_c O_c / sum_c w_c l_c with
//   w_c = 2^(m_c + bias[q][head][row(t_c)] - M);  W[q,t] = mean_heads sum_{c in t} w_c l_c / denom_head
struct MhaCombineParams {
  int L, T, chunks_per_frame;
  const float* part_o;     // [n_chunks][L][256]
  const float* part_m;     // [n_chunks][8][L]
  const float* part_l;
  const float* tpe_bias;   // [L][8][4] log2 units, or nullptr
  int tpe_rows[RMEM_MAX_BANK];
  __nv_bfloat16* out;      // [L][ld_out] at column out_col0
  int ld_out, out_col0;
  float* attn_mass;        // [L][T] or nullptr
};

__global__ void __launch_bounds__(256) mha_combine_kernel(const MhaCombineParams p) {
  constexpr int kMaxChunks = RMEM_MAX_BANK * RMEM_MAX_CHUNKS_PER_FRAME;
  const int q = blockIdx.x;
  const int n_chunks = p.T * p.chunks_per_frame;
  __shared__ float w_sh[mha::kHeads][kMaxChunks];
  __shared__ float mass_sh[mha::kHeads][RMEM_MAX_BANK];
  const int head = threadIdx.x >> 5, lane = threadIdx.x & 31;   // warp = head
  {
    float mm[2], ll[2], M = -INFINITY;
#pragma unroll
    for (int k = 0; k < 2; ++k) {
      const int cidx = lane + 32 * k;
      mm[k] = -INFINITY;
      ll[k] = 0.f;
      if (cidx < n_chunks) {
        const int t = cidx / p.chunks_per_frame;
        const size_t ml = (static_cast<size_t>(cidx) * mha::kHeads + head) * p.L + q;
        float m = p.part_m[ml];
        const float l = p.part_l[ml];
        if (p.tpe_bias) m += p.tpe_bias[(static_cast<size_t>(q) * mha::kHeads + head) * 4 + p.tpe_rows[t]];
        if (l > 0.f) { mm[k] = m; ll[k] = l; }
      }
      M = fmaxf(M, mm[k]);
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) M = fmaxf(M, __shfl_xor_sync(0xffffffffu, M, o));
    float w[2], denom = 0.f;
#pragma unroll
    for (int k = 0; k < 2; ++k) {
      w[k] = (mm[k] == -INFINITY) ? 0.f : exp2f(mm[k] - M);
      denom += w[k] * ll[k];
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) denom += __shfl_xor_sync(0xffffffffu, denom, o);
    const float inv = 1.f / denom;
    if (lane < RMEM_MAX_BANK) mass_sh[head][lane] = 0.f;
    __syncwarp();
#pragma unroll
    for (int k = 0; k < 2; ++k) {
      const int cidx = lane + 32 * k;
      if (cidx < n_chunks) {
        w_sh[head][cidx] = w[k] * inv;
        atomicAdd(&mass_sh[head][cidx / p.chunks_per_frame], w[k] * ll[k] * inv);
      }
    }
  }
  __syncthreads();
  if (p.attn_mass && threadIdx.x < p.T) {
    float s = 0.f;
#pragma unroll
    for (int h = 0; h < mha::kHeads; ++h) s += mass_sh[h][threadIdx.x];
    p.attn_mass[static_cast<size_t>(q) * p.T + threadIdx.x] = s * (1.f / mha::kHeads);
  }
  const int c = threadIdx.x;                       // channel; head = c / 32 = warp
  float acc = 0.f;
  for (int cidx = 0; cidx < n_chunks; ++cidx) {
    const float w = w_sh[head][cidx];
    if (w == 0.f) continue;
    acc += w * p.part_o[(static_cast<size_t>(cidx) * p.L + q) * mha::kC + c];
  }
  p.out[static_cast<size_t>(q) * p.ld_out + p.out_col0 + c] = __float2bfloat16(acc);
}

// ---------------------------------------------------------------------------------------------
// TPE preparation, AOT: q_time = bf16(q + cur); bias[q][head][r] = log2e/sqrt(32) * (q_time_head . mem[r]_head)
__global__ void __launch_bounds__(256)
tpe_prepare_mha_kernel(const __nv_bfloat16* __restrict__ q, int ldq, const float* __restrict__ cur,
                       const float* __restrict__ mem, __nv_bfloat16* __restrict__ q_time, float* __restrict__ bias,
                       int L, float scale) {
  const int row = blockIdx.x * 8 + (threadIdx.x >> 5);
  const int lane = threadIdx.x & 31;
  if (row >= L) return;
  const uint4 raw = *reinterpret_cast<const uint4*>(q + static_cast<size_t>(row) * ldq + lane * 8);
  const __nv_bfloat162* hv = reinterpret_cast<const __nv_bfloat162*>(&raw);
  float v[8];
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    v[2 * i] = __bfloat162float(hv[i].x) + cur[lane * 8 + 2 * i];
    v[2 * i + 1] = __bfloat162float(hv[i].y) + cur[lane * 8 + 2 * i + 1];
  }
  uint4 pk;
  __nv_bfloat162* o = reinterpret_cast<__nv_bfloat162*>(&pk);
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    o[i] = __floats2bfloat162_rn(v[2 * i], v[2 * i + 1]);
    v[2 * i] = __bfloat162float(o[i].x);           // the bias uses the rounded q' the tensor cores see
    v[2 * i + 1] = __bfloat162float(o[i].y);
  }
  *reinterpret_cast<uint4*>(q_time + static_cast<size_t>(row) * mha::kC + lane * 8) = pk;
#pragma unroll
  for (int r = 0; r < RMEM_TPE_ROWS; ++r) {
    float d = 0.f;
#pragma unroll
    for (int i = 0; i < 8; ++i) d += v[i] * mem[r * mha::kC + lane * 8 + i];
    d += __shfl_xor_sync(0xffffffffu, d, 1);       // 4 lanes = one head (32 channels)
    d += __shfl_xor_sync(0xffffffffu, d, 2);
    if ((lane & 3) == 0) bias[(static_cast<size_t>(row) * mha::kHeads + (lane >> 2)) * 4 + r] = d * scale;
  }
}

// ---------------------------------------------------------------------------------------------
static int mha_encode(CUtensorMap* map, const void* base, int rank, const uint64_t* dims,
                      const uint64_t* strides_bytes, const uint32_t* box) {
  auto fn = get_tensor_map_encoder();
  if (!fn) return set_error(RMEM_ERR_DRIVER, "cuTensorMapEncodeTiled entry point not found");
  const uint32_t es[5] = {1, 1, 1, 1, 1};
  CUresult r = fn(map, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, rank, const_cast<void*>(base), dims, strides_bytes,
                  box, es, CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B,
                  CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) return set_error(RMEM_ERR_DRIVER, "cuTensorMapEncodeTiled (mha) failed");
  return RMEM_OK;
}

int mha_partial_launch(const void* q, int ldq, const void* k, int ldk, int64_t k_slot_stride, const void* v,
                       int ldv, int64_t v_slot_stride, const int32_t* order_dev, int T, int L, int n_slots,
                       int chunks_per_frame, float* part_o, float* part_m, float* part_l,
                       cudaStream_t stream) {
  using namespace mha;
  if (T < 1 || T > RMEM_MAX_BANK || chunks_per_frame < 1 || chunks_per_frame > RMEM_MAX_CHUNKS_PER_FRAME ||
      L < 1 || n_slots < T || (ldq % 8) || (ldk % 8) || (ldv % 8) || (k_slot_stride % 8) || (v_slot_stride % 8))
    return set_error(RMEM_ERR_ARG, "mha: bad T / L / chunks_per_frame / n_slots / leading dimension");
  CUtensorMap tm_q, tm_k, tm_v;
  {
    const uint64_t dims[3] = {64, static_cast<uint64_t>(L), kC / 64};
    const uint64_t strides[2] = {static_cast<uint64_t>(ldq) * 2, 128};
    const uint32_t box[3] = {64, kBlockM, 1};
    if (int e = mha_encode(&tm_q, q, 3, dims, strides, box)) return e;
  }
  {
    const uint64_t dims[4] = {64, static_cast<uint64_t>(L), kC / 64, static_cast<uint64_t>(n_slots)};
    const uint64_t strides[3] = {static_cast<uint64_t>(ldk) * 2, 128,
                                 static_cast<uint64_t>(n_slots > 1 ? k_slot_stride : static_cast<int64_t>(L) * ldk) * 2};
    const uint32_t box[4] = {64, kBlockN, 1, 1};
    if (int e = mha_encode(&tm_k, k, 4, dims, strides, box)) return e;
  }
  {
    const uint64_t dims[4] = {64, static_cast<uint64_t>(L), kC / 64, static_cast<uint64_t>(n_slots)};
    const uint64_t strides[3] = {static_cast<uint64_t>(ldv) * 2, 128,
                                 static_cast<uint64_t>(n_slots > 1 ? v_slot_stride : static_cast<int64_t>(L) * ldv) * 2};
    const uint32_t box[4] = {64, kBlockN, 1, 1};
    if (int e = mha_encode(&tm_v, v, 4, dims, strides, box)) return e;
  }
  static SmemOptIn optin;
  if (int r = ensure_dynamic_smem(reinterpret_cast<const void*>(mha_partial_kernel), &optin, kSmemTotal,
                                  "mha: cudaFuncSetAttribute"))
    return r;
  Params p;
  p.L = L;
  p.n_chunks = T * chunks_per_frame;
  p.chunks_per_frame = chunks_per_frame;
  p.order = order_dev;
  p.part_o = part_o;
  p.part_m = part_m;
  p.part_l = part_l;
  p.scale_log2 = 1.4426950408889634f / sqrtf(static_cast<float>(kHeadDim));
  dim3 grid((L + kBlockM - 1) / kBlockM, kHeads / 2, p.n_chunks);
  mha_partial_kernel<<<grid, kThreads, kSmemTotal, stream>>>(tm_q, tm_k, tm_v, p);
  return check_launch("mha_partial_kernel");
}

int mha_combine_launch(const float* part_o, const float* part_m, const float* part_l, const float* tpe_bias,
                       const int32_t* tpe_rows_host, int T, int L, int chunks_per_frame, void* out, int ld_out,
                       int out_col0, float* attn_mass, cudaStream_t stream) {
  if (T < 1 || T > RMEM_MAX_BANK || chunks_per_frame < 1 || chunks_per_frame > RMEM_MAX_CHUNKS_PER_FRAME)
    return set_error(RMEM_ERR_ARG, "mha_combine: bad T / chunks_per_frame");
  MhaCombineParams p;
  p.L = L; p.T = T; p.chunks_per_frame = chunks_per_frame;
  p.part_o = part_o; p.part_m = part_m; p.part_l = part_l;
  p.tpe_bias = tpe_bias;
  for (int t = 0; t < RMEM_MAX_BANK; ++t) p.tpe_rows[t] = (tpe_rows_host && t < T) ? tpe_rows_host[t] : 0;
  p.out = static_cast<__nv_bfloat16*>(out);
  p.ld_out = ld_out; p.out_col0 = out_col0;
  p.attn_mass = attn_mass;
  mha_combine_kernel<<<L, 256, 0, stream>>>(p);
  return check_launch("mha_combine_kernel");
}

int tpe_prepare_mha_launch(const void* q, int ldq, const float* cur, const float* mem, void* q_time, float* bias,
                           int L, cudaStream_t stream) {
  if (L < 1 || (ldq % 8)) return set_error(RMEM_ERR_ARG, "tpe_prepare_mha: bad L / ldq");
  tpe_prepare_mha_kernel<<<(L + 7) / 8, 256, 0, stream>>>(
      static_cast<const __nv_bfloat16*>(q), ldq, cur, mem, static_cast<__nv_bfloat16*>(q_time), bias, L,
      1.4426950408889634f / sqrtf(static_cast<float>(mha::kHeadDim)));
  return check_launch("tpe_prepare_mha_kernel");
}

}  // namespace rmem
