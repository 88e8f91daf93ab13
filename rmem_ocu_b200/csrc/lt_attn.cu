// Long-term (global) gated attention of the current frame's tokens against the restricted memory
// bank -- DeAOT shape: 1 head, d_qk = 128, d_v = 1024 (V | ID_V), bf16 in, fp32 accumulate.
//
// Replaces the explicit softmax(QK^T)V path of GatedPropagation.forward
// (reference aot_plus/networks/layers/attention.py:174-193) together with the attention-weight
// recording of GatedPropagationModule.forward (layers/transformer.py:1185-1192) and the temporal
// positional embedding on the bank keys (transformer.py:1140-1172).
//
// Decomposition (split-KV by memory frame):
//   rmem_lt_attn_partial : grid (q_tile, dv_split, chunk).  One CTA owns 128 queries x 256 value
//       channels x one key chunk (a memory frame, or a part of one).  TMA streams 128-key K and V
//       tiles through two 2-stage rings; one elected thread issues tcgen05.mma: S = Q K^T into a
//       double-buffered 128-column TMEM tile, O += P V into a 256-column TMEM accumulator with P
//       read straight from TMEM (the bf16 probabilities overwrite the S tile in place, so P never
//       touches shared memory).  Eight softmax warps -- two per TMEM lane quarter, each thread
//       owning half of one query row -- read S with tcgen05.ld, exchange the row max through
//       shared memory, exponentiate with packed f32x2 FMAs + MUFU.EX2, rescale O lazily and
//       write P with tcgen05.st.  Output: un-normalised O, running max m (log2 units) and sum l.
//   rmem_lt_attn_combine : merges the chunk partials, adds the per-frame TPE bias
//       (q'.e_row -- a rank-1 term, so K is never rewritten), emits the per-frame attention mass
//       W[q,t] (= record_attn_weight), applies the gate U and writes bf16.
#include <cstdlib>
#include "sm100_ptx.cuh"
#include "rmem_internal.h"

namespace rmem {
using namespace sm100;

constexpr int kBlockM = 128;      // queries per CTA
constexpr int kBlockN = 128;      // keys per pipeline tile
constexpr int kDK = 128;          // d_att
constexpr int kDV = 1024;         // expand_d_vu (V | ID_V)
constexpr int kDVTile = 256;      // value channels per CTA (TMEM accumulator columns)
constexpr int kDVSplits = kDV / kDVTile;
constexpr int kKStages = 2, kVStages = 2;
constexpr int kSoftmaxWarps = 8;
constexpr int kThreads = 64 + 32 * kSoftmaxWarps;   // warp0 TMA, warp1 MMA, warps2-9 softmax/epilogue

constexpr uint32_t kQBytes = kBlockM * kDK * 2;            // 32 KB  [2][128][64]
constexpr uint32_t kKBytes = kBlockN * kDK * 2;            // 32 KB  [2][128][64]
constexpr uint32_t kVBytes = kBlockN * kDVTile * 2;        // 64 KB  [4][128][64]
constexpr uint32_t kXchBytes = 2 * 2 * kBlockM * 4;        // row-max exchange, one buffer per S buffer
constexpr uint32_t kSmemData = kQBytes + kKStages * kKBytes + kVStages * kVBytes;
constexpr uint32_t kSmemTotal = kSmemData + kXchBytes + 256 /*barriers*/;
static_assert(kSmemTotal <= 232448, "exceeds the 227 KB shared memory of an sm_100 CTA");

constexpr uint32_t kTmemCols = 512;
constexpr uint32_t kTmemS0 = 0;      // S / P buffers: 2 x 128 columns (P = 64 columns, aliased on S)
constexpr uint32_t kTmemO = 256;     // O: 256 columns

constexpr float kRescaleThreshold = 8.0f;  // log2 units
constexpr int kPolyEvery = 3;              // every 3rd pair of exponentials runs on the FMA pipes (~1/3)

struct LtParams {
  int L;                 // tokens per frame
  int n_chunks;          // total key chunks = T * chunks_per_frame
  int chunks_per_frame;
  const int32_t* order;  // [T] logical bank position -> physical slot (device)
  float* part_o;         // [n_chunks][L][1024]
  float* part_m;         // [n_chunks][L]
  float* part_l;         // [n_chunks][L]
  float scale_log2;      // log2(e) / sqrt(d_att)
  // direct mode (n_chunks == 1: the one chunk holds every key, so the softmax is complete in this CTA): the epilogue
  // normalises, applies the gate and writes bf16 -- no partials, no combine launch
  const __nv_bfloat16* direct_gate;   // [L][1024] or nullptr
  __nv_bfloat16* direct_out;          // [L][1024]; nullptr = partial mode
  // direct mode on a 2-CTA cluster (gridDim.z = 2, cluster dims (1, 1, 2)): the CTAs take one half of the keys each;
  // CTA 1 ships its un-normalised O, m and l into CTA 0's idle K / V rings through distributed shared memory and
  // CTA 0 merges, normalises, gates and stores -- half the key tiles per CTA, twice the CTAs (104 instead of 52)
  int direct_pair;
};

// cluster primitives used by the paired direct mode
__device__ __forceinline__ uint32_t lt_cluster_rank() {
  uint32_t r;
  asm volatile("mov.u32 %0, %%cluster_ctarank;" : "=r"(r));
  return r;
}
__device__ __forceinline__ void lt_cluster_sync() {
  asm volatile("barrier.cluster.arrive.release.aligned;\n\tbarrier.cluster.wait.acquire.aligned;" ::: "memory");
}
__device__ __forceinline__ uint32_t lt_map_to_cta(uint32_t smem_addr, uint32_t rank) {
  uint32_t r;
  asm volatile("mapa.shared::cluster.u32 %0, %1, %2;" : "=r"(r) : "r"(smem_addr), "r"(rank));
  return r;
}
__device__ __forceinline__ void lt_st_cluster_f4(uint32_t addr, float4 v) {
  asm volatile("st.shared::cluster.v4.f32 [%0], {%1, %2, %3, %4};" ::"r"(addr), "f"(v.x), "f"(v.y), "f"(v.z), "f"(v.w)
               : "memory");
}
__device__ __forceinline__ void lt_st_cluster_f1(uint32_t addr, float v) {
  asm volatile("st.shared::cluster.f32 [%0], %1;" ::"r"(addr), "f"(v) : "memory");
}

__global__ void __launch_bounds__(kThreads, 1)
lt_attn_partial_kernel(const __grid_constant__ CUtensorMap tm_q,
                       const __grid_constant__ CUtensorMap tm_k,
                       const __grid_constant__ CUtensorMap tm_v,
                       const __grid_constant__ CUtensorMap tm_o, const LtParams p) {
  // 128B-swizzled TMA / UMMA tiles need a 1024-byte aligned base; the dynamic shared window of a kernel
  // without static shared memory starts 1024-aligned, the declared alignment makes that a requirement.
  extern __shared__ __align__(1024) uint8_t smem[];
  uint8_t* sQ = smem;
  uint8_t* sK = sQ + kQBytes;
  uint8_t* sV = sK + kKStages * kKBytes;
  float* xch = reinterpret_cast<float*>(sV + kVStages * kVBytes);   // [2][2][128]
  uint64_t* bars = reinterpret_cast<uint64_t*>(xch + 2 * 2 * kBlockM);
  uint64_t* q_full = bars + 0;
  uint64_t* k_full = bars + 1;             // [2]
  uint64_t* k_empty = bars + 3;            // [2]
  uint64_t* v_full = bars + 5;             // [2]
  uint64_t* v_empty = bars + 7;            // [2]
  uint64_t* s_full = bars + 9;             // [2]
  uint64_t* p_full = bars + 11;            // [2]
  uint64_t* pv_done = bars + 13;           // [2]
  uint64_t* o_full = bars + 15;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + 16);

  // warp-uniform warp index (shuffle from lane 0), so that the MMA warp's operands live in uniform registers
  const int warp = __shfl_sync(0xffffffffu, static_cast<int>(threadIdx.x >> 5), 0);
  const int lane = threadIdx.x & 31;
  const int q_tile = blockIdx.x;
  const int dv_split = blockIdx.y;
  const int chunk = blockIdx.z;

  // key range of this chunk
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
    tma_prefetch_desc(&tm_o);
    mbar_init(q_full, 1);
    for (int i = 0; i < 2; ++i) {
      mbar_init(k_full + i, 1);
      mbar_init(k_empty + i, 1);
      mbar_init(v_full + i, 1);
      mbar_init(v_empty + i, 1);
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
  const uint32_t tmem_base = __shfl_sync(0xffffffffu, *tmem_slot, 0);
  pdl_wait();        // programmatic dependent launch: see rmem_internal.h
  pdl_trigger();

  if (warp == 0) {
    // ------------------------------------------------------------------ TMA producer
    if (lane == 0 && n_tiles > 0) {   // never leave a bulk copy in flight when there is no consumer
      const int slot = p.order ? p.order[frame_pos] : 0;
      mbar_expect_tx(q_full, kQBytes);
      tma_load_3d(sQ, &tm_q, q_full, 0, q_tile * kBlockM, 0);
      for (int j = 0; j < n_tiles; ++j) {
        const int st = j & 1;
        const uint32_t ph = ((j >> 1) - 1) & 1;
        if (j >= 2) mbar_wait(k_empty + st, ph);
        mbar_expect_tx(k_full + st, kKBytes);
        tma_load_4d(sK + st * kKBytes, &tm_k, k_full + st, 0, key_begin + j * kBlockN, 0, slot);
        if (j >= 2) mbar_wait(v_empty + st, ph);
        mbar_expect_tx(v_full + st, kVBytes);
        tma_load_4d(sV + st * kVBytes, &tm_v, v_full + st, 0, key_begin + j * kBlockN, dv_split * 4, slot);
      }
    }
    __syncwarp();                                                  // lanes 1..31 wait for the producing lane
    if (p.direct_pair) { lt_cluster_sync(); lt_cluster_sync(); }   // the two barriers of the paired epilogue (below)
  } else if (warp == 1) {
    // ------------------------------------------------------------------ MMA issuer
    // (One issuing warp: giving Q K^T and P V a warp each, as the stream-K kernel does, changed the self-attention from
    // 25.3 to 26.9 us -- here the softmax warps set the pace.)
    // The whole warp runs the loop and the waits; one elected lane issues.  (Inside an `if (lane == 0)` region the
    // compiler has to assume per-thread operands and wraps every tcgen05.mma in an ELECT / R2UR.BROADCAST loop,
    // ~90 cycles per MMA.)
    if (n_tiles > 0) {
      constexpr uint32_t idesc_qk = make_idesc_bf16(kBlockM, kBlockN, 0, 0);
      constexpr uint32_t idesc_pv = make_idesc_bf16(kBlockM, kDVTile, 0, 1);
      const uint32_t q_addr = smem_u32(sQ);
      auto issue_qk = [&](int j) {
        const int st = j & 1;
        mbar_wait(k_full + st, (j >> 1) & 1);
        tc_fence_after();
        const uint32_t k_addr = smem_u32(sK + st * kKBytes);
        const uint32_t d = tmem_base + kTmemS0 + (j & 1) * kBlockN;
        if (elect_one()) {
#pragma unroll
          for (int ks = 0; ks < kDK / 16; ++ks) {
            const uint32_t blk = ks >> 2, within = ks & 3;
            const uint64_t a = make_desc_sw128(q_addr + blk * (kBlockM * 128) + within * 32, 16, 1024);
            const uint64_t b = make_desc_sw128(k_addr + blk * (kBlockN * 128) + within * 32, 16, 1024);
            mma_bf16_ss(d, a, b, idesc_qk, ks > 0);
          }
          tc_commit(k_empty + st);
          tc_commit(s_full + (j & 1));
        }
        __syncwarp();
      };
      mbar_wait(q_full, 0);
      issue_qk(0);
      for (int j = 0; j < n_tiles; ++j) {
        // S(j+1) goes into the buffer P(j-1) lived in: PV(j-1) was issued before it, and tcgen05.mma
        // instructions of one thread execute in issue order.
        if (j + 1 < n_tiles) issue_qk(j + 1);
        mbar_wait(p_full + (j & 1), (j >> 1) & 1);
        const int st = j & 1;
        mbar_wait(v_full + st, (j >> 1) & 1);
        tc_fence_after();
        const uint32_t v_addr = smem_u32(sV + st * kVBytes);
        const uint32_t p_tmem = tmem_base + kTmemS0 + (j & 1) * kBlockN;   // bf16 P: 64 columns
        const uint32_t d = tmem_base + kTmemO;
        if (elect_one()) {
#pragma unroll
          for (int ks = 0; ks < kBlockN / 16; ++ks) {
            const uint64_t b = make_desc_sw128(v_addr + ks * (16 * 128), kBlockN * 128, 1024);
            mma_bf16_ts(d, p_tmem + ks * 8, b, idesc_pv, (j > 0) || (ks > 0));
          }
          tc_commit(v_empty + st);
          tc_commit(pv_done + (j & 1));
        }
        __syncwarp();
      }
      if (elect_one()) tc_commit(o_full);
      __syncwarp();
    }
    if (p.direct_pair) { lt_cluster_sync(); lt_cluster_sync(); }   // the two barriers of the paired epilogue (below)
  } else {
    // ------------------------------------------------------------------ softmax + epilogue
    const int quarter = warp & 3;                 // TMEM lane quarter this warp may access
    const int hf = (warp - 2) >> 2;               // which half of the row's columns this thread owns
    const int row = quarter * 32 + lane;          // query row inside the tile
    const uint32_t lane_addr = static_cast<uint32_t>(quarter * 32) << 16;
    const uint32_t pair_bar = 1 + quarter;        // named barrier shared by the two warps of a quarter
    const int q_global = q_tile * kBlockM + row;
    float m_used = -INFINITY;
    float l_sum = 0.f;
    const float c = p.scale_log2;

    for (int j = 0; j < n_tiles; ++j) {
      const int b = j & 1;
      mbar_wait(s_full + b, (j >> 1) & 1);
      tc_fence_after();
#ifdef LT_EXP_NOSOFTMAX
      tc_fence_before();
      mbar_arrive(p_full + b);
      continue;
#endif
      uint32_t sr[64];
      {
        uint32_t(&lo)[32] = *reinterpret_cast<uint32_t(*)[32]>(&sr[0]);
        uint32_t(&hi)[32] = *reinterpret_cast<uint32_t(*)[32]>(&sr[32]);
        const uint32_t s_addr = tmem_base + lane_addr + kTmemS0 + b * kBlockN + hf * 64;
        tmem_ld32(s_addr, lo);
        tmem_ld32(s_addr + 32, hi);
        tmem_ld_wait();
      }
#ifdef LT_EXP_LDONLY
      if (sr[0] == 0x12345678u && sr[63] == 0x9abcdef0u) l_sum += 1.f;
      tc_fence_before();
      mbar_arrive(p_full + b);
      continue;
#endif
      const int valid = key_len - j * kBlockN - hf * 64;   // valid columns in this half (may be <= 0)
      if (valid < 64) {                                       // warp-uniform: only the chunk's last tile
#pragma unroll
        for (int i = 0; i < 64; ++i)
          if (i >= valid) sr[i] = 0xff800000u;                // -inf
      }
      float mx = -INFINITY;
#pragma unroll
      for (int i = 0; i < 64; i += 2)
        mx = fmaxf(mx, fmaxf(__uint_as_float(sr[i]), __uint_as_float(sr[i + 1])));
      // row max across the two halves (the partner thread lives in warp +-4)
      xch[(b * 2 + hf) * kBlockM + row] = mx;
      named_bar_sync(pair_bar, 64);
      mx = fmaxf(mx, xch[(b * 2 + (hf ^ 1)) * kBlockM + row]) * c;
      // lazy rescale: keep the reference max unless it grew by more than the threshold
      const bool grow = mx > m_used + kRescaleThreshold;
      float scale = 1.f;
      float m_new = m_used;
      if (grow) {
        m_new = mx;
        scale = fast_exp2(m_used - m_new);            // 0 when m_used == -inf
      }
      if (j > 0 && __any_sync(0xffffffffu, grow)) {    // same decision in both warps of the pair
        mbar_wait(pv_done + ((j - 1) & 1), ((j - 1) >> 1) & 1);
        tc_fence_after();
        const uint32_t o_addr = tmem_base + lane_addr + kTmemO + hf * (kDVTile / 2);
#pragma unroll 1
        for (int cb = 0; cb < kDVTile / 64; ++cb) {
          uint32_t orow[32];
          tmem_ld32(o_addr + cb * 32, orow);
          tmem_ld_wait();
#pragma unroll
          for (int i = 0; i < 32; ++i) orow[i] = __float_as_uint(__uint_as_float(orow[i]) * scale);
          tmem_st32(o_addr + cb * 32, orow);
        }
      }
      l_sum *= scale;
      m_used = m_new;
      const float neg_m = -m_used;
      float acc0 = 0.f, acc1 = 0.f;
      uint32_t packed[32];
#pragma unroll
      for (int i = 0; i < 64; i += 2) {
        float x0, x1;
        fma_f32x2(x0, x1, __uint_as_float(sr[i]), __uint_as_float(sr[i + 1]), c, c, neg_m, neg_m);
        float p0, p1;
        if ((i >> 1) % kPolyEvery == kPolyEvery - 1) {
          exp2_poly_x2(p0, p1, x0, x1);             // FMA / ALU pipes
        } else {
          p0 = fast_exp2(x0);                       // MUFU.EX2
          p1 = fast_exp2(x1);
        }
        add_f32x2(acc0, acc1, acc0, acc1, p0, p1);
        packed[i >> 1] = pack_bf16x2(p0, p1);
      }
      l_sum += acc0 + acc1;
      // P (bf16, 64 columns) overwrites the first half of this S buffer; both partner threads have
      // their S values in registers (they passed tcgen05.wait::ld before the pair barrier).
      tmem_st32(tmem_base + lane_addr + kTmemS0 + b * kBlockN + hf * 32, packed);
      tmem_st_wait();
      tc_fence_before();
      mbar_arrive(p_full + b);
    }

    // epilogue: O (fp32, un-normalised) -> part_o ; m, l -> part_m / part_l
    if (n_tiles > 0) {
      mbar_wait(o_full, 0);
      tc_fence_after();
    }
    // l of the two halves: exchanged through the buffer the last tile did NOT use (its last readers
    // passed the last tile's pair barrier)
    const int xb = (n_tiles & 1);                 // last tile used buffer (n_tiles - 1) & 1
    xch[(xb * 2 + hf) * kBlockM + row] = l_sum;
    named_bar_sync(pair_bar, 64);
    const float l_tot = l_sum + xch[(xb * 2 + (hf ^ 1)) * kBlockM + row];
    const bool row_ok = q_global < p.L;
    // O tile -> shared memory (the K / V rings are idle now: o_full means every MMA has completed) as
    // eight [128 rows][32 fp32] blocks in the 128B-swizzled layout, then one thread issues TMA stores:
    // full-line writes, rows beyond L clipped by the tensor map.
    uint8_t* stage = sK;
    const uint32_t o_addr = tmem_base + lane_addr + kTmemO + hf * (kDVTile / 2);
    if (p.direct_pair) {
      // staging in CTA 0's K / V rings: O as [64 column quads][128 rows] float4 (a warp's 32 rows are 512 contiguous
      // bytes), then m [128] and l [128]
      const uint32_t rank = lt_cluster_rank();
      const uint32_t stage_addr = smem_u32(stage);
      constexpr uint32_t kStageML = kBlockM * kDVTile * 4;
      lt_cluster_sync();                 // A: both CTAs have passed o_full -- CTA 0's rings are idle
      if (rank == 1) {
        const uint32_t remote = lt_map_to_cta(stage_addr, 0);
#pragma unroll 1
        for (int cb = 0; cb < kDVTile / 64; ++cb) {
          uint32_t orow[32];
          if (n_tiles > 0) {
            tmem_ld32(o_addr + cb * 32, orow);
            tmem_ld_wait();
          } else {
#pragma unroll
            for (int i = 0; i < 32; ++i) orow[i] = 0u;
          }
#pragma unroll
          for (int v = 0; v < 8; ++v) {
            const int c4 = hf * 32 + cb * 8 + v;               // column quad of the 256-column tile
            lt_st_cluster_f4(remote + (static_cast<uint32_t>(c4) * kBlockM + row) * 16,
                             make_float4(__uint_as_float(orow[4 * v]), __uint_as_float(orow[4 * v + 1]),
                                         __uint_as_float(orow[4 * v + 2]), __uint_as_float(orow[4 * v + 3])));
          }
        }
        if (hf == 0) {
          lt_st_cluster_f1(remote + kStageML + row * 4, n_tiles > 0 ? m_used : -INFINITY);
          lt_st_cluster_f1(remote + kStageML + (kBlockM + row) * 4, n_tiles > 0 ? l_tot : 0.f);
        }
      }
      lt_cluster_sync();                 // B: CTA 1's stores are visible in CTA 0's shared memory
      if (rank == 0) {
        const float* ml = reinterpret_cast<const float*>(stage + kStageML);
        const float m1 = ml[row], l1 = ml[kBlockM + row];
        const float m0 = n_tiles > 0 ? m_used : -INFINITY, l0 = n_tiles > 0 ? l_tot : 0.f;
        const float M = fmaxf(m0, m1);
        const float w0 = m0 == -INFINITY ? 0.f : exp2f(m0 - M), w1 = m1 == -INFINITY ? 0.f : exp2f(m1 - M);
        const float denom = w0 * l0 + w1 * l1;
        const float inv = denom > 0.f ? 1.f / denom : 0.f;
        const float s0 = w0 * inv, s1 = w1 * inv;
        const size_t base = static_cast<size_t>(q_global) * kDV + dv_split * kDVTile + hf * (kDVTile / 2);
        const float4* st4p = reinterpret_cast<const float4*>(stage);
#pragma unroll 1
        for (int cb = 0; cb < kDVTile / 64; ++cb) {
          uint32_t orow[32];
          if (n_tiles > 0) {
            tmem_ld32(o_addr + cb * 32, orow);
            tmem_ld_wait();
          } else {
#pragma unroll
            for (int i = 0; i < 32; ++i) orow[i] = 0u;
          }
          if (row_ok) {
            uint4 g[4];
            if (p.direct_gate) {
#pragma unroll
              for (int v = 0; v < 4; ++v)
                g[v] = *reinterpret_cast<const uint4*>(p.direct_gate + base + cb * 32 + v * 8);
            }
#pragma unroll
            for (int v = 0; v < 4; ++v) {
              const float4 r0 = st4p[(hf * 32 + cb * 8 + 2 * v) * kBlockM + row];
              const float4 r1 = st4p[(hf * 32 + cb * 8 + 2 * v + 1) * kBlockM + row];
              const float other[8] = {r0.x, r0.y, r0.z, r0.w, r1.x, r1.y, r1.z, r1.w};
              uint32_t w[4];
#pragma unroll
              for (int e = 0; e < 4; ++e) {
                float a = __uint_as_float(orow[v * 8 + 2 * e]) * s0 + other[2 * e] * s1;
                float b = __uint_as_float(orow[v * 8 + 2 * e + 1]) * s0 + other[2 * e + 1] * s1;
                if (p.direct_gate) {
                  const uint32_t gw = (&g[v].x)[e];
                  const __nv_bfloat162 g2 = *reinterpret_cast<const __nv_bfloat162*>(&gw);
                  a *= __bfloat162float(g2.x);
                  b *= __bfloat162float(g2.y);
                }
                w[e] = pack_bf16x2(a, b);
              }
              *reinterpret_cast<uint4*>(p.direct_out + base + cb * 32 + v * 8) = make_uint4(w[0], w[1], w[2], w[3]);
            }
          }
        }
      }
    } else if (p.direct_out != nullptr) {
      // this thread: row q_global, value channels [dv_split * 256 + hf * 128, +128): 64-byte runs per 32 channels
      const float inv = (n_tiles > 0 && l_tot > 0.f) ? 1.f / l_tot : 0.f;
      const size_t base = static_cast<size_t>(q_global) * kDV + dv_split * kDVTile + hf * (kDVTile / 2);
#pragma unroll 1
      for (int cb = 0; cb < kDVTile / 64; ++cb) {
        uint32_t orow[32];
        if (n_tiles > 0) {
          tmem_ld32(o_addr + cb * 32, orow);
          tmem_ld_wait();
        } else {
#pragma unroll
          for (int i = 0; i < 32; ++i) orow[i] = 0u;
        }
        if (row_ok) {
          uint4 g[4];
          if (p.direct_gate) {
#pragma unroll
            for (int v = 0; v < 4; ++v)
              g[v] = *reinterpret_cast<const uint4*>(p.direct_gate + base + cb * 32 + v * 8);
          }
#pragma unroll
          for (int v = 0; v < 4; ++v) {
            uint32_t w[4];
#pragma unroll
            for (int e = 0; e < 4; ++e) {
              float a = __uint_as_float(orow[v * 8 + 2 * e]) * inv;
              float b = __uint_as_float(orow[v * 8 + 2 * e + 1]) * inv;
              if (p.direct_gate) {
                const uint32_t gw = (&g[v].x)[e];
                const __nv_bfloat162 g2 = *reinterpret_cast<const __nv_bfloat162*>(&gw);
                a *= __bfloat162float(g2.x);
                b *= __bfloat162float(g2.y);
              }
              w[e] = pack_bf16x2(a, b);
            }
            *reinterpret_cast<uint4*>(p.direct_out + base + cb * 32 + v * 8) = make_uint4(w[0], w[1], w[2], w[3]);
          }
        }
      }
    } else {
#pragma unroll 1
    for (int cb = 0; cb < kDVTile / 64; ++cb) {
      uint32_t orow[32];
      if (n_tiles > 0) {
        tmem_ld32(o_addr + cb * 32, orow);
        tmem_ld_wait();
      } else {
#pragma unroll
        for (int i = 0; i < 32; ++i) orow[i] = 0u;
      }
      uint8_t* blk = stage + (hf * 4 + cb) * (kBlockM * 128);
#pragma unroll
      for (int ch = 0; ch < 8; ++ch)
        *reinterpret_cast<uint4*>(blk + sw128_offset(row, ch)) =
            make_uint4(orow[4 * ch], orow[4 * ch + 1], orow[4 * ch + 2], orow[4 * ch + 3]);
    }
    fence_proxy_async_smem();
    named_bar_sync(5, 32 * kSoftmaxWarps);
    if (warp == 2 && lane == 0) {
#pragma unroll 1
      for (int cbk = 0; cbk < kDVTile / 32; ++cbk)
        tma_store_3d(&tm_o, stage + cbk * (kBlockM * 128), dv_split * kDVTile + cbk * 32, q_tile * kBlockM, chunk);
      tma_store_commit();
      tma_store_wait_read();
    }
    if (row_ok && dv_split == 0 && hf == 0) {
      p.part_m[static_cast<size_t>(chunk) * p.L + q_global] = m_used;
      p.part_l[static_cast<size_t>(chunk) * p.L + q_global] = l_tot;
    }
    }  // partial mode
    tc_fence_before();
  }

  __syncthreads();
  if (warp == 1) {
    tc_fence_after();
    tmem_dealloc<kTmemCols>(tmem_base);
  }
}

// ---------------------------------------------------------------------------------------------
// combine: out[q,:] = gate[q,:] * sum_c w_c O_c[q,:] / sum_c w_c l_c,  w_c = 2^(m_c + bias[q,row(t_c)] - M)
//          W[q,t]   = sum_{c in frame t} w_c l_c / sum_c w_c l_c
struct CombineParams {
  int L, T, chunks_per_frame;
  const float* part_o;
  const float* part_m;
  const float* part_l;
  const float* tpe_bias;      // [L][4] (log2 units) or nullptr
  int tpe_rows[RMEM_MAX_BANK];  // TPE table row per logical bank position
  const __nv_bfloat16* gate;  // [L][1024] or nullptr
  __nv_bfloat16* out;         // [L][1024]
  float* attn_mass;           // [L][T] or nullptr
};

constexpr int kMaxChunks = RMEM_MAX_BANK * RMEM_MAX_CHUNKS_PER_FRAME;   // 64

__global__ void __launch_bounds__(256)
lt_attn_combine_kernel(const CombineParams p) {
  pdl_wait();        // programmatic dependent launch: see rmem_internal.h
  pdl_trigger();
  const int q = blockIdx.x;
  const int n_chunks = p.T * p.chunks_per_frame;
  __shared__ float w_sh[kMaxChunks];      // w_c / denom
  __shared__ float wl_sh[kMaxChunks];     // w_c l_c / denom
  if (threadIdx.x < 32) {
    // lane c (and c + 32) owns one chunk; max / sum through warp shuffles
    const int lane = threadIdx.x;
    float mm[2], ll[2];
    float M = -INFINITY;
#pragma unroll
    for (int k = 0; k < 2; ++k) {
      const int cidx = lane + 32 * k;
      mm[k] = -INFINITY;
      ll[k] = 0.f;
      if (cidx < n_chunks) {
        const int t = cidx / p.chunks_per_frame;
        float m = p.part_m[static_cast<size_t>(cidx) * p.L + q];
        const float l = p.part_l[static_cast<size_t>(cidx) * p.L + q];
        if (p.tpe_bias) m += p.tpe_bias[q * 4 + p.tpe_rows[t]];
        if (l > 0.f) { mm[k] = m; ll[k] = l; }    // empty chunk: stays -inf / 0
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
#pragma unroll
    for (int k = 0; k < 2; ++k) {
      const int cidx = lane + 32 * k;
      if (cidx < n_chunks) {
        w_sh[cidx] = w[k] * inv;
        wl_sh[cidx] = w[k] * ll[k] * inv;
      }
    }
    __syncwarp();
    if (p.attn_mass && lane < p.T) {
      float sacc = 0.f;
      for (int k = 0; k < p.chunks_per_frame; ++k) sacc += wl_sh[lane * p.chunks_per_frame + k];
      p.attn_mass[static_cast<size_t>(q) * p.T + lane] = sacc;
    }
  }
  __syncthreads();
  const int col = threadIdx.x * 4;
  float4 acc = make_float4(0.f, 0.f, 0.f, 0.f);
  for (int cidx = 0; cidx < n_chunks; ++cidx) {
    const float w = w_sh[cidx];
    if (w == 0.f) continue;
    const float4 o = *reinterpret_cast<const float4*>(
        p.part_o + (static_cast<size_t>(cidx) * p.L + q) * kDV + col);
    acc.x += w * o.x; acc.y += w * o.y; acc.z += w * o.z; acc.w += w * o.w;
  }
  if (p.gate) {
    const uint2 g = *reinterpret_cast<const uint2*>(p.gate + static_cast<size_t>(q) * kDV + col);
    const __nv_bfloat162 g01 = *reinterpret_cast<const __nv_bfloat162*>(&g.x);
    const __nv_bfloat162 g23 = *reinterpret_cast<const __nv_bfloat162*>(&g.y);
    acc.x *= __bfloat162float(g01.x); acc.y *= __bfloat162float(g01.y);
    acc.z *= __bfloat162float(g23.x); acc.w *= __bfloat162float(g23.y);
  }
  const __nv_bfloat162 o01 = __floats2bfloat162_rn(acc.x, acc.y);
  const __nv_bfloat162 o23 = __floats2bfloat162_rn(acc.z, acc.w);
  uint2 packed;
  packed.x = *reinterpret_cast<const uint32_t*>(&o01);
  packed.y = *reinterpret_cast<const uint32_t*>(&o23);
  *reinterpret_cast<uint2*>(p.out + static_cast<size_t>(q) * kDV + col) = packed;
}

// ---------------------------------------------------------------------------------------------
static int encode_map(CUtensorMap* map, const void* base, int rank, const uint64_t* dims,
                      const uint64_t* strides_bytes /*rank-1*/, const uint32_t* box) {
  auto fn = get_tensor_map_encoder();
  if (!fn) return set_error(RMEM_ERR_DRIVER, "cuTensorMapEncodeTiled entry point not found");
  uint32_t elem_strides[5] = {1, 1, 1, 1, 1};
  CUresult r = fn(map, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, rank, const_cast<void*>(base), dims,
                  strides_bytes, box, elem_strides, CU_TENSOR_MAP_INTERLEAVE_NONE,
                  CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                  CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) return set_error(RMEM_ERR_DRIVER, "cuTensorMapEncodeTiled failed");
  return RMEM_OK;
}

int lt_attn_partial_launch(const void* q, const void* kbank, const void* vbank,
                           const int32_t* order_dev, int T, int L, int n_slots,
                           int chunks_per_frame, float* part_o, float* part_m, float* part_l,
                           cudaStream_t stream) {
  if (T < 1 || T > RMEM_MAX_BANK || chunks_per_frame < 1 ||
      chunks_per_frame > RMEM_MAX_CHUNKS_PER_FRAME || L < 1 || n_slots < T)
    return set_error(RMEM_ERR_ARG, "lt_attn: bad T / L / chunks_per_frame / n_slots");
  CUtensorMap tm_q, tm_k, tm_v;
  {
    const uint64_t dims[3] = {64, static_cast<uint64_t>(L), 2};
    const uint64_t strides[2] = {kDK * 2, 128};
    const uint32_t box[3] = {64, kBlockM, 2};
    if (int e = encode_map(&tm_q, q, 3, dims, strides, box)) return e;
  }
  {
    const uint64_t dims[4] = {64, static_cast<uint64_t>(L), 2, static_cast<uint64_t>(n_slots)};
    const uint64_t strides[3] = {kDK * 2, 128, static_cast<uint64_t>(L) * kDK * 2};
    const uint32_t box[4] = {64, kBlockN, 2, 1};
    if (int e = encode_map(&tm_k, kbank, 4, dims, strides, box)) return e;
  }
  {
    const uint64_t dims[4] = {64, static_cast<uint64_t>(L), kDV / 64, static_cast<uint64_t>(n_slots)};
    const uint64_t strides[3] = {kDV * 2, 128, static_cast<uint64_t>(L) * kDV * 2};
    const uint32_t box[4] = {64, kBlockN, kDVTile / 64, 1};
    if (int e = encode_map(&tm_v, vbank, 4, dims, strides, box)) return e;
  }
  CUtensorMap tm_o;
  {
    auto fn = get_tensor_map_encoder();
    if (!fn) return set_error(RMEM_ERR_DRIVER, "cuTensorMapEncodeTiled entry point not found");
    const uint64_t dims[3] = {kDV, static_cast<uint64_t>(L), static_cast<uint64_t>(T * chunks_per_frame)};
    const uint64_t strides[2] = {kDV * 4, static_cast<uint64_t>(L) * kDV * 4};
    const uint32_t box[3] = {32, kBlockM, 1};
    const uint32_t es[3] = {1, 1, 1};
    CUresult r = fn(&tm_o, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 3, part_o, dims, strides, box, es,
                    CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B,
                    CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_SUCCESS) return set_error(RMEM_ERR_DRIVER, "cuTensorMapEncodeTiled (partial O) failed");
  }
  static SmemOptIn optin;
  if (int r = ensure_dynamic_smem(reinterpret_cast<const void*>(lt_attn_partial_kernel), &optin, kSmemTotal,
                                  "lt_attn: cudaFuncSetAttribute"))
    return r;
  LtParams p;
  p.L = L;
  p.n_chunks = T * chunks_per_frame;
  p.chunks_per_frame = chunks_per_frame;
  p.order = order_dev;
  p.part_o = part_o;
  p.part_m = part_m;
  p.part_l = part_l;
  p.scale_log2 = 1.4426950408889634f / sqrtf(static_cast<float>(kDK));
  p.direct_gate = nullptr;
  p.direct_out = nullptr;
  p.direct_pair = 0;
  dim3 grid((L + kBlockM - 1) / kBlockM, kDVSplits, p.n_chunks);
  launch_k(lt_attn_partial_kernel, grid, kThreads, kSmemTotal, stream, tm_q, tm_k, tm_v, tm_o, p);
  return check_launch("lt_attn_partial_kernel");
}

// One frame of keys (a layer's self-attention, or a bank holding one frame), complete in one launch:
// out = gate * softmax(q k^T / sqrt(d)) v as bf16.  Grid: (query tiles, 4 value splits).
int lt_attn_direct_launch(const void* q, const void* k, const void* v, int L, const void* gate, void* out,
                          cudaStream_t stream) {
  if (L < 1) return set_error(RMEM_ERR_ARG, "lt_attn_direct: bad L");
  CUtensorMap tm_q, tm_k, tm_v;
  {
    const uint64_t dims[3] = {64, static_cast<uint64_t>(L), 2};
    const uint64_t strides[2] = {kDK * 2, 128};
    const uint32_t box[3] = {64, kBlockM, 2};
    if (int e = encode_map(&tm_q, q, 3, dims, strides, box)) return e;
  }
  {
    const uint64_t dims[4] = {64, static_cast<uint64_t>(L), 2, 1};
    const uint64_t strides[3] = {kDK * 2, 128, static_cast<uint64_t>(L) * kDK * 2};
    const uint32_t box[4] = {64, kBlockN, 2, 1};
    if (int e = encode_map(&tm_k, k, 4, dims, strides, box)) return e;
  }
  {
    const uint64_t dims[4] = {64, static_cast<uint64_t>(L), kDV / 64, 1};
    const uint64_t strides[3] = {kDV * 2, 128, static_cast<uint64_t>(L) * kDV * 2};
    const uint32_t box[4] = {64, kBlockN, kDVTile / 64, 1};
    if (int e = encode_map(&tm_v, v, 4, dims, strides, box)) return e;
  }
  static SmemOptIn optin;
  if (int r = ensure_dynamic_smem(reinterpret_cast<const void*>(lt_attn_partial_kernel), &optin, kSmemTotal,
                                  "lt_attn_direct: cudaFuncSetAttribute"))
    return r;
  LtParams p;
  p.L = L;
  p.n_chunks = 1;
  p.chunks_per_frame = 1;
  p.order = nullptr;                 // one slot: position 0
  p.part_o = nullptr;
  p.part_m = nullptr;
  p.part_l = nullptr;
  p.scale_log2 = 1.4426950408889634f / sqrtf(static_cast<float>(kDK));
  p.direct_gate = static_cast<const __nv_bfloat16*>(gate);
  p.direct_out = static_cast<__nv_bfloat16*>(out);
  // more than four key tiles: split the keys over a 2-CTA cluster (RMEM_LT_SINGLE_PAIR=0 keeps one CTA per tile)
  static const bool pair_on = [] { const char* e = getenv("RMEM_LT_SINGLE_PAIR"); return !(e && e[0] == '0'); }();
  p.direct_pair = (pair_on && L > 4 * kBlockN) ? 1 : 0;
  if (!p.direct_pair) {
    dim3 grid((L + kBlockM - 1) / kBlockM, kDVSplits, 1);
    launch_k(lt_attn_partial_kernel, grid, kThreads, kSmemTotal, stream, tm_q, tm_k, tm_v, tm_q /*unused*/, p);
    return check_launch("lt_attn_partial_kernel (direct)");
  }
  p.n_chunks = 2;
  p.chunks_per_frame = 2;              // the existing chunk logic: keys [0, per) and [per, L), per a multiple of 128
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = dim3((L + kBlockM - 1) / kBlockM, kDVSplits, 2);
  cfg.blockDim = dim3(kThreads);
  cfg.dynamicSmemBytes = kSmemTotal;
  cfg.stream = stream;
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeClusterDimension;
  attr[0].val.clusterDim.x = 1;
  attr[0].val.clusterDim.y = 1;
  attr[0].val.clusterDim.z = 2;
  cfg.attrs = attr;
  cfg.numAttrs = 1;
  cudaLaunchKernelEx(&cfg, lt_attn_partial_kernel, tm_q, tm_k, tm_v, tm_q /*unused*/, p);
  return check_launch("lt_attn_partial_kernel (direct, paired)");
}

int lt_attn_combine_launch(const float* part_o, const float* part_m, const float* part_l,
                           const float* tpe_bias, const int32_t* tpe_rows_host, int T, int L,
                           int chunks_per_frame, const void* gate, void* out, float* attn_mass,
                           cudaStream_t stream) {
  if (T < 1 || T > RMEM_MAX_BANK || chunks_per_frame < 1 ||
      chunks_per_frame > RMEM_MAX_CHUNKS_PER_FRAME)
    return set_error(RMEM_ERR_ARG, "lt_combine: bad T / chunks_per_frame");
  CombineParams p;
  p.L = L; p.T = T; p.chunks_per_frame = chunks_per_frame;
  p.part_o = part_o; p.part_m = part_m; p.part_l = part_l;
  p.tpe_bias = tpe_bias;
  for (int t = 0; t < RMEM_MAX_BANK; ++t) p.tpe_rows[t] = (tpe_rows_host && t < T) ? tpe_rows_host[t] : 0;
  p.gate = static_cast<const __nv_bfloat16*>(gate);
  p.out = static_cast<__nv_bfloat16*>(out);
  p.attn_mass = attn_mass;
  launch_k(lt_attn_combine_kernel, L, 256, 0, stream, p);
  return check_launch("lt_attn_combine_kernel");
}

}  // namespace rmem
