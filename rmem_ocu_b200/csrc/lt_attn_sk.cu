// Long-term gated attention, DeAOT shape (1 head, d_qk = 128, d_v = 1024 = V | ID_V), second generation:
// a PERSISTENT, STREAM-K scheduled kernel on 2-CTA clusters that share their K / V tiles by TMA multicast.
//
// Same operator as lt_attn.cu (GatedPropagation.forward, layers/attention.py:174-193; temporal positional
// embedding transformer.py:1140-1172; attention-weight recording :1185-1192; gate attention.py:206); what
// changes is how the work is laid over the chip:
//
//   * work unit = (query-tile PAIR, value split, 128-key tile).  A pair is two adjacent 128-query tiles, one per
//     CTA of a 2-CTA cluster; a value split is 256 of the 1024 value channels (the TMEM accumulator).  The units of
//     one (pair, split) run over ALL frames of the bank: T * ceil(L/128) key tiles.
//   * the U units are cut into n_clusters equal contiguous ranges (stream-K): every cluster gets U / n_clusters
//     tiles whatever T and L are, so there is no ragged last wave (the first-generation grid of 260 CTAs on 148 SMs
//     ran 1.76 waves), and the prologue (barrier init, TMEM allocation, first TMA round trip) is paid once per SM
//     instead of once per 13 key tiles.
//   * both CTAs of a cluster walk the same key tiles at the same time, so each loads HALF of every K and V tile and
//     multicasts it into both CTAs' shared memory: L2 -> SM traffic per MMA cycle halves (the first generation moved
//     341 MB per launch for 19 MB of operands, 8 TB/s -- at a higher MMA duty cycle that stream alone would exceed
//     what L2 can deliver).  MMAs stay cta_group::1: each CTA owns its queries' S, P and O in its own TMEM.
//   * a range may start and end in the middle of a (pair, split): every maximal run of tiles of one (pair, split)
//     inside a range is a SEGMENT and produces one partial (un-normalised O, running max m, sum l); the combine
//     kernel merges the 2-4 segments of each (pair, split).  The temporal-PE bias is added to the scores here (it is
//     constant per frame and per query), so segments may span frames.
//   * per-frame attention masses W[q, t] (record_attn_weight) come from per-(segment, frame) partial sums written
//     by the value-split-0 CTAs.
//
// Pipeline inside a CTA: warp 0 TMA producer, warp 1 issues S = Q K^T (into a double-buffered TMEM tile), warp 2
// issues O += P V (P read from TMEM) -- two issuers, because one thread's serial stream of barrier waits,
// descriptors, MMAs and commits took ~2900 cycles per tile for 1536 cycles of tensor work -- and 8 softmax warps (two
// threads per query row, lazy rescale, part of the exponentials on the FMA pipes).  The rings, S buffers and
// barriers run continuously across segments; only the O accumulator is drained (TMEM -> registers -> global, column-major so that a warp's
// 32 rows make one 128-byte line) at the end of a segment.
#include <mutex>
#include "sm100_ptx.cuh"
#include "rmem_internal.h"

namespace rmem {
using namespace sm100;

namespace sk {

#ifndef LT_SK_POLY
#define LT_SK_POLY 4
#endif

constexpr int kBlockM = 128;      // queries per CTA
constexpr int kBlockN = 128;      // keys per tile
constexpr int kDK = 128;
constexpr int kDV = 1024;
constexpr int kDVTile = 256;      // value channels per CTA
constexpr int kDVSplits = kDV / kDVTile;
#ifndef LT_SK_ROW_THREADS
#define LT_SK_ROW_THREADS 2
#endif
constexpr int kRowThreads = LT_SK_ROW_THREADS;   // threads per query row: each owns 128 / kRowThreads of a tile's keys
constexpr int kSoftmaxWarps = 4 * kRowThreads;
constexpr int kColsS = kBlockN / kRowThreads;    // S columns (keys) per thread
constexpr int kThreads = 96 + 32 * kSoftmaxWarps;   // producer, QK issuer, PV issuer, 8 softmax warps
constexpr int kMaxSeg = 4;        // segments one CTA may produce (host picks n_clusters so that this holds)

constexpr uint32_t kQBytes = kBlockM * kDK * 2;            // 32 KB  [2][128][64]
constexpr uint32_t kKBytes = kBlockN * kDK * 2;            // 32 KB  [2][128][64], one [128][64] block per CTA of the cluster
constexpr uint32_t kVBytes = kBlockN * kDVTile * 2;        // 64 KB  [4][128][64], two blocks per CTA
constexpr uint32_t kXchBytes = 2 * kRowThreads * kBlockM * 2;   // row-max exchange (bf16), one buffer per S buffer
constexpr uint32_t kSmemData = kQBytes + 2 * kKBytes + 2 * kVBytes;
constexpr uint32_t kSmemTotal = kSmemData + kXchBytes + 256;
static_assert(kRowThreads == 2 || kRowThreads == 4, "two or four threads per query row");
static_assert(kSmemTotal <= 232448, "exceeds the 227 KB shared memory of an sm_100 CTA");

constexpr uint32_t kTmemCols = 512;
constexpr uint32_t kTmemS0 = 0;
constexpr uint32_t kTmemO = 256;
constexpr float kRescaleThreshold = 8.0f;
constexpr int kPolyEvery = LT_SK_POLY;     // every n-th pair of exponentials runs on the FMA pipes (MUFU: 16 / clk / SM);
                                           // 3, 4, 5 and 8 all measure 35.5 us at L = 1620, T = 5: the exponentials are not the limiter

// words per segment slot of the workspace
// O partials are bf16: two adjacent columns per 32-bit word, column-pair-major [128][128] so that a warp's 32 rows
// make one 128-byte line.  (The final output is bf16 too; an un-normalised partial rounded once more adds a
// relative 2^-9 before 2-4 of them are averaged.  fp32 partials doubled the bytes of the segment drain -- which all
// 148 SMs issue at the same moment at the end of the kernel -- and of the combine step.)
constexpr size_t kPartOWords = static_cast<size_t>(kDVTile / 2) * kBlockM;
constexpr int kStatRows = 1 + kRowThreads;                                         // m | l of each of the row's threads
constexpr size_t kPartMLWords = kStatRows * kBlockM;
constexpr size_t kPartFWords = static_cast<size_t>(RMEM_MAX_BANK) * kStatRows * kBlockM;   // the same per frame

struct Params {
  int L, T;
  int S;             // independent sequences batched in this launch (their tensors stacked along a leading dim)
  int n_full;        // full query-tile pairs per sequence (two different 128-query tiles, one per CTA)
  int eps;           // work entries per sequence = 4 * n_full + (odd number of query tiles ? 2 : 0)
  int order_stride;  // int32 words between the `order` arrays of consecutive sequences
  int tpf;           // key tiles per frame = ceil(L / 128)
  int P;             // key tiles per entry = T * tpf
  long long U;       // units = S * eps * P
  int n_clusters;
  int max_seg;       // segment slots per CTA in the workspace (<= kMaxSeg)
  const int32_t* order;
  const float* tpe_bias;           // [L][4] log2 units, or nullptr
  int tpe_rows[RMEM_MAX_BANK];
  float* part_o;     // [n_ctas * max_seg][256][128]
  float* part_ml;    // [n_ctas * max_seg][3][128]
  float* part_f;     // [n_ctas * max_seg][16][3][128] or nullptr (no recording)
  float scale_log2;
};

__device__ __forceinline__ uint32_t cluster_ctarank() {
  uint32_t r;
  asm volatile("mov.u32 %0, %%cluster_ctarank;" : "=r"(r));
  return r;
}
__device__ __forceinline__ void cluster_sync_all() {
  asm volatile("barrier.cluster.arrive.release.aligned;\n\tbarrier.cluster.wait.acquire.aligned;" ::: "memory");
}
// TMA load multicast to the CTAs of `mask`: data and the mbarrier signal land at the same CTA-relative offsets.
__device__ __forceinline__ void tma_load_4d_mc(void* dst, const CUtensorMap* m, uint64_t* bar, int c0, int c1, int c2,
                                               int c3, uint16_t mask) {
  asm volatile(
      "cp.async.bulk.tensor.4d.shared::cluster.global.tile.mbarrier::complete_tx::bytes.multicast::cluster "
      "[%0], [%1, {%3, %4, %5, %6}], [%2], %7;" ::"r"(smem_u32(dst)),
      "l"(reinterpret_cast<uint64_t>(m)), "r"(smem_u32(bar)), "r"(c0), "r"(c1), "r"(c2), "r"(c3), "h"(mask)
      : "memory");
}
__device__ __forceinline__ void tma_load_5d_mc(void* dst, const CUtensorMap* m, uint64_t* bar, int c0, int c1, int c2,
                                               int c3, int c4, uint16_t mask) {
  asm volatile(
      "cp.async.bulk.tensor.5d.shared::cluster.global.tile.mbarrier::complete_tx::bytes.multicast::cluster "
      "[%0], [%1, {%3, %4, %5, %6, %7}], [%2], %8;" ::"r"(smem_u32(dst)),
      "l"(reinterpret_cast<uint64_t>(m)), "r"(smem_u32(bar)), "r"(c0), "r"(c1), "r"(c2), "r"(c3), "r"(c4), "h"(mask)
      : "memory");
}
// tcgen05.commit arriving on the barrier at this CTA-relative offset in every CTA of `mask`.
__device__ __forceinline__ void tc_commit_mc(uint64_t* bar, uint16_t mask) {
  asm volatile(
      "tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.multicast::cluster.b64 [%0], %1;" ::"r"(
          smem_u32(bar)),
      "h"(mask)
      : "memory");
}

#ifdef LT_SK_TRACE
// developer instrumentation (-DLT_SK_TRACE): SM clock stamps of one CTA's three roles, per tile
constexpr int kTraceTiles = 64, kTraceCols = 16;
__device__ long long g_sk_trace[3][kTraceTiles][kTraceCols];
#define SK_TR(role, i, col) do { if (blockIdx.x == LT_SK_TRACE && (i) < kTraceTiles && (threadIdx.x & 31) == 0) g_sk_trace[role][i][col] = clock64(); } while (0)
#else
#define SK_TR(role, i, col) do {} while (0)
#endif

__host__ __device__ __forceinline__ long long range_start(long long c, long long U, int n_clusters) {
  return c * U / n_clusters;
}

// Position of a CTA's i-th tile in the unit space, advanced incrementally (no divisions in the per-tile loops).
struct TileIt {
  int pair, tip, frame_pos, tif;     // (pair, split) index; tile inside it; bank position; tile inside the frame
  __device__ __forceinline__ void init(long long u0, int P, int tpf) {
    pair = static_cast<int>(u0 / P);
    tip = static_cast<int>(u0 - static_cast<long long>(pair) * P);
    frame_pos = tip / tpf;
    tif = tip - frame_pos * tpf;
  }
  __device__ __forceinline__ void next(int P, int tpf) {
    ++tip; ++tif;
    if (tif == tpf) { tif = 0; ++frame_pos; }
    if (tip == P) { tip = 0; frame_pos = 0; ++pair; }
  }
};

// A work ENTRY is what a cluster processes over one run of P key tiles.  For a full pair of query tiles it is
// (pair, value split): CTA r holds query tile 2 * pair + r, both CTAs the same value split, and K and V tiles are shared
// by multicast.  When the number of query tiles is odd (13 at 480p), the last tile would pair with a phantom -- 1/14 of
// all MMA and softmax work spent on padding -- so it gets entries of another kind instead: BOTH CTAs hold that tile and
// CTA r takes value split 2 * k + r (k = 0, 1): K tiles are still shared, each CTA loads its own V tile, S = Q K^T is
// computed twice (as it is across the value splits anyway) and the tile costs 2 P units instead of 4 P.
struct Entry {
  int seq, q_tile, split, qid;   // qid: changes exactly when the CTA's query tile changes
  bool odd;
};
__device__ __forceinline__ Entry decode_entry(int e, int rank, int eps, int n_full) {
  Entry en;
  en.seq = e / eps;
  const int es = e - en.seq * eps;
  en.odd = es >= 4 * n_full;
  if (!en.odd) {
    en.q_tile = 2 * (es >> 2) + rank;
    en.split = es & 3;
    en.qid = en.seq * (n_full + 1) + (es >> 2);
  } else {
    en.q_tile = 2 * n_full;
    en.split = 2 * (es - 4 * n_full) + rank;
    en.qid = en.seq * (n_full + 1) + n_full;
  }
  return en;
}

__global__ void __cluster_dims__(2, 1, 1) __launch_bounds__(kThreads, 1)
lt_attn_sk_kernel(const __grid_constant__ CUtensorMap tm_q, const __grid_constant__ CUtensorMap tm_k,
                  const __grid_constant__ CUtensorMap tm_v, const Params p) {
  extern __shared__ __align__(1024) uint8_t smem[];
  uint8_t* sQ = smem;
  uint8_t* sK = sQ + kQBytes;
  uint8_t* sV = sK + 2 * kKBytes;
  __nv_bfloat16* xch = reinterpret_cast<__nv_bfloat16*>(sV + 2 * kVBytes);   // [2][4][128]
  uint64_t* bars = reinterpret_cast<uint64_t*>(sV + 2 * kVBytes + kXchBytes);
  uint64_t* q_full = bars + 0;
  uint64_t* k_full = bars + 1;             // [2]
  uint64_t* k_empty = bars + 3;            // [2]  count 2: one commit from each CTA of the cluster
  uint64_t* v_full = bars + 5;             // [2]
  uint64_t* v_empty = bars + 7;            // [2]  count 2
  uint64_t* s_full = bars + 9;             // [2]
  uint64_t* p_full = bars + 11;            // [2]
  uint64_t* pv_done = bars + 13;           // [2]
  uint64_t* o_full = bars + 15;
  uint64_t* o_drained = bars + 16;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + 17);

  // warp-uniform by construction (a shuffle from lane 0): the issuing warps below keep their operands in uniform
  // registers, so each tcgen05.mma / commit is ONE instruction instead of an elect + broadcast loop
  const int warp = __shfl_sync(0xffffffffu, static_cast<int>(threadIdx.x >> 5), 0);
  const int lane = threadIdx.x & 31;
  const uint32_t rank = cluster_ctarank();
  const int cluster = blockIdx.x >> 1;
  const long long u0 = range_start(cluster, p.U, p.n_clusters);
  const int n = static_cast<int>(range_start(cluster + 1, p.U, p.n_clusters) - u0);   // tiles of this CTA

  if (warp == 0 && lane == 0) {
    if (smem_u32(smem) & 1023u) asm volatile("trap;");
    tma_prefetch_desc(&tm_q);
    tma_prefetch_desc(&tm_k);
    tma_prefetch_desc(&tm_v);
    mbar_init(q_full, 1);
    for (int i = 0; i < 2; ++i) {
      mbar_init(k_full + i, 1);
      mbar_init(k_empty + i, 2);
      mbar_init(v_full + i, 1);
      mbar_init(v_empty + i, 2);
      mbar_init(s_full + i, 1);
      mbar_init(p_full + i, 32 * kSoftmaxWarps);
      mbar_init(pv_done + i, 1);
    }
    mbar_init(o_full, 1);
    mbar_init(o_drained, 32 * kSoftmaxWarps);
    mbar_fence_init();
  }
  if (warp == 1) tmem_alloc<kTmemCols>(tmem_slot);
  tc_fence_before();
  __syncthreads();
  cluster_sync_all();               // the peer's barriers exist before anything multicasts into them
  tc_fence_after();
  const uint32_t tmem_base = __shfl_sync(0xffffffffu, *tmem_slot, 0);
  pdl_wait();        // programmatic dependent launch: see rmem_internal.h
  pdl_trigger();

  if (warp == 0) {
    // ------------------------------------------------------------------ TMA producer
    if (lane == 0) {
      TileIt it;
      it.init(u0, p.P, p.tpf);
      int q_pair_loaded = -1;
      int frame_cached = -1, slot = 0;
      for (int i = 0; i < n; ++i) {
        const Entry en = decode_entry(it.pair, static_cast<int>(rank), p.eps, p.n_full);
        const int gq = en.qid, dv_split = en.split, seq = en.seq;
        const int st = i & 1;
        if (gq != q_pair_loaded) {
          // the queries change: the last QK^T that read the old ones has completed once tile i-1 released its K
          if (i > 0) mbar_wait(k_empty + ((i - 1) & 1), ((i - 1) >> 1) & 1);
          mbar_expect_tx(q_full, kQBytes);
          tma_load_4d(sQ, &tm_q, q_full, 0, en.q_tile * kBlockM, 0, seq);
          q_pair_loaded = gq;
          frame_cached = -1;
        }
        if (it.frame_pos != frame_cached) {
          slot = p.order[static_cast<size_t>(seq) * p.order_stride + it.frame_pos];
          frame_cached = it.frame_pos;
        }
        const uint32_t ph = ((i >> 1) - 1) & 1;
        SK_TR(0, i, 0);
        if (i >= 2) mbar_wait(k_empty + st, ph);
        SK_TR(0, i, 1);
        mbar_expect_tx(k_full + st, kKBytes);
        tma_load_5d_mc(sK + st * kKBytes + rank * (kKBytes / 2), &tm_k, k_full + st, 0, it.tif * kBlockN,
                       static_cast<int>(rank), slot, seq, 0x3);
        if (i >= 2) mbar_wait(v_empty + st, ph);
        SK_TR(0, i, 2);
        mbar_expect_tx(v_full + st, kVBytes);
        if (!en.odd) {
          tma_load_5d_mc(sV + st * kVBytes + rank * (kVBytes / 2), &tm_v, v_full + st, 0, it.tif * kBlockN,
                         dv_split * 4 + 2 * static_cast<int>(rank), slot, seq, 0x3);
        } else {         // the CTAs hold different value splits: each fetches both halves of its own V tile
          tma_load_5d(sV + st * kVBytes, &tm_v, v_full + st, 0, it.tif * kBlockN, dv_split * 4, slot, seq);
          tma_load_5d(sV + st * kVBytes + kVBytes / 2, &tm_v, v_full + st, 0, it.tif * kBlockN, dv_split * 4 + 2, slot,
                      seq);
        }
        it.next(p.P, p.tpf);
      }
    }
  } else if (warp == 1) {
    // ------------------------------------------------------------------ MMA issuer 1: S = Q K^T
    // (its own warp: one thread's serial issue stream -- barrier waits, descriptors, 16 MMAs and 4 commits per
    // tile -- was the bottleneck of the first generation: ~2900 cycles per tile for 1536 cycles of tensor work)
    // The whole warp runs the loop (waits included); one elected lane issues.  Inside an `if (lane == 0)` region the
    // compiler must assume per-thread operands and wraps every tcgen05.mma in an ELECT / R2UR.BROADCAST loop
    // (~90 cycles per MMA, measured); with warp-uniform control flow the descriptors live in uniform registers.
    if (n > 0) {
      constexpr uint32_t idesc_qk = make_idesc_bf16(kBlockM, kBlockN, 0, 0);
      const uint64_t q_desc = make_desc_sw128(smem_u32(sQ), 16, 1024);
      const uint64_t k_desc0 = make_desc_sw128(smem_u32(sK), 16, 1024);
      TileIt it;
      it.init(u0, p.P, p.tpf);
      int q_pair_seen = -1, q_loads = 0;
      for (int j = 0; j < n; ++j) {
        const int qid = decode_entry(it.pair, static_cast<int>(rank), p.eps, p.n_full).qid;
        if (qid != q_pair_seen) {
          mbar_wait(q_full, q_loads & 1);
          ++q_loads;
          q_pair_seen = qid;
        }
        const int st = j & 1;
        // S buffer st last held P(j-2): its aggregation must have completed
        if (j >= 2) mbar_wait(pv_done + st, ((j - 2) >> 1) & 1);
        SK_TR(1, j, 0);
        mbar_wait(k_full + st, (j >> 1) & 1);
        SK_TR(1, j, 1);
        tc_fence_after();
        const uint64_t k_desc = k_desc0 + static_cast<uint64_t>((st * kKBytes) >> 4);
        const uint32_t d = tmem_base + kTmemS0 + st * kBlockN;
        if (elect_one()) {
#pragma unroll
          for (int ks = 0; ks < kDK / 16; ++ks) {
            // 64-channel block (ks >> 2) of the [2][128][64] tile, 32 bytes per K step inside it
            const uint64_t off = static_cast<uint64_t>(((ks >> 2) * (kBlockM * 128) + (ks & 3) * 32) >> 4);
            mma_bf16_ss(d, q_desc + off, k_desc + off, idesc_qk, ks > 0);
          }
          tc_commit_mc(k_empty + st, 0x3);          // both CTAs' producers write this slot
          tc_commit(s_full + st);
        }
        __syncwarp();
        SK_TR(1, j, 2);
        it.next(p.P, p.tpf);
      }
    }
  } else if (warp == 2) {
    // ------------------------------------------------------------------ MMA issuer 2: O += P V
    if (n > 0) {
      constexpr uint32_t idesc_pv = make_idesc_bf16(kBlockM, kDVTile, 0, 1);
      const uint64_t v_desc0 = make_desc_sw128(smem_u32(sV), kBlockN * 128, 1024);
      TileIt it;
      it.init(u0, p.P, p.tpf);
      int segs_done = 0;
      for (int i = 0; i < n; ++i) {
        const bool first = (i == 0) || (it.tip == 0);
        const bool last = (i == n - 1) || (it.tip == p.P - 1);
        const int st = i & 1;
        SK_TR(1, i, 3);
        mbar_wait(v_full + st, (i >> 1) & 1);
        if (first && segs_done > 0) mbar_wait(o_drained, (segs_done - 1) & 1);   // the accumulator has been read out
        SK_TR(1, i, 4);
        mbar_wait(p_full + st, (i >> 1) & 1);
        SK_TR(1, i, 5);
        tc_fence_after();
        const uint64_t v_desc = v_desc0 + static_cast<uint64_t>((st * kVBytes) >> 4);
        const uint32_t p_tmem = tmem_base + kTmemS0 + st * kBlockN;   // bf16 P: 64 columns
        const uint32_t d = tmem_base + kTmemO;
        if (elect_one()) {
#pragma unroll
          for (int ks = 0; ks < kBlockN / 16; ++ks)
            mma_bf16_ts(d, p_tmem + ks * 8, v_desc + static_cast<uint64_t>((ks * 16 * 128) >> 4), idesc_pv,
                        (!first) || (ks > 0));
          tc_commit_mc(v_empty + st, 0x3);
          tc_commit(pv_done + st);
          if (last) tc_commit(o_full);
        }
        __syncwarp();
        if (last) ++segs_done;
        SK_TR(1, i, 6);
        it.next(p.P, p.tpf);
      }
    }
  } else {
    // ------------------------------------------------------------------ softmax + segment epilogue
    // 16 warps: warp (quarter, part) owns TMEM lanes [32 quarter, +32) = 32 query rows and keys [32 part, +32) of
    // every 128-key tile; the four threads of a row agree on the running max through shared memory.
    const int quarter = warp & 3;                 // TMEM lane quarter this warp may access
    const int part = (warp - 3) >> 2;             // which 32 of the tile's 128 columns this thread owns
    const int row = quarter * 32 + lane;
    const uint32_t lane_addr = static_cast<uint32_t>(quarter * 32) << 16;
    const uint32_t row_bar = 1 + quarter;         // named barrier of the 4 warps that share these 32 rows
    const float c = p.scale_log2;
    const bool tracer = (warp == 3 && lane == 0);
    float m_used = -INFINITY, l_sum = 0.f, l_frame = 0.f, bias = 0.f;
    float4 bias4 = make_float4(0.f, 0.f, 0.f, 0.f);
    int q_global = 0, prev_frame = -1, segs_done = 0;
    size_t slot = 0;
    bool rec = false;
    TileIt it;
    it.init(u0, p.P, p.tpf);
    const int pair0 = it.pair;
    (void)tracer;

    for (int i = 0; i < n; ++i) {
      const int frame_pos = it.frame_pos;
      const bool first = (i == 0) || (it.tip == 0);
      const bool last = (i == n - 1) || (it.tip == p.P - 1);
      if (first || frame_pos != prev_frame) {
        if (first) {
          m_used = -INFINITY; l_sum = 0.f;
          const Entry en = decode_entry(it.pair, static_cast<int>(rank), p.eps, p.n_full);
          const int seq = en.seq;
          q_global = en.q_tile * kBlockM + row;                                                // row inside its sequence
          slot = static_cast<size_t>(blockIdx.x) * p.max_seg + (it.pair - pair0);
          rec = p.part_f != nullptr && en.split == 0;
          bias4 = make_float4(0.f, 0.f, 0.f, 0.f);
          if (p.tpe_bias && q_global < p.L)
            bias4 = *reinterpret_cast<const float4*>(p.tpe_bias + (static_cast<size_t>(seq) * p.L + q_global) * 4);
        } else if (rec) {     // close the previous frame's partial mass (relative to the running max at this point)
          float* f = p.part_f + (slot * RMEM_MAX_BANK + prev_frame) * (kStatRows * kBlockM);
          if (part == 0) f[row] = m_used;
          f[(1 + part) * kBlockM + row] = l_frame;
        }
        l_frame = 0.f;
        prev_frame = frame_pos;
        const int trow = p.tpe_rows[frame_pos];
        bias = trow == 0 ? bias4.x : trow == 1 ? bias4.y : trow == 2 ? bias4.z : bias4.w;
      }

      const int b = i & 1;
      if (tracer) SK_TR(2, i, 0);
      mbar_wait(s_full + b, (i >> 1) & 1);
      if (tracer) SK_TR(2, i, 1);
      tc_fence_after();
      uint32_t sr[kColsS];
#pragma unroll
      for (int cb = 0; cb < kColsS / 32; ++cb) {
        uint32_t(&blk)[32] = *reinterpret_cast<uint32_t(*)[32]>(&sr[cb * 32]);
        tmem_ld32(tmem_base + lane_addr + kTmemS0 + b * kBlockN + part * kColsS + cb * 32, blk);
      }
      tmem_ld_wait();
      const int valid = p.L - it.tif * kBlockN - part * kColsS;   // valid columns of this thread's share (>= kColsS: all)
      if (valid < kColsS) {                                     // warp-uniform: only a frame's last tile
#pragma unroll
        for (int k = 0; k < kColsS; ++k)
          if (k >= valid) sr[k] = 0xff800000u;                  // -inf
      }
      float mx = -INFINITY;
#pragma unroll
      for (int k = 0; k < kColsS; k += 2)
        mx = fmaxf(mx, fmaxf(__uint_as_float(sr[k]), __uint_as_float(sr[k + 1])));
      if (tracer) SK_TR(2, i, 2);
      // the row's four threads exchange their maxima as bf16: every thread reads all four ROUNDED values (its own
      // included), so all four agree exactly on the reference point; it may sit a rounding step below the true
      // maximum, which only makes some probabilities marginally larger than 1
      __nv_bfloat16* xb = xch + (b * kRowThreads) * kBlockM + row;
      xb[part * kBlockM] = __float2bfloat16_rn(mx);
      named_bar_sync(row_bar, 32 * kRowThreads);
      if (tracer) SK_TR(2, i, 3);
      mx = fmaxf(__bfloat162float(xb[0]), __bfloat162float(xb[kBlockM]));
      if (kRowThreads == 4)
        mx = fmaxf(mx, fmaxf(__bfloat162float(xb[2 * kBlockM]), __bfloat162float(xb[3 * kBlockM])));
      mx = mx * c + bias;
      const bool grow = mx > m_used + kRescaleThreshold;
      float scale = 1.f;
      float m_new = m_used;
      if (grow) {
        m_new = mx;
        scale = fast_exp2(m_used - m_new);            // 0 when m_used == -inf
      }
      if (!first && __any_sync(0xffffffffu, grow)) {   // same decision in the four warps of these rows
        mbar_wait(pv_done + ((i - 1) & 1), ((i - 1) >> 1) & 1);
        tc_fence_after();
        const uint32_t o_addr = tmem_base + lane_addr + kTmemO + part * (kDVTile / kRowThreads);
#pragma unroll 1
        for (int cb = 0; cb < kDVTile / kRowThreads / 32; ++cb) {
          uint32_t orow[32];
          tmem_ld32(o_addr + cb * 32, orow);
          tmem_ld_wait();
#pragma unroll
          for (int k = 0; k < 32; ++k) orow[k] = __float_as_uint(__uint_as_float(orow[k]) * scale);
          tmem_st32(o_addr + cb * 32, orow);
        }
      }
      l_sum *= scale;
      l_frame *= scale;
      m_used = m_new;
      const float off = bias - m_used;
      float acc0 = 0.f, acc1 = 0.f;
      uint32_t packed[kColsS / 2];
#pragma unroll
      for (int k = 0; k < kColsS; k += 2) {
        float x0, x1;
        fma_f32x2(x0, x1, __uint_as_float(sr[k]), __uint_as_float(sr[k + 1]), c, c, off, off);
        float p0, p1;
        if ((k >> 1) % kPolyEvery == kPolyEvery - 1) {
          exp2_poly_x2(p0, p1, x0, x1);             // FMA / ALU pipes
        } else {
          p0 = fast_exp2(x0);                       // MUFU.EX2
          p1 = fast_exp2(x1);
        }
        add_f32x2(acc0, acc1, acc0, acc1, p0, p1);
        packed[k >> 1] = pack_bf16x2(p0, p1);
      }
      l_sum += acc0 + acc1;
      l_frame += acc0 + acc1;
      if (tracer) SK_TR(2, i, 4);
      // P (bf16, 64 columns) overwrites the first half of this S buffer: every thread of the row has its S values
      // in registers (they all passed tcgen05.wait::ld before the row barrier)
      if constexpr (kColsS == 32) {
        tmem_st16(tmem_base + lane_addr + kTmemS0 + b * kBlockN + part * 16,
                  *reinterpret_cast<const uint32_t(*)[16]>(&packed[0]));
      } else {
        tmem_st32(tmem_base + lane_addr + kTmemS0 + b * kBlockN + part * 32,
                  *reinterpret_cast<const uint32_t(*)[32]>(&packed[0]));
      }
      tmem_st_wait();
      tc_fence_before();
      mbar_arrive(p_full + b);
      if (tracer) SK_TR(2, i, 5);

      if (last) {
        // ---- segment epilogue: O (fp32, un-normalised) -> part_o, column-major; m, l -> part_ml
        mbar_wait(o_full, segs_done & 1);
        tc_fence_after();
        const bool row_ok = q_global < p.L;
        constexpr int kCols = kDVTile / kRowThreads;          // accumulator columns per thread (64 or 128)
        uint32_t* po = reinterpret_cast<uint32_t*>(p.part_o) + slot * kPartOWords +
                       static_cast<size_t>(part * (kCols / 2)) * kBlockM + row;
        const uint32_t o_addr = tmem_base + lane_addr + kTmemO + part * kCols;
#pragma unroll 1
        for (int cb = 0; cb < kCols / 64; ++cb) {
          uint32_t o0[32], o1[32];
          tmem_ld32(o_addr + cb * 64, o0);
          tmem_ld32(o_addr + cb * 64 + 32, o1);
          tmem_ld_wait();
          if (cb == kCols / 64 - 1) {              // everything is in registers: the MMA thread may overwrite O
            tc_fence_before();
            mbar_arrive(o_drained);
          }
#ifndef LT_SK_EXP_NOSTORE
          if (row_ok) {
#pragma unroll
            for (int k = 0; k < 16; ++k)
              po[(cb * 32 + k) * kBlockM] = pack_bf16x2(__uint_as_float(o0[2 * k]), __uint_as_float(o0[2 * k + 1]));
#pragma unroll
            for (int k = 0; k < 16; ++k)
              po[(cb * 32 + 16 + k) * kBlockM] = pack_bf16x2(__uint_as_float(o1[2 * k]), __uint_as_float(o1[2 * k + 1]));
          }
#endif
        }
        if (row_ok) {
          float* ml = p.part_ml + slot * kPartMLWords;
          if (part == 0) ml[row] = m_used;
          ml[(1 + part) * kBlockM + row] = l_sum;
          if (rec) {
            float* f = p.part_f + (slot * RMEM_MAX_BANK + frame_pos) * (kStatRows * kBlockM);
            if (part == 0) f[row] = m_used;
            f[(1 + part) * kBlockM + row] = l_frame;
          }
        }
        ++segs_done;
        if (tracer) SK_TR(2, i, 6);
      }
      it.next(p.P, p.tpf);
    }
    tc_fence_before();
  }

  __syncthreads();
  cluster_sync_all();               // the peer may still multicast into / arrive on this CTA until it is done too
  if (warp == 1) {
    tc_fence_after();
    tmem_dealloc<kTmemCols>(tmem_base);
  }
}

// ---------------------------------------------------------------------------------------------
// combine: for every (query tile, 64-column block) merge the segments of its (pair, split):
//   out[q, c] = gate[q, c] * sum_s w_s O_s[q, c] / sum_s w_s l_s,   w_s = 2^(m_s - M)
//   W[q, t]   = sum_{s covering frame t} 2^(mf_{s,t} - M) lf_{s,t} / sum_s w_s l_s        (split 0 only)
struct CombineParams {
  int L, T, tpf, P;
  int n_q_tiles, n_full, eps;  // per sequence: query tiles, full pairs, work entries (see Entry)
  long long U;
  int n_clusters, max_seg;
  const float* part_o;
  const float* part_ml;
  const float* part_f;
  const __nv_bfloat16* gate;   // [L][1024] or nullptr
  __nv_bfloat16* out;          // [L][1024]
  float* attn_mass;            // [L][T] or nullptr
};

constexpr int kCbCols = 64;     // columns per combine CTA
static_assert(RMEM_MAX_BANK <= kCbCols + 1, "the mass CTAs reuse the staging tile as [128][T]");
constexpr int kMaxSegPerPair = 16;

__global__ void __launch_bounds__(256)
lt_attn_sk_combine_kernel(const CombineParams p) {
  pdl_wait();        // programmatic dependent launch: see rmem_internal.h
  pdl_trigger();
  __shared__ float tile[kBlockM][kCbCols + 1];
  __shared__ int seg_slot[kMaxSegPerPair], seg_a[kMaxSegPerPair], seg_b[kMaxSegPerPair];
  __shared__ int n_seg_sh;
  const int seq = blockIdx.x / p.n_q_tiles;                // blockIdx.x = (sequence, query tile)
  const int q_tile = blockIdx.x - seq * p.n_q_tiles;
  const bool mass_cta = blockIdx.y == kDV / kCbCols;       // the extra row of CTAs computes W[q, t] only
  const int dv_split = mass_cta ? 0 : blockIdx.y >> 2;
  const int col0 = (blockIdx.y & 3) * kCbCols;            // inside the split's 256 columns
  // the entry that produced (q_tile, dv_split) and the CTA of the cluster that held it (see Entry)
  const bool odd_tile = q_tile >= 2 * p.n_full;
  const int pair = seq * p.eps + (odd_tile ? 4 * p.n_full + (dv_split >> 1) : (q_tile >> 1) * 4 + dv_split);
  const int rank = odd_tile ? (dv_split & 1) : (q_tile & 1);
  const size_t seq_row0 = static_cast<size_t>(seq) * p.L;  // first row of this sequence in gate / out / attn_mass
  // the gate does not depend on anything computed here: fetch it first (two threads per row, 64 bytes each)
  const int rr = threadIdx.x >> 1, hh = threadIdx.x & 1;
  const int qq = q_tile * kBlockM + rr;
  const size_t out_off = (seq_row0 + qq) * kDV + dv_split * kDVTile + col0 + hh * 32;
  uint4 g[4];
  if (!mass_cta && p.gate && qq < p.L) {
#pragma unroll
    for (int k = 0; k < 4; ++k) g[k] = *reinterpret_cast<const uint4*>(p.gate + out_off + k * 8);
  }
  if (threadIdx.x < 32) {
    // the segments of this (pair, split): clusters whose ranges intersect [pair * P, (pair + 1) * P).  Lane k looks at
    // cluster c0 + k (c0 = the cluster whose range contains the first tile); the hits are compacted with a ballot.
    // 32-bit arithmetic: the launcher checks U * n_clusters < 2^31.
    const int U = static_cast<int>(p.U), nc = p.n_clusters;
    const int lo = pair * p.P, hi = lo + p.P;
    int c0 = static_cast<int>((static_cast<unsigned>(lo) * static_cast<unsigned>(nc)) / static_cast<unsigned>(U));
    while (c0 + 1 < nc && ((c0 + 1) * U) / nc <= lo) ++c0;
    while (c0 > 0 && (c0 * U) / nc > lo) --c0;
    const int c = c0 + static_cast<int>(threadIdx.x);
    bool hit = false;
    int a = 0, b = 0, sl = 0;
    if (c < nc) {      // an entry spans at most n_clusters / n_entries + 2 <= 32 clusters (checked by the launcher)
      const int s = (c * U) / nc, e = ((c + 1) * U) / nc;
      a = s > lo ? s : lo;
      b = e < hi ? e : hi;
      hit = b > a;
      sl = (2 * c + rank) * p.max_seg + (pair - s / p.P);
    }
    const unsigned m = __ballot_sync(0xffffffffu, hit);
    if (hit) {
      const int k = __popc(m & ((1u << threadIdx.x) - 1u));
      if (k < kMaxSegPerPair) { seg_slot[k] = sl; seg_a[k] = a - lo; seg_b[k] = b - lo; }
    }
    if (threadIdx.x == 0) n_seg_sh = min(__popc(m), kMaxSegPerPair);
  }
  if (mass_cta) {
    float* mass_sh = &tile[0][0];                          // [128][T] partial masses
    for (int i = threadIdx.x; i < kBlockM * p.T; i += 256) mass_sh[i] = 0.f;
  }
  __syncthreads();
  const int ns = n_seg_sh;
  const int r = threadIdx.x & (kBlockM - 1), h = threadIdx.x >> 7;       // row of the tile, column half
  const int q = q_tile * kBlockM + r;
  float M = -INFINITY, denom = 0.f, inv = 0.f;
  float w[kMaxSegPerPair];
  if (q < p.L) {
    float mm[kMaxSegPerPair], ll[kMaxSegPerPair];
#pragma unroll 1
    for (int s0 = 0; s0 < ns; s0 += 4) {
      float t[4][kStatRows];
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        const float* ml = p.part_ml + static_cast<size_t>(seg_slot[min(s0 + j, ns - 1)]) * kPartMLWords;
#pragma unroll
        for (int k = 0; k < kStatRows; ++k) t[j][k] = ml[k * kBlockM + r];
      }
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        if (s0 + j < ns) {
          mm[s0 + j] = t[j][0];
          float lsum = 0.f;
#pragma unroll
          for (int k = 1; k < kStatRows; ++k) lsum += t[j][k];
          ll[s0 + j] = lsum;
          M = fmaxf(M, t[j][0]);
        }
      }
    }
#pragma unroll 1
    for (int s = 0; s < ns; ++s) {
      w[s] = exp2f(mm[s] - M);
      denom += w[s] * ll[s];
    }
    inv = 1.f / denom;
  }
  if (mass_cta) {
    // W[q, t] = sum over the (segment, frame) pieces of frame t; the row's two threads take alternate segments
    float* mass_sh = &tile[0][0];
    if (q < p.L) {
      for (int s = h; s < ns; s += 2) {
        const int t0 = seg_a[s] / p.tpf, t1 = (seg_b[s] - 1) / p.tpf;
        for (int t = t0; t <= t1; ++t) {
          const float* f = p.part_f + (static_cast<size_t>(seg_slot[s]) * RMEM_MAX_BANK + t) * (kStatRows * kBlockM);
          float lsum = 0.f;
#pragma unroll
          for (int k = 1; k < kStatRows; ++k) lsum += f[k * kBlockM + r];
          const float v = exp2f(f[r] - M) * lsum;
          atomicAdd(mass_sh + r * p.T + t, v * inv);
        }
      }
    }
    __syncthreads();
    // [128][T] is contiguous in attn_mass as well (row-major [L][T]): coalesced copy
    const int n_valid = min(kBlockM, p.L - q_tile * kBlockM) * p.T;
    for (int i = threadIdx.x; i < n_valid; i += 256)
      p.attn_mass[(seq_row0 + static_cast<size_t>(q_tile) * kBlockM) * p.T + i] = mass_sh[i];
    return;
  }
  if (q < p.L) {
    float acc[kCbCols / 2];
#pragma unroll
    for (int k = 0; k < kCbCols / 2; ++k) acc[k] = 0.f;
    // four segments per round, all their loads issued before the first use: one memory round trip per round
    // instead of one per segment (a (pair, split) has 2-4 segments on this path's shapes)
#pragma unroll 1
    for (int s0 = 0; s0 < ns; s0 += 4) {
      uint32_t v[4][kCbCols / 4];
      float ws[4];
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        const int s = min(s0 + j, ns - 1);
        ws[j] = s0 + j < ns ? w[s] * inv : 0.f;
        const uint32_t* po = reinterpret_cast<const uint32_t*>(p.part_o) + static_cast<size_t>(seg_slot[s]) * kPartOWords +
                             static_cast<size_t>((col0 + h * (kCbCols / 2)) / 2) * kBlockM + r;
#pragma unroll
        for (int k = 0; k < kCbCols / 4; ++k) v[j][k] = po[static_cast<size_t>(k) * kBlockM];
      }
#pragma unroll
      for (int j = 0; j < 4; ++j) {
#pragma unroll
        for (int k = 0; k < kCbCols / 4; ++k) {
          acc[2 * k] += ws[j] * __uint_as_float(v[j][k] << 16);
          acc[2 * k + 1] += ws[j] * __uint_as_float(v[j][k] & 0xffff0000u);
        }
      }
    }
#pragma unroll
    for (int k = 0; k < kCbCols / 2; ++k) tile[r][h * (kCbCols / 2) + k] = acc[k];
  }
  __syncthreads();
  // rows of the staged tile -> gate -> bf16, two threads per row (64 bytes each, 128 contiguous bytes per row)
  if (qq < p.L) {
    uint4 o[4];
#pragma unroll
    for (int k = 0; k < 4; ++k) {
      uint32_t* ow = reinterpret_cast<uint32_t*>(&o[k]);
      const __nv_bfloat162* gw = reinterpret_cast<const __nv_bfloat162*>(&g[k]);
#pragma unroll
      for (int e = 0; e < 4; ++e) {
        float a = tile[rr][hh * 32 + k * 8 + 2 * e], b = tile[rr][hh * 32 + k * 8 + 2 * e + 1];
        if (p.gate) { a *= __bfloat162float(gw[e].x); b *= __bfloat162float(gw[e].y); }
        ow[e] = pack_bf16x2(a, b);
      }
      *reinterpret_cast<uint4*>(p.out + out_off + k * 8) = o[k];
    }
  }
}

// ---------------------------------------------------------------------------------------------
struct Plan {
  int tpf, P, n_full, eps, n_entries, n_clusters, max_seg, n_ctas;
  long long U;
};

static int sm_count() {
  static std::mutex mu;
  static int cached[kMaxDevices];
  int dev = 0;
  if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= kMaxDevices) return 148;
  std::lock_guard<std::mutex> lock(mu);
  if (!cached[dev]) {
    int n = 0;
    if (cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess || n < 2) n = 148;
    cached[dev] = n;
  }
  return cached[dev];
}

static Plan make_plan(int L, int T, int S, int n_sms, int n_clusters_req) {
  Plan pl;
  pl.tpf = (L + kBlockN - 1) / kBlockN;
  pl.P = T * pl.tpf;
  const int q_tiles = (L + kBlockM - 1) / kBlockM;
  pl.n_full = q_tiles / 2;
  pl.eps = kDVSplits * pl.n_full + (q_tiles & 1 ? kDVSplits / 2 : 0);
  pl.n_entries = S * pl.eps;
  pl.U = static_cast<long long>(pl.n_entries) * pl.P;
  int nc = n_sms / 2;                       // one cluster per SM pair: everything is resident at once
  if (n_clusters_req > 0 && n_clusters_req < nc) nc = n_clusters_req;   // leave SMs to kernels on other streams
  // tiny problems: no more clusters than units, and few enough that an entry's segments fit the combine step's scan
  // window (an entry spans at most nc / n_entries + 2 clusters; the window is 32)
  if (nc > pl.U) nc = static_cast<int>(pl.U);
  if (nc > 30 * pl.n_entries) nc = 30 * pl.n_entries;
  if (nc < 1) nc = 1;
  // a range of R tiles touches at most (R + P - 2) / P + 1 (pair, split)s: keep that within kMaxSeg
  for (;;) {
    const long long R = (pl.U + nc - 1) / nc;
    const long long touched = R <= 1 ? 1 : (R + pl.P - 2) / pl.P + 1;
    if (touched <= kMaxSeg) { pl.max_seg = static_cast<int>(touched); break; }
    nc *= 2;
  }
  pl.n_clusters = nc;
  pl.n_ctas = 2 * nc;
  return pl;
}

static size_t slot_words(bool) { return kPartOWords + kPartMLWords + kPartFWords; }

}  // namespace sk

#ifdef LT_SK_TRACE
extern "C" int rmem_dev_sk_trace(long long* out_host) {
  return cudaMemcpyFromSymbol(out_host, sk::g_sk_trace, sizeof(sk::g_sk_trace)) == cudaSuccess ? 0 : -2;
}
#endif

int64_t lt_attn_sk_workspace_bytes(int L, int T, int S, int n_clusters) {
  if (L < 1 || T < 1 || T > RMEM_MAX_BANK || S < 1) return -1;
  const sk::Plan pl = sk::make_plan(L, T, S, sk::sm_count(), n_clusters);
  return static_cast<int64_t>(pl.n_ctas) * pl.max_seg * sk::slot_words(true) * 4;
}

static int sk_encode(CUtensorMap* map, const void* base, int rank, const uint64_t* dims, const uint64_t* strides_bytes,
                     const uint32_t* box) {
  auto fn = get_tensor_map_encoder();
  if (!fn) return set_error(RMEM_ERR_DRIVER, "cuTensorMapEncodeTiled entry point not found");
  uint32_t elem_strides[5] = {1, 1, 1, 1, 1};
  CUresult r = fn(map, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, rank, const_cast<void*>(base), dims, strides_bytes, box,
                  elem_strides, CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B,
                  CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) return set_error(RMEM_ERR_DRIVER, "cuTensorMapEncodeTiled (lt_attn_sk) failed");
  return RMEM_OK;
}

static void sk_split_workspace(const sk::Plan& pl, void* workspace, float** part_o, float** part_ml, float** part_f) {
  const size_t n_slots = static_cast<size_t>(pl.n_ctas) * pl.max_seg;
  float* w = static_cast<float*>(workspace);
  *part_o = w;
  *part_ml = w + n_slots * sk::kPartOWords;
  *part_f = *part_ml + n_slots * sk::kPartMLWords;
}

int lt_attn_sk_partial_launch(const void* q, const void* kbank, const void* vbank, const int32_t* order_dev,
                              int order_stride, int T, int L, int n_slots, int S, const float* tpe_bias,
                              const int32_t* tpe_rows_host, int record, int n_clusters, void* workspace,
                              int64_t workspace_bytes, cudaStream_t stream) {
  using namespace sk;
  if (T < 1 || T > RMEM_MAX_BANK || L < 1 || n_slots < T || S < 1 || (S > 1 && order_stride < T))
    return set_error(RMEM_ERR_ARG, "lt_attn_sk: bad T / L / n_slots / batch");
  const Plan pl = make_plan(L, T, S, sm_count(), n_clusters);
  if (!workspace || workspace_bytes < static_cast<int64_t>(pl.n_ctas) * pl.max_seg * slot_words(true) * 4)
    return set_error(RMEM_ERR_ARG, "lt_attn_sk: workspace smaller than rmem_lt_attn_deaot_sk_workspace_bytes(L, T)");
  if (pl.U * (pl.n_clusters + 1) >= (1ll << 31))
    return set_error(RMEM_ERR_ARG, "lt_attn_sk: T * L too large for the 32-bit tile arithmetic of the combine step");
  if (pl.n_clusters / pl.n_entries + 2 > 32 || T > kMaxSegPerPair)
    return set_error(RMEM_ERR_ARG, "lt_attn_sk: a (pair, split) would span more clusters than the combine step scans");
  CUtensorMap tm_q, tm_k, tm_v;
  {
    const uint64_t dims[4] = {64, static_cast<uint64_t>(L), 2, static_cast<uint64_t>(S)};
    const uint64_t strides[3] = {kDK * 2, 128, static_cast<uint64_t>(L) * kDK * 2};
    const uint32_t box[4] = {64, kBlockM, 2, 1};
    if (int e = sk_encode(&tm_q, q, 4, dims, strides, box)) return e;
  }
  {
    const uint64_t dims[5] = {64, static_cast<uint64_t>(L), 2, static_cast<uint64_t>(n_slots), static_cast<uint64_t>(S)};
    const uint64_t strides[4] = {kDK * 2, 128, static_cast<uint64_t>(L) * kDK * 2,
                                 static_cast<uint64_t>(n_slots) * L * kDK * 2};
    const uint32_t box[5] = {64, kBlockN, 1, 1, 1};              // one 64-channel block: half a K tile per CTA
    if (int e = sk_encode(&tm_k, kbank, 5, dims, strides, box)) return e;
  }
  {
    const uint64_t dims[5] = {64, static_cast<uint64_t>(L), kDV / 64, static_cast<uint64_t>(n_slots),
                              static_cast<uint64_t>(S)};
    const uint64_t strides[4] = {kDV * 2, 128, static_cast<uint64_t>(L) * kDV * 2,
                                 static_cast<uint64_t>(n_slots) * L * kDV * 2};
    const uint32_t box[5] = {64, kBlockN, 2, 1, 1};              // two of the split's four blocks per CTA
    if (int e = sk_encode(&tm_v, vbank, 5, dims, strides, box)) return e;
  }
  static SmemOptIn optin;
  if (int r = ensure_dynamic_smem(reinterpret_cast<const void*>(lt_attn_sk_kernel), &optin, kSmemTotal,
                                  "lt_attn_sk: cudaFuncSetAttribute"))
    return r;
  Params p;
  p.L = L; p.T = T; p.tpf = pl.tpf; p.P = pl.P; p.U = pl.U; p.n_clusters = pl.n_clusters; p.max_seg = pl.max_seg;
  p.S = S; p.n_full = pl.n_full; p.eps = pl.eps; p.order_stride = order_stride;
  p.order = order_dev;
  p.tpe_bias = tpe_bias;
  for (int t = 0; t < RMEM_MAX_BANK; ++t) p.tpe_rows[t] = (tpe_bias && tpe_rows_host && t < T) ? tpe_rows_host[t] : 0;
  float* part_f = nullptr;
  sk_split_workspace(pl, workspace, &p.part_o, &p.part_ml, &part_f);
  p.part_f = record ? part_f : nullptr;
  p.scale_log2 = 1.4426950408889634f / sqrtf(static_cast<float>(kDK));
  launch_k(lt_attn_sk_kernel, pl.n_ctas, kThreads, kSmemTotal, stream, tm_q, tm_k, tm_v, p);
  return check_launch("lt_attn_sk_kernel");
}

int lt_attn_sk_combine_launch(int T, int L, int S, const void* gate, void* out, float* attn_mass, int n_clusters,
                              const void* workspace, int64_t workspace_bytes, cudaStream_t stream) {
  using namespace sk;
  if (T < 1 || T > RMEM_MAX_BANK || L < 1 || S < 1) return set_error(RMEM_ERR_ARG, "lt_attn_sk_combine: bad T / L / batch");
  const Plan pl = make_plan(L, T, S, sm_count(), n_clusters);
  if (!workspace || workspace_bytes < static_cast<int64_t>(pl.n_ctas) * pl.max_seg * slot_words(true) * 4)
    return set_error(RMEM_ERR_ARG, "lt_attn_sk_combine: workspace too small");
  CombineParams p;
  p.L = L; p.T = T; p.tpf = pl.tpf; p.P = pl.P; p.U = pl.U; p.n_clusters = pl.n_clusters; p.max_seg = pl.max_seg;
  p.n_q_tiles = (L + kBlockM - 1) / kBlockM; p.n_full = pl.n_full; p.eps = pl.eps;
  float *po, *pml, *pf;
  sk_split_workspace(pl, const_cast<void*>(workspace), &po, &pml, &pf);
  p.part_o = po; p.part_ml = pml; p.part_f = pf;
  p.gate = static_cast<const __nv_bfloat16*>(gate);
  p.out = static_cast<__nv_bfloat16*>(out);
  p.attn_mass = attn_mass;
  dim3 grid(S * ((L + kBlockM - 1) / kBlockM), kDV / kCbCols + (attn_mass ? 1 : 0));
  launch_k(lt_attn_sk_combine_kernel, grid, 256, 0, stream, p);
  return check_launch("lt_attn_sk_combine_kernel");
}

}  // namespace rmem
