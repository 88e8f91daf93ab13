// Restricted long-term memory bank: device-resident slot table, append, and the relevance +
// freshness (UCB) eviction rule -- reference restrict_long_memories,
// aot_plus/networks/layers/transformer.py:324-436 (AOT) / :880-991 (DeAOT), and
// update_long_term_memory, :306-322 / :859-878.
//
// The bank tensors are fixed-capacity [n_slots][L][C] buffers; `order` maps the logical position
// (oldest first -- what the temporal positional embedding is indexed by) to a physical slot, so an
// eviction is an index shuffle and never moves feature data.  All scalars the Python reference
// keeps in dicts keyed by frame id (stored_attn_weight_dict, stored_frame_times) are kept per slot:
// frame ids are unique, a dict entry dies with its frame, so the two are equivalent.
#include "rmem_internal.h"

namespace rmem {

struct BankState {
  int32_t order[RMEM_MAX_BANK];
  int32_t frame_id[RMEM_MAX_BANK];
  float ema[RMEM_MAX_BANK];
  int32_t ema_valid[RMEM_MAX_BANK];
  int32_t times[RMEM_MAX_BANK];
  int32_t last_drop;       // logical index dropped by the last score_evict (-1: nothing dropped)
  int32_t last_drop_frame; // its frame id
  float last_margin;       // second-smallest minus smallest objective among candidates
  int32_t n_evictions;
  int32_t last_choice;     // argmin of the last scoring pass even when nothing had to be dropped
  int32_t cur_frame;       // device-side frame_step (so a captured CUDA graph stamps the right frame id)
  int32_t pad[2];
};
static_assert(sizeof(BankState) == RMEM_BANK_STATE_WORDS * 4, "bank state layout");

__global__ void bank_init_kernel(BankState* st, int frame_step) {
  const int i = threadIdx.x;
  if (i < RMEM_MAX_BANK) {
    st->order[i] = i;
    st->frame_id[i] = (i == 0) ? frame_step : -1;
    st->ema[i] = 0.f;
    st->ema_valid[i] = 0;
    st->times[i] = 0;
  }
  if (i == 0) {
    st->last_drop = -1; st->last_drop_frame = -1; st->last_margin = 0.f; st->n_evictions = 0;
    st->last_choice = -1; st->cur_frame = frame_step; st->pad[0] = st->pad[1] = 0;
  }
}

__global__ void bank_tick_kernel(BankState* st) { st[blockIdx.x].cur_frame += 1; }   // one state per sequence

int bank_tick_launch(void* state, cudaStream_t stream) { return bank_tick_batched_launch(state, 1, stream); }

int bank_tick_batched_launch(void* state, int S, cudaStream_t stream) {
  static_assert(sizeof(BankState) == RMEM_BANK_STATE_WORDS * 4, "bank states are stacked RMEM_BANK_STATE_WORDS apart");
  if (S < 1) return set_error(RMEM_ERR_ARG, "bank_tick: bad batch");
  bank_tick_kernel<<<S, 1, 0, stream>>>(static_cast<BankState*>(state));
  return check_launch("bank_tick_kernel");
}

int bank_init_launch(void* state, int frame_step, cudaStream_t stream) {
  bank_init_kernel<<<1, 32, 0, stream>>>(static_cast<BankState*>(state), frame_step);
  return check_launch("bank_init_kernel");
}

// ------------------------------------------------------------------------------------------------
struct AppendParams {
  const uint4* src[8];
  uint4* dst[8];
  long long row_vec[8];  // uint4 per slot
  int n;
};

__global__ void __launch_bounds__(256)
bank_append_kernel(BankState* st, int T, int frame_step, const AppendParams p) {
  const int slot = st->order[T];
  const int t = blockIdx.y;
  const long long n = p.row_vec[t];
  const uint4* s = p.src[t];
  uint4* d = p.dst[t] + static_cast<long long>(slot) * n;
  for (long long i = blockIdx.x * static_cast<long long>(blockDim.x) + threadIdx.x; i < n;
       i += static_cast<long long>(gridDim.x) * blockDim.x)
    d[i] = s[i];
  if (blockIdx.x == 0 && t == 0 && threadIdx.x == 0)
    st->frame_id[slot] = frame_step >= 0 ? frame_step : st->cur_frame;
}

int bank_append_launch(void* state, int T, int frame_step, const void* const* src, void* const* dst,
                       const int64_t* row_bytes, int n_tensors, cudaStream_t stream) {
  if (n_tensors < 1 || n_tensors > 8 || T < 1 || T >= RMEM_MAX_BANK)
    return set_error(RMEM_ERR_ARG, "bank_append: bad tensor count or bank length");
  AppendParams p;
  p.n = n_tensors;
  for (int i = 0; i < n_tensors; ++i) {
    if (row_bytes[i] % 16) return set_error(RMEM_ERR_ARG, "bank_append: slot size must be a multiple of 16 bytes");
    p.src[i] = static_cast<const uint4*>(src[i]);
    p.dst[i] = static_cast<uint4*>(dst[i]);
    p.row_vec[i] = row_bytes[i] / 16;
  }
  dim3 grid(148, n_tensors);
  bank_append_kernel<<<grid, 256, 0, stream>>>(static_cast<BankState*>(state), T, frame_step, p);
  return check_launch("bank_append_kernel");
}

// ------------------------------------------------------------------------------------------------
// One CTA. Phase 1: s_t = sum_q W[q,t] * fg[q] (fp64 accumulation of fp32 products).
// Phase 2 (thread 0): the scalar rule, in fp32 with explicitly rounded operations so no FMA
// contraction changes a decision.
constexpr int kEvictThreads = 512;

__global__ void __launch_bounds__(kEvictThreads)
bank_score_evict_kernel(BankState* st, const float* __restrict__ W, const float* __restrict__ fg, int L,
                        int T_old, int capacity) {
  __shared__ double red[kEvictThreads / 32][RMEM_MAX_BANK];
  __shared__ float s_sh[RMEM_MAX_BANK];
  double acc[RMEM_MAX_BANK];
#pragma unroll
  for (int t = 0; t < RMEM_MAX_BANK; ++t) acc[t] = 0.0;
  for (int q = threadIdx.x; q < L; q += kEvictThreads) {
    const float f = fg ? fg[q] : 1.f;
    const float* w = W + static_cast<size_t>(q) * T_old;
#pragma unroll
    for (int t = 0; t < RMEM_MAX_BANK; ++t)
      if (t < T_old) acc[t] += static_cast<double>(__fmul_rn(w[t], f));
  }
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
#pragma unroll
  for (int t = 0; t < RMEM_MAX_BANK; ++t) {
    double v = acc[t];
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    if (lane == 0) red[warp][t] = v;
  }
  __syncthreads();
  if (threadIdx.x < RMEM_MAX_BANK) {
    double v = 0.0;
    for (int wv = 0; wv < kEvictThreads / 32; ++wv) v += red[wv][threadIdx.x];
    s_sh[threadIdx.x] = static_cast<float>(v);
  }
  __syncthreads();
  if (threadIdx.x != 0) return;

  const int T_new = T_old + 1;
  float total = 0.f;
  for (int t = 0; t < T_old; ++t) total = __fadd_rn(total, s_sh[t]);
  float r[RMEM_MAX_BANK];
  for (int t = 0; t < T_old; ++t) {
    const int slot = st->order[t];
    float v = __fdiv_rn(s_sh[t], total);
    if (st->ema_valid[slot]) v = __fadd_rn(__fmul_rn(0.2f, st->ema[slot]), __fmul_rn(0.8f, v));
    st->ema[slot] = v;
    st->ema_valid[slot] = 1;
    r[t] = v;
  }
  for (int t = 0; t < T_new; ++t) st->times[st->order[t]] += 1;
  float n[RMEM_MAX_BANK];
  float sum_n = 0.f;
  for (int t = 0; t < T_old; ++t) {
    n[t] = (t == 0) ? static_cast<float>(T_old) : static_cast<float>(st->times[st->order[t]]);
    sum_n = __fadd_rn(sum_n, n[t]);
  }
  const float log_sum = static_cast<float>(log(static_cast<double>(sum_n)));
  int best = 1;
  float best_v = INFINITY, second_v = INFINITY;
  for (int t = 1; t < T_old; ++t) {
    const float bonus = __fmul_rn(1.5f, __fsqrt_rn(__fdiv_rn(log_sum, __fadd_rn(n[t], 8.f))));
    const float o = __fadd_rn(r[t], bonus);
    if (o < best_v) { second_v = best_v; best_v = o; best = t; }
    else if (o < second_v) { second_v = o; }
  }
  st->last_choice = best;
  st->last_margin = second_v - best_v;
  if (T_new > capacity) {
    const int slot = st->order[best];
    st->last_drop = best;
    st->last_drop_frame = st->frame_id[slot];
    for (int t = best; t < T_new - 1; ++t) st->order[t] = st->order[t + 1];
    st->order[T_new - 1] = slot;
    st->frame_id[slot] = -1;
    st->ema_valid[slot] = 0;
    st->ema[slot] = 0.f;
    st->times[slot] = 0;
    st->n_evictions += 1;
  } else {
    st->last_drop = -1;
    st->last_drop_frame = -1;
  }
}

int bank_score_evict_launch(void* state, const float* attn_mass, const float* fg, int L, int T_old,
                            int capacity, cudaStream_t stream) {
  if (T_old < 1 || T_old + 1 > RMEM_MAX_BANK || capacity < 1 || L < 1)
    return set_error(RMEM_ERR_ARG, "bank_score_evict: bad T_old / capacity");
  bank_score_evict_kernel<<<1, kEvictThreads, 0, stream>>>(static_cast<BankState*>(state), attn_mass, fg,
                                                           L, T_old, capacity);
  return check_launch("bank_score_evict_kernel");
}

}  // namespace rmem
