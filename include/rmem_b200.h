/* rmem_b200.h -- C ABI of librmem_b200.so: the B200 (sm_100a) kernels behind the RMem per-frame
 * memory-reading hot path of AOT / DeAOT (reference: Bardli/RMem_ocu, aot_plus/networks/...).
 *
 * The reference has no FFI of its own (it is pure PyTorch); these entry points are what a binding
 * for this path binds -- one per operator the reference evaluates on the path -- and each cites
 * the reference code it replaces.  INTEGRATION.md shows the ctypes stub a maintainer would add.
 *
 * Conventions
 *   - every pointer is a DEVICE pointer unless its name ends in _host; the caller owns all buffers,
 *     the library never allocates device memory; workspaces are passed in;
 *   - every call is asynchronous on `stream` (a cudaStream_t passed as void*) and re-entrant across
 *     streams; no global mutable state besides a thread-local error string;
 *   - return value: RMEM_OK (0) or a negative RMEM_ERR_* code; rmem_last_error() explains it;
 *   - bf16 = IEEE bfloat16 (uint16 storage), f32 = IEEE binary32; row-major, innermost dim last.
 */
#ifndef RMEM_B200_H_
#define RMEM_B200_H_

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define RMEM_ABI_VERSION 1

#define RMEM_OK 0
#define RMEM_ERR_ARG (-1)     /* invalid argument (shape / capacity)           */
#define RMEM_ERR_CUDA (-2)    /* a CUDA runtime call or kernel launch failed  */
#define RMEM_ERR_DRIVER (-3)  /* driver entry point (tensor map encode) failed */

#define RMEM_MAX_BANK 16             /* max frames in the long-term bank (former + latter + 1) */
#define RMEM_MAX_CHUNKS_PER_FRAME 4  /* split-KV granularity inside one memory frame          */
#define RMEM_TPE_ROWS 4              /* rows of mem_pos_emb (TEMPORAL_POSITIONAL_EMBEDDING_SLOT_4) */

int rmem_abi_version(void);
const char* rmem_last_error(void);

/* ---------------------------------------------------------------------------------------------
 * Long-term gated attention, DeAOT (1 head, d_att 128, value width 1024 = V | ID_V).
 * Replaces GatedPropagation.forward explicit path, layers/attention.py:174-193, as called from
 * GatedPropagationModule.forward, layers/transformer.py:1183-1184, with the bank layout of
 * DualBranchGPM.long_term_memories (transformer.py:859-878) turned into fixed slots.
 *
 *   q        bf16 [L][128]            current-frame queries, cur_pos_emb already added
 *   kbank    bf16 [n_slots][L][128]   memory keys, one slot per stored frame (no TPE added)
 *   vbank    bf16 [n_slots][L][1024]  memory values (V | ID_V)
 *   order    i32  [T]                 logical bank position (oldest first) -> slot
 *   part_o   f32  [T*cpf][L][1024]    un-normalised partial outputs per key chunk
 *   part_m   f32  [T*cpf][L]          running max per chunk (log2 units, scaled by 1/sqrt(128))
 *   part_l   f32  [T*cpf][L]          softmax denominators per chunk
 */
int rmem_lt_attn_deaot_partial(const void* q, const void* kbank, const void* vbank,
                               const int32_t* order, int T, int L, int n_slots,
                               int chunks_per_frame, float* part_o, float* part_m, float* part_l,
                               void* stream);

/* Merge of the chunk partials + temporal positional embedding + recording + gate.
 * Replaces the TPE add of transformer.py:1140-1172 (as a per-frame score bias q'.e_row/sqrt(d)),
 * the softmax normalisation of attention.py:187-193, the attention-weight recording of
 * transformer.py:1185-1192 and the gate `outputs * U` of attention.py:206.
 *
 *   tpe_bias       f32 [L][4]    log2e/sqrt(128) * (q' . mem_pos_emb[r]), or NULL (no TPE)
 *   tpe_rows_host  i32 [T]       HOST array: mem_pos_emb row used by bank position t
 *   gate           bf16 [L][1024] cat_curr_U, or NULL
 *   out            bf16 [L][1024]
 *   attn_mass      f32 [L][T]    record_attn_weight, or NULL
 */
int rmem_lt_attn_deaot_combine(const float* part_o, const float* part_m, const float* part_l,
                               const float* tpe_bias, const int32_t* tpe_rows_host, int T, int L,
                               int chunks_per_frame, const void* gate, void* out,
                               float* attn_mass, void* stream);

#ifdef __cplusplus
}
#endif
#endif /* RMEM_B200_H_ */
