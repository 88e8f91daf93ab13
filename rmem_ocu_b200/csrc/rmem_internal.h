// Internal declarations shared by the kernels and the C-ABI translation unit (capi.cu).
#pragma once
#include <cuda.h>
#include <cuda_runtime.h>
#include <stdint.h>
#include "../../include/rmem_b200.h"

namespace rmem {

// ---- error reporting (thread-local last-error string, never throws across the ABI)
int set_error(int code, const char* msg);
int set_cuda_error(cudaError_t e, const char* where);
int check_launch(const char* kernel_name);

typedef CUresult (*TensorMapEncodeFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*,
                                      const cuuint64_t*, const cuuint64_t*, const cuuint32_t*,
                                      const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                                      CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
TensorMapEncodeFn get_tensor_map_encoder();

// ---- launchers (one per kernel family); all asynchronous on `stream`
int lt_attn_partial_launch(const void* q, const void* kbank, const void* vbank,
                           const int32_t* order_dev, int T, int L, int n_slots,
                           int chunks_per_frame, float* part_o, float* part_m, float* part_l,
                           cudaStream_t stream);
int lt_attn_combine_launch(const float* part_o, const float* part_m, const float* part_l,
                           const float* tpe_bias, const int32_t* tpe_rows_host, int T, int L,
                           int chunks_per_frame, const void* gate, void* out, float* attn_mass,
                           cudaStream_t stream);

}  // namespace rmem
