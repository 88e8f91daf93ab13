// extern "C" surface of librmem_b200.so (declared in include/rmem_b200.h) + error plumbing.
#include <stdio.h>
#include <string.h>
#include "rmem_internal.h"

namespace rmem {

static thread_local char g_err[512] = "";

int set_error(int code, const char* msg) {
  snprintf(g_err, sizeof(g_err), "%s", msg);
  return code;
}
int set_cuda_error(cudaError_t e, const char* where) {
  snprintf(g_err, sizeof(g_err), "%s: %s", where, cudaGetErrorString(e));
  return RMEM_ERR_CUDA;
}
int check_launch(const char* kernel_name) {
  cudaError_t e = cudaPeekAtLastError();
  if (e != cudaSuccess) {
    cudaGetLastError();
    return set_cuda_error(e, kernel_name);
  }
  return RMEM_OK;
}

TensorMapEncodeFn get_tensor_map_encoder() {
  static TensorMapEncodeFn fn = nullptr;
  static bool tried = false;
  if (!tried) {
    tried = true;
    void* p = nullptr;
    cudaDriverEntryPointQueryResult qres;
    cudaError_t e = cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &qres);
    if (e == cudaSuccess && qres == cudaDriverEntryPointSuccess) fn = reinterpret_cast<TensorMapEncodeFn>(p);
  }
  return fn;
}

}  // namespace rmem

using namespace rmem;

extern "C" {

int rmem_abi_version(void) { return RMEM_ABI_VERSION; }
const char* rmem_last_error(void) { return g_err; }

int rmem_lt_attn_deaot_partial(const void* q, const void* kbank, const void* vbank,
                               const int32_t* order, int T, int L, int n_slots,
                               int chunks_per_frame, float* part_o, float* part_m, float* part_l,
                               void* stream) {
  return lt_attn_partial_launch(q, kbank, vbank, order, T, L, n_slots, chunks_per_frame, part_o,
                                part_m, part_l, static_cast<cudaStream_t>(stream));
}

int rmem_lt_attn_deaot_combine(const float* part_o, const float* part_m, const float* part_l,
                               const float* tpe_bias, const int32_t* tpe_rows_host, int T, int L,
                               int chunks_per_frame, const void* gate, void* out,
                               float* attn_mass, void* stream) {
  return lt_attn_combine_launch(part_o, part_m, part_l, tpe_bias, tpe_rows_host, T, L,
                                chunks_per_frame, gate, out, attn_mass,
                                static_cast<cudaStream_t>(stream));
}

}  // extern "C"
