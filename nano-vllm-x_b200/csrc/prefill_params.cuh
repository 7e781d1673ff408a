// Parameter block shared by the prefill kernels (one definition: both translation units see the same layout).
#pragma once
#include <cuda_bf16.h>
#include <stdint.h>

namespace b200 {
struct PrefillParams {
  const __nv_bfloat16* q;
  const __nv_bfloat16* k_cache;
  const __nv_bfloat16* v_cache;
  const int32_t* qo_indptr;
  const int32_t* kv_indptr;
  const int32_t* kv_indices;
  const int32_t* last_page_len;
  __nv_bfloat16* out;
  int64_t q_row_stride, out_row_stride, kv_slot_stride;
  int B, Hq, Hkv, group, log2_ps, causal;
  float scale_log2;
};
int launch_prefill_mma(const PrefillParams& p, int total_q, cudaStream_t st);
int launch_prefill_sm100(const PrefillParams& p, int total_q, void* workspace, size_t workspace_bytes, cudaStream_t st);
size_t prefill_sm100_workspace_bytes(int B, int total_q, int group);
bool prefill_sm100_supports(int group);
}  // namespace b200
