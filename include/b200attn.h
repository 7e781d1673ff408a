/*
 * b200attn.h — C ABI of the B200-native attention / KV-cache hot path for nano-vllm-x.
 *
 * One shared library (libb200attn.so, nvcc -gencode arch=compute_100a,code=sm_100a).  Plain
 * pointers and sizes only; every pointer is a DEVICE pointer owned by the caller (PyTorch in
 * the shipped host side) unless the name ends in `_host`.  The library owns no device memory
 * (one exception: the fixed scratch array of b200_decode_noalloc, section 3b),
 * never allocates, never synchronises and launches only on the stream it is given, so the
 * entry points that take all their memory from the caller can be captured in a CUDA graph
 * (reference: core/cuda_graph.py:27-39, core/model_runner.py:407-416).
 *
 * Return value: 0 on success, non-zero on error; b200_last_error() then returns a
 * thread-local message.  Asynchronous faults surface on the caller's next sync, as with the
 * reference's Triton / FlashInfer launches.
 *
 * Citations are into /root/reference/nanovllmx/.
 */
#ifndef B200ATTN_H_
#define B200ATTN_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef void* b200_stream_t; /* cudaStream_t */

#define B200ATTN_ABI_VERSION 1

/* dtype codes for K/V/Q/O element type */
#define B200_DTYPE_BF16 0
#define B200_DTYPE_FP16 1
#define B200_DTYPE_FP32 2 /* b200_sample logits only */

int b200_abi_version(void);
const char* b200_last_error(void);
/* Number of SMs of the current device (used by the host to size workspaces). <0 on error. */
int b200_num_sms(void);

/* ------------------------------------------------------------------------------------------
 * (1) KV slot write.  Replaces the Triton kernel `store_kvcache_kernel` + launcher
 *     `store_kvcache` (core/kv_cache.py:12-44), called through KVCachePool.set_kv_cache
 *     (core/kv_cache.py:90-99) from Attention.forward (layers/attention.py:347-352).
 *
 *     k, v        : [T, row_elems] rows, element stride 1, row strides k_row_stride /
 *                   v_row_stride IN ELEMENTS (v is a strided view of the qkv GEMM output,
 *                   models/qwen3.py:123)
 *     k_cache,    : [num_slots, row_elems] contiguous layer view of the pool
 *     v_cache       (core/kv_cache.py:69-79), row_elems = Hkv_local * head_dim
 *     slots       : [T] int64 (layers/attention.py:134-141); slot < 0 (the reference uses -1,
 *                   core/kv_cache.py:25-26) => row is skipped
 *     elem_bytes  : 2 (bf16 / fp16); rows are moved as raw bytes, so the copy is bit-exact
 * ---------------------------------------------------------------------------------------- */
int b200_store_kv(const void* k, int64_t k_row_stride, const void* v, int64_t v_row_stride,
                  void* k_cache, void* v_cache, const int64_t* slots, int32_t T,
                  int32_t row_elems, int32_t elem_bytes, b200_stream_t stream);

/* ------------------------------------------------------------------------------------------
 * (2) Device-side metadata build.  Replaces the torch/host index construction of
 *     AttentionBackend.build_metadata (layers/attention.py:62-101, 134-141) and
 *     build_metadata_for_cuda_graph_replay (layers/attention.py:219-275).
 *
 *     The worker keeps a device-resident slot table `slot_table[row][col]` (int32): row =
 *     one live sequence, col = token position, value = KV slot id.  Each step the host
 *     sends ONE packed int32 message (`step_msg`, already on the device) instead of the
 *     reference's 2*B per-sequence H2D copies and B `.item()` syncs:
 *
 *       step_msg[0*B .. 1*B)  row        slot-table row of sequence i
 *       step_msg[1*B .. 2*B)  kv_len     len(seq.kv_indices)
 *       step_msg[2*B .. 3*B)  cached_len seq.cached_kv_len   (new slots = [cached_len, kv_len))
 *       step_msg[3*B .. 4*B)  q_len      len(seq.token_ids) - seq.cached_kv_len
 *       step_msg[4*B .. 5*B)  wr_col     first table column written this step
 *       step_msg[5*B .. 6*B)  wr_cnt     number of slot ids appended to the row this step
 *       step_msg[6*B .. 7*B)  wr_off     offset of those ids inside the payload
 *       step_msg[7*B .. 7*B+n_payload)   payload: the new slot ids, int32
 *
 *     b200_slot_table_update scatters the payload into the table;
 *     b200_build_meta then produces, bit-exactly, the arrays the reference builds:
 *
 *       qo_indptr       int32 [B+1]           (may be NULL)
 *       kv_indptr       int32 [padded_bs+1]
 *       kv_indices      int32 [kv_indices_cap]  first sum(pages) entries = page table
 *       last_page_len   int32 [padded_bs]     (may be NULL: the graph-replay path of the
 *                                             reference never rewrites it)
 *       out_slots       int64 [out_slots_cap]
 *
 *     page_size must be a power of two.  For page_size > 1 (not defined by the reference,
 *     SURVEY.md a-3) pages are page-aligned runs of slots: entry = slot_of_first_token / ps.
 *
 *     padded_bs >= B: sequences B..padded_bs-1 are CUDA-graph padding with kv_len 1
 *     (layers/attention.py:239-244).  If fill_tail != 0 the reference's whole-buffer fills
 *     are reproduced: kv_indices[sum(pages) .. kv_indices_cap) = 0 (:258) and
 *     out_slots[sum(new) .. out_slots_cap) = -1 (:265).
 * ---------------------------------------------------------------------------------------- */
int b200_slot_table_update(int32_t* slot_table, int64_t table_row_stride, const int32_t* step_msg,
                           int32_t B, int32_t n_payload, b200_stream_t stream);

int b200_build_meta(const int32_t* slot_table, int64_t table_row_stride, const int32_t* step_msg,
                    int32_t B, int32_t padded_bs, int32_t page_size, int32_t fill_tail,
                    int32_t* qo_indptr, int32_t* kv_indptr, int32_t* kv_indices,
                    int64_t kv_indices_cap, int32_t* last_page_len, int64_t* out_slots,
                    int64_t out_slots_cap, int32_t* scratch /* >= padded_bs+1 int32 */,
                    b200_stream_t stream);

/* ------------------------------------------------------------------------------------------
 * (3) Paged GQA decode attention.  Replaces Attention.forward's DECODE branch
 *     (layers/attention.py:365-369) = FlashInfer BatchDecodeWithPagedKVCacheWrapper.plan+run
 *     as constructed at layers/attention.py:54-59, 187-207, 277-288.  No host plan: the
 *     split-KV schedule is computed on the device from kv_indptr, the grid depends only on
 *     the device (one CTA per SM), so a captured launch stays valid for any batch contents.
 *
 *     q, out      : [B, Hq, D], rows q_row_stride / out_row_stride elements apart
 *     k_cache,    : [num_slots, Hkv, D] layer view (NHD), kv_slot_stride elements between
 *     v_cache       token slots, D between heads
 *     kv_indptr   : int32 [B+1]; kv_indices: page ids; last_page_len: int32 [B]
 *     page_size   : power of two; token t of a sequence lives in slot
 *                   kv_indices[indptr + t / ps] * ps + t % ps
 *     workspace   : >= b200_decode_workspace_bytes(B, Hq, Hkv, D) bytes, 16-byte aligned
 *     Only D == 128 and dtype bf16 are built in round 1.
 * ---------------------------------------------------------------------------------------- */
size_t b200_decode_workspace_bytes(int32_t B, int32_t Hq, int32_t Hkv, int32_t D);

int b200_decode(const void* q, int64_t q_row_stride, const void* k_cache, const void* v_cache,
                int64_t kv_slot_stride, const int32_t* kv_indptr, const int32_t* kv_indices,
                const int32_t* last_page_len, void* out, int64_t out_row_stride, void* workspace,
                size_t workspace_bytes, int32_t B, int32_t Hq, int32_t Hkv, int32_t D,
                int32_t page_size, float sm_scale, int32_t dtype, b200_stream_t stream);

/* ------------------------------------------------------------------------------------------
 * (3b) b200_decode without a workspace argument.  NO COUNTERPART IN THE REFERENCE: FlashInfer's
 *     wrappers are given a torch workspace tensor (layers/attention.py:44-48) and every host of the
 *     reference has torch's allocator.  This entry point is for hosts of the C ABI that cannot obtain
 *     device memory when they call: the engine gave all free HBM to the KV pool at start-up
 *     (core/model_runner.py:113-144 sizes the pool that way), cudaMalloc / cudaFree are off limits
 *     because they stall every stream of a running worker, or the caller is a foreign-language
 *     binding / plugin thread with no device allocator at hand.
 *
 *     Same arguments, kernels and results as b200_decode; the split-KV partials go to an array of
 *     B200_DECODE_NOALLOC_SCRATCH_BYTES that is part of the library image.  Consequences:
 *       - every GPU on which the library's module is loaded gives up that many bytes until exit;
 *       - shapes with b200_decode_workspace_bytes(B, Hq, Hkv, D) above the limit are refused
 *         (non-zero return, message with both sizes) -- use b200_decode for those;
 *       - the array is shared by all calls on a device and is not cleared between them (the kernels
 *         write every word they later read, so stale contents are harmless); calls issued on one
 *         stream are ordered, calls that may run at the same time on the GPU race on it;
 *       - the first call on each device resolves the array's device address (cudaGetSymbolAddress,
 *         which may load the module): make that call with the device current and outside any stream
 *         capture.  The call allocates nothing and frees nothing, then or later.
 * ---------------------------------------------------------------------------------------- */
#define B200_DECODE_NOALLOC_SCRATCH_BYTES (8u << 20)

int b200_decode_noalloc(const void* q, int64_t q_row_stride, const void* k_cache, const void* v_cache,
                        int64_t kv_slot_stride, const int32_t* kv_indptr, const int32_t* kv_indices,
                        const int32_t* last_page_len, void* out, int64_t out_row_stride, int32_t B,
                        int32_t Hq, int32_t Hkv, int32_t D, int32_t page_size, float sm_scale,
                        int32_t dtype, b200_stream_t stream);

/* ------------------------------------------------------------------------------------------
 * (4) Ragged paged prefill attention (prefix pages + new tokens, bottom-right causal).
 *     Replaces Attention.forward's PREFILL branch (layers/attention.py:360-364) =
 *     FlashInfer BatchPrefillWithPagedKVCacheWrapper.plan+run (layers/attention.py:49-53,
 *     106-119).  Query j of sequence i (Lq new, Lkv total tokens) attends kv positions
 *     <= Lkv - Lq + j.  K/V of the new tokens must already be in the pool (b200_store_kv).
 *
 *     q, out      : [T, Hq, D]; qo_indptr int32 [B+1]; total_q = T (host-known upper bound of
 *                   qo_indptr[B]; sizes the work list, no device sync)
 *     workspace   : >= b200_prefill_workspace_bytes(B, total_q, Hq, Hkv, D) bytes, 16-byte aligned.
 *                   Holds the device-built work list (one (sequence, query tile) pair per 256 packed
 *                   query rows, heaviest first) and the counter the persistent CTAs draw from; a
 *                   short workspace is an error (no fallback).
 * ---------------------------------------------------------------------------------------- */
size_t b200_prefill_workspace_bytes(int32_t B, int32_t total_q, int32_t Hq, int32_t Hkv, int32_t D);

int b200_prefill(const void* q, int64_t q_row_stride, const void* k_cache, const void* v_cache,
                 int64_t kv_slot_stride, const int32_t* qo_indptr, const int32_t* kv_indptr,
                 const int32_t* kv_indices, const int32_t* last_page_len, void* out,
                 int64_t out_row_stride, void* workspace, size_t workspace_bytes, int32_t B,
                 int32_t total_q, int32_t Hq, int32_t Hkv, int32_t D, int32_t page_size,
                 float sm_scale, int32_t causal, int32_t dtype, b200_stream_t stream);

/* ------------------------------------------------------------------------------------------
 * (f-3) Fused q/k RMSNorm + rotary embedding + KV slot write: the step just ahead of
 *     Attention.forward in Qwen3Attention.forward (models/qwen3.py:122-134).  One launch replaces
 *     q_norm / k_norm (layers/layernorm.py:16-36), rotary_emb (layers/rotary_embedding.py:73-102)
 *     and the slot write of Attention.forward (layers/attention.py:347-352 -> core/kv_cache.py:12-44);
 *     the caller then runs Attention.forward(q, None, None) (attention only).
 *
 *     qkv         : [T, (Hq + 2 Hkv) * D] output of qkv_proj, rows qkv_row_stride elements apart,
 *                   laid out [q heads | k heads | v heads] (models/qwen3.py:123)
 *     positions   : int64 [T] (core/model_runner.py:175-189)
 *     q_weight,   : [D] RMSNorm weights of q_norm / k_norm, activation dtype
 *     k_weight
 *     cos_sin_cache : fp32 [max_pos, D] = [cos(D/2) | sin(D/2)] per position, rows
 *                   cos_sin_row_stride elements apart (RotaryEmbedding.cos_sin_cache,
 *                   layers/rotary_embedding.py:62-71; rotary_dim == head_dim)
 *     q_out       : [T, Hq * D]; may alias the q part of qkv (in place)
 *     k_out       : [T, Hkv * D] or NULL (the normalised+rotated K is otherwise only written to the pool)
 *     k_cache,    : [num_slots, Hkv, D] layer views of the pool; slots int64 [T], slot < 0 =>
 *     v_cache       that token's K/V are not written (CUDA-graph padding, layers/attention.py:265)
 *     Output layout of the rotary step is the reference's INTERLEAVED one (out[2i] = o1[i],
 *     out[2i+1] = o2[i], layers/rotary_embedding.py:24), so pools stay interchangeable.
 *     Only D == 128 and bf16 are built.
 * ---------------------------------------------------------------------------------------- */
int b200_qk_norm_rope_store(const void* qkv, int64_t qkv_row_stride, const int64_t* positions,
                            const void* q_weight, const void* k_weight, float eps,
                            const float* cos_sin_cache, int64_t cos_sin_row_stride, int32_t max_pos,
                            void* q_out, int64_t q_out_row_stride, void* k_out, int64_t k_out_row_stride,
                            void* k_cache, void* v_cache, const int64_t* slots, int32_t T, int32_t Hq,
                            int32_t Hkv, int32_t D, int32_t dtype, b200_stream_t stream);

/* ------------------------------------------------------------------------------------------
 * (f-4) Sampler tail.  Replaces Sampler.forward (layers/sampler.py:19-55) with its helpers
 *     _apply_top_k_top_p (:58-90, a full-vocabulary sort) and _apply_min_p (:93-106), called from
 *     ModelRunner.execute_model (core/model_runner.py:262-272) with the tensors of
 *     prepare_sampling_params (:191-219).  One launch, no sort, no [B, V] temporaries.
 *
 *     logits      : [B, V] bf16 (what compute_logits returns) or fp32, rows row_stride elements apart
 *     temperatures: fp32 [B] or NULL          top_ks : int64 [B] or NULL (>= V = off; the reference
 *     top_ps      : fp32 [B] or NULL                   sends vocab_size for top_k == -1, :203-206)
 *     min_ps      : fp32 [B] or NULL
 *                   all four NULL => greedy argmax (layers/sampler.py:28-36); a row whose temperature
 *                   is <= 0 is greedy too (the reference would divide by zero)
 *     noise       : fp32 [B, V] Exp(1) samples, rows noise_row_stride apart, or NULL.  torch.multinomial
 *                   (layers/sampler.py:53) is argmax(probs / q) with q = empty_like(probs).exponential_(1);
 *                   passing that q reproduces the reference's draw.  NULL => q comes from
 *                   Philox-4x32-10(counter = (column / 4, row, offset), key = seed), word column % 4,
 *                   u = ((word >> 8) + 0.5) * 2^-24, q = -log u.
 *     out_ids     : int64 [B]
 *     Order of operations as the reference: logits / T, top-k (ties at the k-th value kept), top-p
 *     over the top-k survivors in ascending (value, index) order with the last one always kept,
 *     min-p, then the draw.  V <= 2^24.
 * ---------------------------------------------------------------------------------------- */
int b200_sample(const void* logits, int64_t row_stride, int32_t logits_dtype, int32_t B, int32_t V,
                const float* temperatures, const int64_t* top_ks, const float* top_ps,
                const float* min_ps, const float* noise, int64_t noise_row_stride, uint64_t seed,
                uint64_t offset, int64_t* out_ids, b200_stream_t stream);

/* Counters for bench.py's `gpu_launches`: number of kernels this library has launched in this
 * process since load / since the last reset. */
int64_t b200_launch_count(void);
void b200_reset_launch_count(void);

#ifdef __cplusplus
}
#endif
#endif /* B200ATTN_H_ */
