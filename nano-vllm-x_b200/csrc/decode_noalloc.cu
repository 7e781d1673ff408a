// b200_decode_noalloc: the decode attention of decode.cu for hosts that have no device allocator to
// take a workspace from.  No counterpart in the reference (FlashInfer's wrappers are handed a torch
// workspace tensor, layers/attention.py:44-48).
//
// Who this is for.  b200_decode wants a caller-owned workspace for its split-KV partials.  Under PyTorch
// that is one torch.empty at start-up.  Other hosts of this C ABI do not have that luxury:
//   * an engine that sizes its KV pool from "whatever HBM is free after warm-up" (what the reference does,
//     core/model_runner.py:113-144) leaves nothing for a later cudaMalloc, which then fails or, worse,
//     succeeds on one rank and not on another;
//   * a step loop that must not stall: cudaMalloc / cudaFree synchronise the whole device, so a host that
//     attaches to an already-running worker (a second serving front-end, a Go / Java / C binding, a
//     plugin loaded into someone else's process) may be forbidden to call them at all;
//   * code that runs where the allocator is simply not reachable: a callback on a worker thread without
//     the framework's allocator, or a process whose memory budget is enforced by refusing allocations.
// Such callers pass inputs and outputs only; the partials live in the array below, which belongs to the
// library image and is placed by the driver when the module is loaded on a device.
//
// Cost and limit.  Every GPU on which this library's kernels are loaded holds kScratchBytes (8 MiB) of
// HBM until the process exits, used or not.  A call whose shapes need more than that is refused with an
// error that names both numbers (b200_decode_workspace_bytes(B, Hq, Hkv, D) is the requirement; 8 MiB
// covers e.g. B <= 256 at Hq 32 / Hkv 8 and B <= 2800 at Hq 4 / Hkv 1 on a 148-SM part).
//
// Sharing.  There is ONE array per device and every call on that device uses it.  It keeps whatever the
// previous call left in it; that is harmless here because decode_kernel writes every partial, log-sum-exp
// and segment index before decode_merge_kernel reads it (b200_decode is tested on torch.empty workspaces),
// so nothing has to be cleared between calls.  Calls that follow each other on one stream are ordered and
// safe.  Two calls that can overlap on the GPU (different streams, or different threads on different
// streams) write the same partial slots and race: give concurrent callers b200_decode with a workspace
// each, or serialise them.
//
// Set-up.  In host code `g_decode_scratch` is only a handle for the runtime, not a device address.  The
// first call on each device (with that device current) asks the runtime for the real address with
// cudaGetSymbolAddress, checks the result and caches it per device ordinal; that lookup may load the
// module, so make the first call on a device outside any stream capture.  After it the entry point does
// exactly what b200_decode does: two kernel launches on the caller's stream, no allocation, no sync.
#include "common.cuh"

namespace b200 {

constexpr size_t kScratchBytes = B200_DECODE_NOALLOC_SCRATCH_BYTES;
static __device__ __align__(256) unsigned char g_decode_scratch[kScratchBytes];

// Device address of g_decode_scratch on the current device, or nullptr after fail().
static void* decode_scratch_ptr() {
  static std::atomic<void*> cached[64];
  int dev = 0;
  cudaError_t e = cudaGetDevice(&dev);
  if (e != cudaSuccess) {
    fail("b200_decode_noalloc: cudaGetDevice: %s", cudaGetErrorString(e));
    return nullptr;
  }
  const bool slot = dev >= 0 && dev < 64;
  if (slot) {
    if (void* p = cached[dev].load(std::memory_order_acquire)) return p;
  }
  void* p = nullptr;
  e = cudaGetSymbolAddress(&p, g_decode_scratch);
  if (e != cudaSuccess || p == nullptr) {
    (void)cudaGetLastError();
    fail("b200_decode_noalloc: cudaGetSymbolAddress(device %d): %s", dev, cudaGetErrorString(e));
    return nullptr;
  }
  if (slot) cached[dev].store(p, std::memory_order_release);  // same value from every thread: benign if raced
  return p;
}

}  // namespace b200

extern "C" int b200_decode_noalloc(const void* q, int64_t q_row_stride, const void* k_cache, const void* v_cache,
                                   int64_t kv_slot_stride, const int32_t* kv_indptr, const int32_t* kv_indices,
                                   const int32_t* last_page_len, void* out, int64_t out_row_stride, int32_t B,
                                   int32_t Hq, int32_t Hkv, int32_t D, int32_t page_size, float sm_scale,
                                   int32_t dtype, b200_stream_t stream) {
  using namespace b200;
  if (B < 0) return fail("b200_decode_noalloc: B < 0");
  if (B == 0) return 0;
  const size_t need = b200_decode_workspace_bytes(B, Hq, Hkv, D);
  if (need == 0) return fail("b200_decode_noalloc: unsupported shape B %d Hq %d Hkv %d D %d", B, Hq, Hkv, D);
  if (need > kScratchBytes)
    return fail("b200_decode_noalloc: B %d Hq %d Hkv %d needs %zu bytes of scratch, the built-in array holds %zu; "
                "use b200_decode with a caller-owned workspace", B, Hq, Hkv, need, kScratchBytes);
  void* scratch = decode_scratch_ptr();
  if (!scratch) return 1;
  return b200_decode(q, q_row_stride, k_cache, v_cache, kv_slot_stride, kv_indptr, kv_indices, last_page_len, out,
                     out_row_stride, scratch, kScratchBytes, B, Hq, Hkv, D, page_size, sm_scale, dtype, stream);
}
