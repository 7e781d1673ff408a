// Shared device helpers for the b200attn kernels (sm_100a only).
#pragma once
#include <cuda_bf16.h>
#include <cuda_runtime.h>
#include <stdint.h>

#include <atomic>
#include <cstdarg>
#include <cstdio>

#include "../../include/b200attn.h"

namespace b200 {

// ---- host-side error plumbing ---------------------------------------------------------------
extern thread_local char g_err[512];
extern std::atomic<int64_t> g_launches;

inline int fail(const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(g_err, sizeof(g_err), fmt, ap);
  va_end(ap);
  return 1;
}

inline int check_launch(const char* what) {
  cudaError_t e = cudaPeekAtLastError();
  if (e != cudaSuccess) {
    (void)cudaGetLastError();
    return fail("%s: %s", what, cudaGetErrorString(e));
  }
  g_launches.fetch_add(1, std::memory_order_relaxed);
  return 0;
}

int cached_num_sms();

// cudaFuncAttributeMaxDynamicSharedMemorySize is a per-DEVICE attribute of a kernel: set it once per (kernel, device),
// not once per process (a process that launches on a second GPU would otherwise fail with "invalid configuration").
// `done` is the caller's per-kernel flag array (one bit of state per device ordinal).
template <typename K>
inline cudaError_t ensure_dyn_smem(K kernel, int bytes, std::atomic<bool> (&done)[64]) {
  int dev = 0;
  cudaError_t e = cudaGetDevice(&dev);
  if (e != cudaSuccess) return e;
  if (dev >= 0 && dev < 64 && done[dev].load(std::memory_order_acquire)) return cudaSuccess;
  e = cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, bytes);
  if (e == cudaSuccess && dev >= 0 && dev < 64) done[dev].store(true, std::memory_order_release);
  return e;
}

// Programmatic dependent launch (PDL): the kernel may become resident while the previous kernel on
// the stream is still running; it must execute griddep_wait() before touching that kernel's output.
// Capturable in CUDA graphs (becomes a programmatic dependency edge).
template <typename... KArgs, typename... Args>
inline cudaError_t launch_pdl(void (*kernel)(KArgs...), dim3 grid, dim3 block, size_t smem, cudaStream_t st, Args... args) {
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = grid;
  cfg.blockDim = block;
  cfg.dynamicSmemBytes = smem;
  cfg.stream = st;
  cudaLaunchAttribute at[1];
  at[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  at[0].val.programmaticStreamSerializationAllowed = 1;
  cfg.attrs = at;
  cfg.numAttrs = 1;
  return cudaLaunchKernelEx(&cfg, kernel, KArgs(args)...);
}

// ---- device helpers ---------------------------------------------------------------------------
// PDL: block until the previous kernel on the stream has completed and its writes are visible.
__device__ __forceinline__ void griddep_wait() { asm volatile("griddepcontrol.wait;\n" ::: "memory"); }
// PDL: allow the next (programmatically dependent) kernel on the stream to start launching now.
__device__ __forceinline__ void griddep_launch_dependents() { asm volatile("griddepcontrol.launch_dependents;\n" ::: "memory"); }

__device__ __forceinline__ uint32_t smem_u32(const void* p) {
  return static_cast<uint32_t>(__cvta_generic_to_shared(p));
}

// 16-byte async copy global -> shared, L1 bypass; src_bytes == 0 zero-fills the destination.
__device__ __forceinline__ void cp_async_16(uint32_t dst, const void* src, uint32_t src_bytes) {
  asm volatile("cp.async.cg.shared.global [%0], [%1], 16, %2;\n" ::"r"(dst), "l"(src), "r"(src_bytes)
               : "memory");
}
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;\n" ::: "memory"); }
template <int N>
__device__ __forceinline__ void cp_async_wait() {
  asm volatile("cp.async.wait_group %0;\n" ::"n"(N) : "memory");
}

__device__ __forceinline__ void ldmatrix_x4(uint32_t (&r)[4], uint32_t addr) {
  asm volatile("ldmatrix.sync.aligned.m8n8.x4.shared.b16 {%0,%1,%2,%3}, [%4];\n"
               : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3])
               : "r"(addr));
}
__device__ __forceinline__ void ldmatrix_x4_trans(uint32_t (&r)[4], uint32_t addr) {
  asm volatile("ldmatrix.sync.aligned.m8n8.x4.trans.shared.b16 {%0,%1,%2,%3}, [%4];\n"
               : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3])
               : "r"(addr));
}

// D(16x8,f32) += A(16x16,bf16,row) * B(16x8,bf16,col)
__device__ __forceinline__ void mma_bf16_16816(float (&c)[4], const uint32_t (&a)[4], uint32_t b0,
                                               uint32_t b1) {
  asm volatile(
      "mma.sync.aligned.m16n8k16.row.col.f32.bf16.bf16.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, "
      "{%0,%1,%2,%3};\n"
      : "+f"(c[0]), "+f"(c[1]), "+f"(c[2]), "+f"(c[3])
      : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(b0), "r"(b1));
}

__device__ __forceinline__ uint32_t pack_bf16x2(float lo, float hi) {
  __nv_bfloat162 v = __floats2bfloat162_rn(lo, hi);
  return *reinterpret_cast<uint32_t*>(&v);
}

__device__ __forceinline__ float fast_exp2(float x) {
  float y;
  asm volatile("ex2.approx.ftz.f32 %0, %1;\n" : "=f"(y) : "f"(x));
  return y;
}

// exp2(x * s + b) as ONE ordered instruction pair: keeps ptxas from hoisting the FMAs of a whole softmax tile
// above unrelated code (observed: 15 spilled registers in the prefill softmax warps).
__device__ __forceinline__ float exp2_fma(float x, float s, float b) {
  float y;
  asm volatile("{\n\t.reg .f32 t;\n\tfma.rn.f32 t, %1, %2, %3;\n\tex2.approx.ftz.f32 %0, t;\n\t}\n" : "=f"(y) : "f"(x), "f"(s), "f"(b));
  return y;
}

// Single-instruction ordered (asm volatile) forms + a warp barrier as a scheduling fence: the prefill softmax lays
// out its FMA / MUFU / add / pack software pipeline with them (ptxas moves nothing across bar.warp.sync; left to
// itself it spends every FMA as filler for the first MUFUs and then issues ~50 MUFU.EX2 back to back).
__device__ __forceinline__ float fma_ordered(float x, float s, float b) {
  float y;
  asm volatile("fma.rn.f32 %0, %1, %2, %3;\n" : "=f"(y) : "f"(x), "f"(s), "f"(b));
  return y;
}
__device__ __forceinline__ float exp2_ordered(float x) {
  float y;
  asm volatile("ex2.approx.ftz.f32 %0, %1;\n" : "=f"(y) : "f"(x));
  return y;
}
__device__ __forceinline__ float add_ordered(float a, float b) {
  float y;
  asm volatile("add.rn.f32 %0, %1, %2;\n" : "=f"(y) : "f"(a), "f"(b));
  return y;
}
__device__ __forceinline__ uint32_t pack_bf16x2_ordered(float lo, float hi) {
  uint32_t y;
  asm volatile("cvt.rn.bf16x2.f32 %0, %1, %2;\n" : "=r"(y) : "f"(hi), "f"(lo));
  return y;
}
__device__ __forceinline__ void sched_fence() { asm volatile("bar.warp.sync 0xffffffff;\n" ::: "memory"); }

__device__ __forceinline__ float fast_log2(float x) {
  float y;
  asm volatile("lg2.approx.ftz.f32 %0, %1;\n" : "=f"(y) : "f"(x));
  return y;
}

constexpr float kLog2e = 1.4426950408889634f;

// KV length (in tokens) of sequence i from the FlashInfer-style page arrays.
__device__ __forceinline__ int seq_kv_len(const int32_t* __restrict__ kv_indptr,
                                          const int32_t* __restrict__ last_page_len, int i,
                                          int log2_ps) {
  const int pages = kv_indptr[i + 1] - kv_indptr[i];
  if (pages <= 0) return 0;
  return ((pages - 1) << log2_ps) + last_page_len[i];
}

// Block-wide inclusive scan of one int per thread (blockDim.x == NT, multiple of 32, <= 1024).
// `scratch` needs NT/32 + 1 ints.  Returns the inclusive prefix; *total receives the block sum.
template <int NT>
__device__ __forceinline__ int block_inclusive_scan(int v, int* scratch, int* total) {
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  int x = v;
#pragma unroll
  for (int o = 1; o < 32; o <<= 1) {
    int y = __shfl_up_sync(0xffffffffu, x, o);
    if (lane >= o) x += y;
  }
  __syncthreads();  // protect scratch reuse across calls
  if (lane == 31) scratch[warp] = x;
  __syncthreads();
  int base = 0;
#pragma unroll
  for (int w = 0; w < NT / 32; ++w) {
    int s = scratch[w];
    if (w < warp) base += s;
  }
  int tot = 0;
#pragma unroll
  for (int w = 0; w < NT / 32; ++w) tot += scratch[w];
  *total = tot;
  return x + base;
}

}  // namespace b200
