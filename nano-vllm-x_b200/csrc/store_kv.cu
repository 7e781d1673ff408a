// (1) KV slot write: k_cache[slots[t]] = k[t], v_cache[slots[t]] = v[t], slot < 0 skipped.
// Replaces core/kv_cache.py:12-44 (Triton, one program per token, power-of-two row only).
//
// HBM-bound byte copy.  Each thread moves VEC-byte vectors; UNROLL independent loads are issued
// before the stores (MLP), sources go through the read-only path with no L1 allocation, and the
// slot id is a broadcast load shared by all threads of a row.  Rows are raw bytes, so the copy
// is bit-exact for any 2-byte dtype.
#include "common.cuh"

namespace b200 {

template <typename V>
struct VecIO;
template <>
struct VecIO<int4> {
  static __device__ __forceinline__ int4 ld(const void* p) {
    int4 r;
    asm volatile("ld.global.nc.L1::no_allocate.v4.s32 {%0,%1,%2,%3}, [%4];\n"
                 : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w)
                 : "l"(p));
    return r;
  }
  static __device__ __forceinline__ void st(void* p, const int4& v) {
    asm volatile("st.global.v4.s32 [%0], {%1,%2,%3,%4};\n" ::"l"(p), "r"(v.x), "r"(v.y), "r"(v.z),
                 "r"(v.w)
                 : "memory");
  }
};
template <>
struct VecIO<int2> {
  static __device__ __forceinline__ int2 ld(const void* p) { return __ldg(reinterpret_cast<const int2*>(p)); }
  static __device__ __forceinline__ void st(void* p, const int2& v) { *reinterpret_cast<int2*>(p) = v; }
};
template <>
struct VecIO<int> {
  static __device__ __forceinline__ int ld(const void* p) { return __ldg(reinterpret_cast<const int*>(p)); }
  static __device__ __forceinline__ void st(void* p, const int& v) { *reinterpret_cast<int*>(p) = v; }
};
template <>
struct VecIO<short> {
  static __device__ __forceinline__ short ld(const void* p) { return __ldg(reinterpret_cast<const short*>(p)); }
  static __device__ __forceinline__ void st(void* p, const short& v) { *reinterpret_cast<short*>(p) = v; }
};

// One work item = one vector of one row of K or V.  Items are laid out so that consecutive
// threads touch consecutive vectors of the same row (coalesced on both sides).
template <typename V, int UNROLL>
__global__ void __launch_bounds__(256) store_kv_kernel(const char* __restrict__ k, int64_t k_row_bytes,
                                                       const char* __restrict__ v, int64_t v_row_bytes,
                                                       char* __restrict__ k_cache, char* __restrict__ v_cache,
                                                       const int64_t* __restrict__ slots, int T,
                                                       int vec_per_row, int64_t cache_row_bytes) {
  const int64_t items_per_tensor = (int64_t)T * vec_per_row;
  const int64_t total = 2 * items_per_tensor;
  const int64_t stride = (int64_t)gridDim.x * blockDim.x;
  for (int64_t base = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; base < total; base += stride * UNROLL) {
    V val[UNROLL];
    char* dst[UNROLL];
#pragma unroll
    for (int u = 0; u < UNROLL; ++u) {
      const int64_t e = base + (int64_t)u * stride;
      dst[u] = nullptr;
      if (e < total) {
        const bool is_v = e >= items_per_tensor;
        const int64_t r = is_v ? e - items_per_tensor : e;
        const int t = (int)(r / vec_per_row);
        const int c = (int)(r - (int64_t)t * vec_per_row);
        const int64_t slot = __ldg(slots + t);
        if (slot >= 0) {
          const char* src = (is_v ? v + t * v_row_bytes : k + t * k_row_bytes) + (int64_t)c * sizeof(V);
          val[u] = VecIO<V>::ld(src);
          dst[u] = (is_v ? v_cache : k_cache) + slot * cache_row_bytes + (int64_t)c * sizeof(V);
        }
      }
    }
#pragma unroll
    for (int u = 0; u < UNROLL; ++u)
      if (dst[u]) VecIO<V>::st(dst[u], val[u]);
  }
}

// Row-cooperative variant for the common row sizes: LANES threads own one (token, K|V) row, each moves VPL
// 16-byte vectors of it, and ROWS rows are in flight per thread group (VPL*ROWS == 4 independent loads before
// the first store).  No integer division on the hot path (the flat kernel above needs a 64-bit div per item).
template <int LANES, int VPL, int ROWS>
__global__ void __launch_bounds__(256) store_kv_rows_kernel(const char* __restrict__ k, int64_t k_row_bytes,
                                                            const char* __restrict__ v, int64_t v_row_bytes,
                                                            char* __restrict__ k_cache, char* __restrict__ v_cache,
                                                            const int64_t* __restrict__ slots, int T,
                                                            int64_t cache_row_bytes) {
  const int lane = threadIdx.x % LANES;
  const int groups = (gridDim.x * blockDim.x) / LANES;
  const int total_rows = 2 * T;
  for (int row0 = (blockIdx.x * blockDim.x + threadIdx.x) / LANES; row0 < total_rows; row0 += groups * ROWS) {
    int4 val[ROWS][VPL];
    char* dst[ROWS];
#pragma unroll
    for (int r = 0; r < ROWS; ++r) {
      const int row = row0 + r * groups;
      dst[r] = nullptr;
      if (row < total_rows) {
        const bool is_v = row >= T;
        const int t = is_v ? row - T : row;
        const int64_t slot = __ldg(slots + t);
        if (slot >= 0) {
          const char* src = (is_v ? v + t * v_row_bytes : k + t * k_row_bytes) + lane * 16;
#pragma unroll
          for (int c = 0; c < VPL; ++c) val[r][c] = VecIO<int4>::ld(src + c * LANES * 16);
          dst[r] = (is_v ? v_cache : k_cache) + slot * cache_row_bytes + lane * 16;
        }
      }
    }
#pragma unroll
    for (int r = 0; r < ROWS; ++r)
      if (dst[r]) {
#pragma unroll
        for (int c = 0; c < VPL; ++c) VecIO<int4>::st(dst[r] + c * LANES * 16, val[r][c]);
      }
  }
}

template <int LANES, int VPL, int ROWS>
static int launch_store_rows(const void* k, int64_t kb, const void* v, int64_t vb, void* k_cache, void* v_cache,
                             const int64_t* slots, int T, int64_t row_bytes, cudaStream_t st) {
  const int64_t groups_needed = (2ll * T + ROWS - 1) / ROWS;
  int64_t blocks = (groups_needed * LANES + 255) / 256;
  const int64_t cap = (int64_t)cached_num_sms() * 8;
  if (blocks > cap) blocks = cap;
  if (blocks < 1) blocks = 1;
  store_kv_rows_kernel<LANES, VPL, ROWS><<<(unsigned)blocks, 256, 0, st>>>(
      (const char*)k, kb, (const char*)v, vb, (char*)k_cache, (char*)v_cache, slots, T, row_bytes);
  return check_launch("b200_store_kv");
}

template <typename V>
static int launch_store(const void* k, int64_t k_row_bytes, const void* v, int64_t v_row_bytes, void* k_cache,
                        void* v_cache, const int64_t* slots, int T, int64_t row_bytes, cudaStream_t st) {
  constexpr int UNROLL = 4;
  const int vec_per_row = (int)(row_bytes / sizeof(V));
  const int64_t total = 2ll * T * vec_per_row;
  const int threads = 256;
  int64_t blocks = (total + (int64_t)threads * UNROLL - 1) / ((int64_t)threads * UNROLL);
  const int64_t cap = (int64_t)cached_num_sms() * 16;  // grid-stride beyond 16 CTAs per SM
  if (blocks > cap) blocks = cap;
  if (blocks < 1) blocks = 1;
  store_kv_kernel<V, UNROLL><<<(unsigned)blocks, threads, 0, st>>>(
      (const char*)k, k_row_bytes, (const char*)v, v_row_bytes, (char*)k_cache, (char*)v_cache, slots, T,
      vec_per_row, row_bytes);
  return check_launch("b200_store_kv");
}

}  // namespace b200

extern "C" int b200_store_kv(const void* k, int64_t k_row_stride, const void* v, int64_t v_row_stride,
                             void* k_cache, void* v_cache, const int64_t* slots, int32_t T, int32_t row_elems,
                             int32_t elem_bytes, b200_stream_t stream) {
  using namespace b200;
  if (T < 0 || row_elems <= 0 || elem_bytes <= 0) return fail("b200_store_kv: bad sizes T=%d row_elems=%d", T, row_elems);
  if (T == 0) return 0;
  if (!k || !v || !k_cache || !v_cache || !slots) return fail("b200_store_kv: null pointer");
  // stride(-1)==1 and stride(-2)==head_dim are the caller's contract (core/kv_cache.py:39-41);
  // what can be checked here is that a row stride is not shorter than a row.
  if (k_row_stride < row_elems || v_row_stride < row_elems)
    return fail("b200_store_kv: row stride (%lld,%lld) < row_elems %d", (long long)k_row_stride,
                (long long)v_row_stride, row_elems);
  const int64_t row_bytes = (int64_t)row_elems * elem_bytes;
  const int64_t kb = k_row_stride * elem_bytes, vb = v_row_stride * elem_bytes;
  cudaStream_t st = (cudaStream_t)stream;
  const uintptr_t align_or = (uintptr_t)k | (uintptr_t)v | (uintptr_t)k_cache | (uintptr_t)v_cache |
                             (uintptr_t)row_bytes | (uintptr_t)kb | (uintptr_t)vb;
  if ((align_or & 15) == 0) {
    switch (row_bytes) {   // Hkv_local * 128 * 2 B: 2 KiB (8 heads) ... 256 B (1 head)
      case 2048: return launch_store_rows<32, 4, 1>(k, kb, v, vb, k_cache, v_cache, slots, T, row_bytes, st);
      case 1024: return launch_store_rows<32, 2, 2>(k, kb, v, vb, k_cache, v_cache, slots, T, row_bytes, st);
      case 512: return launch_store_rows<32, 1, 4>(k, kb, v, vb, k_cache, v_cache, slots, T, row_bytes, st);
      case 256: return launch_store_rows<16, 1, 4>(k, kb, v, vb, k_cache, v_cache, slots, T, row_bytes, st);
      default: return launch_store<int4>(k, kb, v, vb, k_cache, v_cache, slots, T, row_bytes, st);
    }
  }
  if ((align_or & 7) == 0) return launch_store<int2>(k, kb, v, vb, k_cache, v_cache, slots, T, row_bytes, st);
  if ((align_or & 3) == 0) return launch_store<int>(k, kb, v, vb, k_cache, v_cache, slots, T, row_bytes, st);
  if ((align_or & 1) == 0) return launch_store<short>(k, kb, v, vb, k_cache, v_cache, slots, T, row_bytes, st);
  return fail("b200_store_kv: rows are not 2-byte aligned");
}
