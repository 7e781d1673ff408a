// (f-3) Fused per-head q/k RMSNorm + rotary embedding + KV slot write: the step just ahead of
// Attention.forward.  Replaces, per layer, the reference's two Inductor RMSNorm kernels
// (layers/layernorm.py:16-36 via models/qwen3.py:125-132), the Inductor RoPE kernel
// (layers/rotary_embedding.py:73-102) and the Triton slot write (core/kv_cache.py:12-44 via
// layers/attention.py:347-352) by ONE kernel that reads the qkv GEMM output once and writes q (in place
// or to a separate buffer) and the K/V pool rows once.
//
// HBM bound.  Algorithmic bytes per token per layer: (Hq + 2 Hkv) * D * 2 read + the same written
// + 8 (position) + 8 (slot) ; the cos/sin row (D fp32) is shared by all heads of a token and stays in L1/L2.
//
// Arithmetic (fp32, one rounding per reference module output, which is what the reference's compiled
// modules produce on the GPU):
//   y  = bf16( (x * rsqrt(mean(x^2) + eps)) * w )                                layernorm.py:24-33
//   c, s = bf16(cos_sin_cache[pos][0:D/2]), bf16(cos_sin_cache[pos][D/2:D])      rotary_embedding.py:16-17, 92-93
//   o1 = y[0:D/2] * c - y[D/2:D] * s ;  o2 = y[D/2:D] * c + y[0:D/2] * s         :19-22
//   out[2i] = bf16(o1[i]), out[2i+1] = bf16(o2[i])     (INTERLEAVED, torch.stack(...,-1).flatten(-2))   :24
// Products of two bf16 values are exact in fp32, so the rotary part is bit-reproducible; only the
// reduction order of mean(x^2) and rsqrt's last ulp can differ from the CPU oracle.
//
// Mapping: 8 lanes own one (token, head) row of D = 128 elements.  Lane l holds x[8l..8l+7] (first half) and
// x[64+8l..64+8l+7] (second half) -> two 16-byte loads, and produces out[16l..16l+15] -> two 16-byte stores.
// A warp = 4 rows side by side; one warp work item = (token, batch of 4*U consecutive heads), so the position,
// slot id and the lane's cos/sin values (rounded to bf16 once, kept packed in 8 registers) are fetched once per
// 4*U rows and all U row loads are in flight before the first use.  V rows ride along as plain copies.
#include "common.cuh"

namespace b200 {

namespace {

// not .nc: q may be rewritten in place (q_out == qkv), each row by the lanes that read it
__device__ __forceinline__ int4 ld_16(const void* p) {
  int4 r;
  asm volatile("ld.global.L1::no_allocate.v4.s32 {%0,%1,%2,%3}, [%4];\n"
               : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w)
               : "l"(p)
               : "memory");
  return r;
}
__device__ __forceinline__ void st_16(void* p, const int4& v) {
  asm volatile("st.global.v4.s32 [%0], {%1,%2,%3,%4};\n" ::"l"(p), "r"(v.x), "r"(v.y), "r"(v.z), "r"(v.w) : "memory");
}

__device__ __forceinline__ float2 bf2_to_f2(uint32_t u) {
  return make_float2(__uint_as_float(u << 16), __uint_as_float(u & 0xffff0000u));
}
__device__ __forceinline__ float round_bf16(float x) { return __bfloat162float(__float2bfloat16_rn(x)); }

}  // namespace

constexpr int QK_LANES = 8;    // lanes per (token, head) row
constexpr int QK_SIDE = 4;     // rows side by side in a warp

template <int U>
__global__ void __launch_bounds__(256, 3) qk_norm_rope_store_kernel(
    const char* qkv, int64_t qkv_row_bytes, const int64_t* __restrict__ positions,
    const __nv_bfloat16* __restrict__ q_weight, const __nv_bfloat16* __restrict__ k_weight, float eps,
    const float* __restrict__ cos_sin, int64_t cos_sin_row_elems, int max_pos, char* q_out,
    int64_t q_out_row_bytes, char* __restrict__ k_out, int64_t k_out_row_bytes, char* __restrict__ k_cache,
    char* __restrict__ v_cache, int64_t cache_row_bytes, const int64_t* __restrict__ slots, int T, int Hq,
    int Hkv, int batches_per_token) {
  constexpr int D = 128, HB = D * 2;    // head row bytes
  const int lane = threadIdx.x & (QK_LANES - 1);
  const int side = (threadIdx.x & 31) / QK_LANES;
  const int heads = Hq + 2 * Hkv;
  const int items = T * batches_per_token;
  const int warps = (gridDim.x * blockDim.x) >> 5;

  // norm weights of this lane's 16 elements, packed bf16 pairs: [0..3] first half, [4..7] second half
  uint32_t wq[8], wk[8];
  {
    const int4 a = __ldg(reinterpret_cast<const int4*>(q_weight + 8 * lane));
    const int4 b = __ldg(reinterpret_cast<const int4*>(q_weight + 64 + 8 * lane));
    const int4 c = __ldg(reinterpret_cast<const int4*>(k_weight + 8 * lane));
    const int4 d = __ldg(reinterpret_cast<const int4*>(k_weight + 64 + 8 * lane));
    wq[0] = a.x; wq[1] = a.y; wq[2] = a.z; wq[3] = a.w; wq[4] = b.x; wq[5] = b.y; wq[6] = b.z; wq[7] = b.w;
    wk[0] = c.x; wk[1] = c.y; wk[2] = c.z; wk[3] = c.w; wk[4] = d.x; wk[5] = d.y; wk[6] = d.z; wk[7] = d.w;
  }

  for (int item = (blockIdx.x * blockDim.x + threadIdx.x) >> 5; item < items; item += warps) {   // warp-uniform
    const int t = item / batches_per_token;
    const int h0 = (item - t * batches_per_token) * (QK_SIDE * U) + side;
    const char* row = qkv + (int64_t)t * qkv_row_bytes;
    int4 lo[U], hi[U];
#pragma unroll
    for (int u = 0; u < U; ++u) {
      const int h = h0 + u * QK_SIDE;
      if (h < heads) {
        lo[u] = ld_16(row + (int64_t)h * HB + 16 * lane);
        hi[u] = ld_16(row + (int64_t)h * HB + 128 + 16 * lane);
      } else {
        lo[u] = make_int4(0, 0, 0, 0);
        hi[u] = lo[u];
      }
    }
    const int64_t slot = __ldg(slots + t);
    uint32_t cs[4], sn[4];      // bf16-rounded cos / sin of elements 8*lane .. 8*lane+7, packed pairs
    if (h0 < Hq + Hkv) {        // this lane has at least one q or k row in the batch
      int64_t p = __ldg(positions + t);
      p = p < 0 ? 0 : (p >= max_pos ? max_pos - 1 : p);   // the reference would fault; stay in bounds
      const float4* cr = reinterpret_cast<const float4*>(cos_sin + p * cos_sin_row_elems + 8 * lane);
      const float4 c0 = __ldg(cr), c1 = __ldg(cr + 1), s0 = __ldg(cr + 16), s1 = __ldg(cr + 17);
      cs[0] = pack_bf16x2(c0.x, c0.y); cs[1] = pack_bf16x2(c0.z, c0.w); cs[2] = pack_bf16x2(c1.x, c1.y); cs[3] = pack_bf16x2(c1.z, c1.w);
      sn[0] = pack_bf16x2(s0.x, s0.y); sn[1] = pack_bf16x2(s0.z, s0.w); sn[2] = pack_bf16x2(s1.x, s1.y); sn[3] = pack_bf16x2(s1.z, s1.w);
    }
#pragma unroll
    for (int u = 0; u < U; ++u) {
      const int h = h0 + u * QK_SIDE;
      const uint32_t xl[4] = {(uint32_t)lo[u].x, (uint32_t)lo[u].y, (uint32_t)lo[u].z, (uint32_t)lo[u].w};
      const uint32_t xh[4] = {(uint32_t)hi[u].x, (uint32_t)hi[u].y, (uint32_t)hi[u].z, (uint32_t)hi[u].w};
      float ss = 0.f;
#pragma unroll
      for (int i = 0; i < 4; ++i) {
        const float2 a = bf2_to_f2(xl[i]), b = bf2_to_f2(xh[i]);
        ss += a.x * a.x; ss += a.y * a.y; ss += b.x * b.x; ss += b.y * b.y;
      }
#pragma unroll
      for (int o = QK_LANES / 2; o > 0; o >>= 1) ss += __shfl_xor_sync(0xffffffffu, ss, o);   // whole warp is here
      if (h >= heads) continue;
      if (h >= Hq + Hkv) {          // V: plain copy into the pool row
        if (slot >= 0) {
          char* dst = v_cache + slot * cache_row_bytes + (int64_t)(h - Hq - Hkv) * HB;
          st_16(dst + 16 * lane, lo[u]);
          st_16(dst + 128 + 16 * lane, hi[u]);
        }
        continue;
      }
      const bool is_q = h < Hq;
      char* dst0 = is_q ? q_out + (int64_t)t * q_out_row_bytes + (int64_t)h * HB
                        : (slot >= 0 ? k_cache + slot * cache_row_bytes + (int64_t)(h - Hq) * HB : nullptr);
      char* dst1 = (!is_q && k_out) ? k_out + (int64_t)t * k_out_row_bytes + (int64_t)(h - Hq) * HB : nullptr;
      if (!dst0 && !dst1) continue;
      const float rs = rsqrtf(ss * (1.0f / D) + eps);
      uint32_t ow[8];
#pragma unroll
      for (int i = 0; i < 4; ++i) {
        const float2 a = bf2_to_f2(xl[i]), b = bf2_to_f2(xh[i]);
        const float2 wa = bf2_to_f2(is_q ? wq[i] : wk[i]), wb = bf2_to_f2(is_q ? wq[4 + i] : wk[4 + i]);
        const float2 c = bf2_to_f2(cs[i]), s = bf2_to_f2(sn[i]);
        const float y1x = round_bf16((a.x * rs) * wa.x), y1y = round_bf16((a.y * rs) * wa.y);
        const float y2x = round_bf16((b.x * rs) * wb.x), y2y = round_bf16((b.y * rs) * wb.y);
        // products of two bf16 values are exact in fp32: one rounding in the add, FMA or not
        ow[2 * i] = pack_bf16x2(y1x * c.x - y2x * s.x, y2x * c.x + y1x * s.x);
        ow[2 * i + 1] = pack_bf16x2(y1y * c.y - y2y * s.y, y2y * c.y + y1y * s.y);
      }
      const int4 o0 = make_int4((int)ow[0], (int)ow[1], (int)ow[2], (int)ow[3]);
      const int4 o1 = make_int4((int)ow[4], (int)ow[5], (int)ow[6], (int)ow[7]);
      if (dst0) { st_16(dst0 + 32 * lane, o0); st_16(dst0 + 32 * lane + 16, o1); }
      if (dst1) { st_16(dst1 + 32 * lane, o0); st_16(dst1 + 32 * lane + 16, o1); }
    }
  }
}

template <int U>
static int launch_qk(const void* qkv, int64_t qkv_row_stride, const int64_t* positions, const void* q_weight,
                     const void* k_weight, float eps, const float* cos_sin_cache, int64_t cos_sin_row_stride,
                     int32_t max_pos, void* q_out, int64_t q_out_row_stride, void* k_out, int64_t k_out_row_stride,
                     void* k_cache, void* v_cache, const int64_t* slots, int32_t T, int32_t Hq, int32_t Hkv,
                     cudaStream_t st) {
  constexpr int D = 128;
  const int heads = Hq + 2 * Hkv;
  const int bpt = (heads + QK_SIDE * U - 1) / (QK_SIDE * U);
  const int64_t items = (int64_t)T * bpt;          // one warp each
  int64_t blocks = (items + 7) / 8;
  const int64_t cap = (int64_t)cached_num_sms() * 8;
  if (blocks > cap) blocks = cap;
  if (blocks < 1) blocks = 1;
  qk_norm_rope_store_kernel<U><<<(unsigned)blocks, 256, 0, st>>>(
      (const char*)qkv, qkv_row_stride * 2, positions, (const __nv_bfloat16*)q_weight, (const __nv_bfloat16*)k_weight,
      eps, cos_sin_cache, cos_sin_row_stride, max_pos, (char*)q_out, q_out_row_stride * 2, (char*)k_out,
      k_out_row_stride * 2, (char*)k_cache, (char*)v_cache, (int64_t)Hkv * D * 2, slots, T, Hq, Hkv, bpt);
  return check_launch("b200_qk_norm_rope_store");
}

}  // namespace b200

extern "C" int b200_qk_norm_rope_store(const void* qkv, int64_t qkv_row_stride, const int64_t* positions,
                                       const void* q_weight, const void* k_weight, float eps,
                                       const float* cos_sin_cache, int64_t cos_sin_row_stride, int32_t max_pos,
                                       void* q_out, int64_t q_out_row_stride, void* k_out,
                                       int64_t k_out_row_stride, void* k_cache, void* v_cache,
                                       const int64_t* slots, int32_t T, int32_t Hq, int32_t Hkv, int32_t D,
                                       int32_t dtype, b200_stream_t stream) {
  using namespace b200;
  if (T < 0 || Hq <= 0 || Hkv <= 0) return fail("b200_qk_norm_rope_store: bad sizes T=%d Hq=%d Hkv=%d", T, Hq, Hkv);
  if (D != 128) return fail("b200_qk_norm_rope_store: only head_dim 128 is built (got %d)", D);
  if (dtype != B200_DTYPE_BF16) return fail("b200_qk_norm_rope_store: only bf16 is built");
  if (T == 0) return 0;
  if (!qkv || !positions || !q_weight || !k_weight || !cos_sin_cache || !q_out || !k_cache || !v_cache || !slots)
    return fail("b200_qk_norm_rope_store: null pointer");
  if (max_pos <= 0 || cos_sin_row_stride < D) return fail("b200_qk_norm_rope_store: bad cos/sin cache shape");
  const int64_t heads = (int64_t)Hq + 2 * Hkv;
  if (qkv_row_stride < heads * D || q_out_row_stride < (int64_t)Hq * D || (k_out && k_out_row_stride < (int64_t)Hkv * D))
    return fail("b200_qk_norm_rope_store: row stride shorter than a row");
  if ((int64_t)T * heads > 0x7fffffffll) return fail("b200_qk_norm_rope_store: T * heads overflows int32");
  const uintptr_t al = (uintptr_t)qkv | (uintptr_t)q_out | (uintptr_t)k_out | (uintptr_t)k_cache | (uintptr_t)v_cache |
                       (uintptr_t)cos_sin_cache | (uintptr_t)(qkv_row_stride * 2) | (uintptr_t)(q_out_row_stride * 2) |
                       (uintptr_t)(k_out_row_stride * 2) | (uintptr_t)(cos_sin_row_stride * 4);
  if (al & 15) return fail("b200_qk_norm_rope_store: pointers and row strides must be 16-byte aligned");
  if (((uintptr_t)q_weight | (uintptr_t)k_weight) & 15) return fail("b200_qk_norm_rope_store: weights must be 16-byte aligned");
  // rows per warp item = 4 * U: pick the U that wastes the fewest row slots for this head count
  auto waste = [&](int u) { const int b = 4 * u; return (int)((heads + b - 1) / b * b - heads); };
  int U = 3;     // 4 and 3 keep the most loads in flight; prefer them unless a smaller batch wastes fewer slots
  if (waste(4) < waste(U)) U = 4;
  if (waste(2) < waste(U)) U = 2;
  if (waste(1) < waste(U)) U = 1;
  cudaStream_t st = (cudaStream_t)stream;
#define B200_QK_ARGS qkv, qkv_row_stride, positions, q_weight, k_weight, eps, cos_sin_cache, cos_sin_row_stride, max_pos, \
                     q_out, q_out_row_stride, k_out, k_out_row_stride, k_cache, v_cache, slots, T, Hq, Hkv, st
  switch (U) {
    case 1: return launch_qk<1>(B200_QK_ARGS);
    case 2: return launch_qk<2>(B200_QK_ARGS);
    case 4: return launch_qk<4>(B200_QK_ARGS);
    default: return launch_qk<3>(B200_QK_ARGS);
  }
#undef B200_QK_ARGS
}
