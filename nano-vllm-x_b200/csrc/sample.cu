// (f-4) Sampler tail: temperature, top-k, top-p, min-p and the multinomial draw in ONE launch, without
// the reference's full-vocabulary sort (nanovllmx/layers/sampler.py:19-106; core/model_runner.py:191-219
// always sends all four tensors, so the reference sorts [B, V] every step).
//
// One CTA (1024 threads) per row.  The row is never sorted: thresholds are found by radix select over an
// order-preserving integer key of the raw logit (8-bit digits; 2 passes for bf16, 4 for fp32):
//   top-k : k-th largest key (count select); ties at the threshold are all kept (sampler.py:72 uses `<`)
//   top-p : ascending (key, index) order, remove the maximal prefix whose probability mass is <= 1 - p
//           (sampler.py:78-83).  Mass is accumulated in 2^-40 fixed point (uint64), so the result is
//           independent of the order in which threads add.  Ties at the cut are removed lowest index first
//           (what a stable ascending sort does), located by a count select over the index.
//   min-p : exp(x - max) < min_p (sampler.py:98-103; the renormalisation cancels)
// The draw is argmax(p_i / q_i) with q ~ Exp(1) — exactly what torch.multinomial(probs, 1) computes — evaluated
// as argmax(x_i - log q_i), lowest index first on ties.  q comes either from a caller-provided noise tensor (the
// values torch.multinomial would draw: parity mode) or from Philox-4x32-10 keyed by (seed, offset, row, column).
// A row with all parameter arrays absent, or temperature <= 0, is greedy: argmax of the raw logits.
#include "common.cuh"

namespace b200 {
namespace {

constexpr int kSampleThreads = 1024;

struct SampleParams {
  const void* logits;
  int64_t row_stride;
  int B, V;
  const float* temperatures;
  const int64_t* top_ks;
  const float* top_ps;
  const float* min_ps;
  const float* noise;
  int64_t noise_row_stride;
  uint64_t seed, offset;
  int64_t* out;
  int limb_bits;       // histogram limb width, see radix_select
  float fix_scale;     // 2^(2 * limb_bits - 1): fixed-point scale of a probability mass
};

template <typename T> struct KeyTraits;
template <> struct KeyTraits<__nv_bfloat16> {
  static constexpr int kBits = 16;
  // order-preserving: negative values flip all bits, positive values flip the sign bit
  __device__ static uint32_t key(__nv_bfloat16 v) {
    uint32_t u = __bfloat16_as_ushort(v);
    return (u & 0x8000u) ? (~u & 0xFFFFu) : (u | 0x8000u);
  }
  __device__ static float value(uint32_t k) {
    uint32_t u = (k & 0x8000u) ? (k & 0x7FFFu) : (~k & 0xFFFFu);
    return __bfloat162float(__ushort_as_bfloat16((unsigned short)u));
  }
};
template <> struct KeyTraits<float> {
  static constexpr int kBits = 32;
  __device__ static uint32_t key(float v) {
    uint32_t u = __float_as_uint(v);
    return (u & 0x80000000u) ? ~u : (u | 0x80000000u);
  }
  __device__ static float value(uint32_t k) {
    uint32_t u = (k & 0x80000000u) ? (k & 0x7FFFFFFFu) : ~k;
    return __uint_as_float(u);
  }
};

__device__ __forceinline__ float to_float(__nv_bfloat16 v) { return __bfloat162float(v); }
__device__ __forceinline__ float to_float(float v) { return v; }

// ---- Philox-4x32-10 ------------------------------------------------------------------------------
__device__ __forceinline__ uint4 philox4x32_10(uint4 c, uint2 k) {
#pragma unroll
  for (int r = 0; r < 10; ++r) {
    const uint32_t hi0 = __umulhi(0xD2511F53u, c.x), lo0 = 0xD2511F53u * c.x;
    const uint32_t hi1 = __umulhi(0xCD9E8D57u, c.z), lo1 = 0xCD9E8D57u * c.z;
    c = make_uint4(hi1 ^ c.y ^ k.x, lo1, hi0 ^ c.w ^ k.y, lo0);
    k.x += 0x9E3779B9u;
    k.y += 0xBB67AE85u;
  }
  return c;
}
__device__ __forceinline__ float word_to_uniform(uint32_t w) {   // (0, 1), 24 bits
  return (static_cast<float>(w >> 8) + 0.5f) * 5.9604644775390625e-08f;
}

// ---- block reductions -----------------------------------------------------------------------------
// max over (a, b) pairs compared lexicographically; every thread receives the result
__device__ __forceinline__ void block_max_pair(uint64_t& v, uint64_t* scratch) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
    const uint64_t y = __shfl_xor_sync(0xffffffffu, v, o);
    v = y > v ? y : v;
  }
  __syncthreads();
  if ((threadIdx.x & 31) == 0) scratch[threadIdx.x >> 5] = v;
  __syncthreads();
  uint64_t m = scratch[0];
#pragma unroll
  for (int w = 1; w < kSampleThreads / 32; ++w) m = scratch[w] > m ? scratch[w] : m;
  v = m;
}

// ---- radix select -----------------------------------------------------------------------------------
// Histograms use NATIVE 32-bit shared atomics only (a 64-bit shared atomicAdd compiles to a CAS spin loop) and are
// striped by lane: cell (digit, lane) is touched by that lane alone within a warp, so neither bank conflicts nor
// same-address serialisation occur however concentrated the digits are (the exponent byte of bf16 logits is).
// A 64-bit weight is split into two limbs of `limb_bits` bits, each summed in its own 32-bit cell; a cell receives
// at most ceil(V / 32) addends, so limb_bits = 32 - ceil(log2(ceil(V / 32))) (capped at 20) cannot overflow.
struct SelectResult {
  uint32_t key;      // smallest key K with W(key <= K) > target; the largest present key if there is none
  uint64_t below;    // W(key < K)
  uint64_t equal;    // W(key == K)
  uint64_t total;    // W(all active elements)
};

struct SelectSmem {
  uint64_t tot[256];
  uint64_t warp_tot[8];
  uint64_t red[kSampleThreads / 32];
  uint64_t below, equal, total;
  int digit, last_nonempty;
};

constexpr int kHistWords = 2 * 256 * 32;     // [limb][digit][lane] uint32 = 64 KiB of dynamic shared memory

// Ascending weighted select over the elements for which elem(i, key, weight) returns true.  BITS = key width
// (multiple of 8).  The target is target_of(total weight), evaluated after the first pass, so a fraction of the
// total needs no pass of its own.  Every thread of the CTA must call it; every thread receives the result.
template <int BITS, bool WEIGHTED, typename F, typename G>
__device__ SelectResult radix_select(int V, int limb_bits, F elem, G target_of, SelectSmem& sm, uint32_t* hist) {
  uint32_t prefix = 0;                       // digits chosen so far, right-aligned
  uint64_t below_total = 0, eq = 0, total = 0, target = 0;
  const int lane = threadIdx.x & 31;
  const uint32_t limb_mask = (1u << limb_bits) - 1u;
#pragma unroll 1
  for (int shift = BITS - 8; shift >= 0; shift -= 8) {
    for (int t = threadIdx.x; t < (WEIGHTED ? kHistWords : kHistWords / 2); t += kSampleThreads) hist[t] = 0;
    if (threadIdx.x == 0) { sm.digit = -1; sm.last_nonempty = -1; }
    __syncthreads();
    for (int i = threadIdx.x; i < V; i += kSampleThreads) {
      uint32_t key; uint64_t w;
      if (!elem(i, key, w)) continue;
      if (shift + 8 < BITS && (key >> (shift + 8)) != prefix) continue;
      const int cell = (((key >> shift) & 0xFF) << 5) | lane;
      if (WEIGHTED) {
        atomicAdd(&hist[cell], (uint32_t)w & limb_mask);
        atomicAdd(&hist[256 * 32 + cell], (uint32_t)(w >> limb_bits));
      } else {
        atomicAdd(&hist[cell], 1u);
      }
    }
    __syncthreads();
    {   // bin totals: thread t sums a rotated quarter of the lanes of bin t / 4 (bank-conflict free)
      const int bin = threadIdx.x >> 2, q = threadIdx.x & 3;
      uint64_t lo = 0, hi = 0;
#pragma unroll
      for (int j = 0; j < 8; ++j) {
        const int l = (q * 8 + j + bin) & 31;
        lo += hist[(bin << 5) | l];
        if (WEIGHTED) hi += hist[256 * 32 + ((bin << 5) | l)];
      }
      uint64_t t = lo + (hi << limb_bits);
      t += __shfl_xor_sync(0xffffffffu, t, 1);
      t += __shfl_xor_sync(0xffffffffu, t, 2);
      if (q == 0) sm.tot[bin] = t;
    }
    __syncthreads();
    // 256-bin inclusive scan by the first 8 warps, then locate the crossing bin
    uint64_t tot = 0, incl = 0;
    if (threadIdx.x < 256) {
      tot = sm.tot[threadIdx.x];
      incl = tot;
#pragma unroll
      for (int o = 1; o < 32; o <<= 1) {
        const uint64_t y = __shfl_up_sync(0xffffffffu, incl, o);
        if (lane >= o) incl += y;
      }
      if (lane == 31) sm.warp_tot[threadIdx.x >> 5] = incl;
    }
    __syncthreads();
    if (shift == BITS - 8) {                  // first pass: the grand total fixes the target
      uint64_t all = 0;
#pragma unroll
      for (int w = 0; w < 8; ++w) all += sm.warp_tot[w];
      total = all;
      target = target_of(all);
    }
    if (threadIdx.x < 256) {
      for (int w = 0; w < (int)(threadIdx.x >> 5); ++w) incl += sm.warp_tot[w];
      if (tot > 0) atomicMax(&sm.last_nonempty, (int)threadIdx.x);
      if (incl > target && incl - tot <= target) { sm.digit = threadIdx.x; sm.below = incl - tot; sm.equal = tot; }
    }
    __syncthreads();
    if (sm.digit < 0) {
      // no crossing: everything weighs <= target; settle on the largest present key
      if ((int)threadIdx.x == sm.last_nonempty) { sm.below = incl - tot; sm.equal = tot; }
      __syncthreads();
      if (threadIdx.x == 0) {
        if (sm.last_nonempty < 0) { sm.below = 0; sm.equal = 0; }
        sm.digit = sm.last_nonempty < 0 ? 0 : sm.last_nonempty;
      }
      target = ~0ull;
      __syncthreads();
    }
    const int d = sm.digit;
    prefix = (prefix << 8) | (uint32_t)d;
    below_total += sm.below;
    if (target != ~0ull) target -= sm.below;
    eq = sm.equal;
    __syncthreads();
  }
  return SelectResult{prefix, below_total, eq, total};
}

// ---- the kernel ---------------------------------------------------------------------------------------
constexpr float kLn2 = 0.6931471805599453f;

// raw logits of columns 4g .. 4g+3 (one 8- / 16-byte load when the row allows it)
template <typename T>
__device__ __forceinline__ void load4(const T* __restrict__ x, int g, int V, bool vec, float (&r)[4]) {
  if (vec && g * 4 + 3 < V) {
    if constexpr (sizeof(T) == 2) {
      const uint2 u = *reinterpret_cast<const uint2*>(x + g * 4);
      const __nv_bfloat162 a = *reinterpret_cast<const __nv_bfloat162*>(&u.x), b = *reinterpret_cast<const __nv_bfloat162*>(&u.y);
      r[0] = __low2float(a); r[1] = __high2float(a); r[2] = __low2float(b); r[3] = __high2float(b);
    } else {
      const float4 f = *reinterpret_cast<const float4*>(x + g * 4);
      r[0] = f.x; r[1] = f.y; r[2] = f.z; r[3] = f.w;
    }
  } else {
#pragma unroll
    for (int c = 0; c < 4; ++c) r[c] = g * 4 + c < V ? to_float(x[g * 4 + c]) : -INFINITY;
  }
}

struct Cut {            // what survives truncation: key >= kth, (key, index) >= (cut_key, cut_idx), e >= min_p
  uint32_t kth_key, cut_key;
  int cut_idx;
  float min_p, maxv, escale;
};

// The draw: argmax_i(x_i / T - ln q_i) over the survivors, lowest index first (sampler.py:48-53).
template <typename T, bool TRUNC>
__device__ __forceinline__ uint64_t draw(const T* __restrict__ x, int V, bool vec, float inv_t, const float* __restrict__ nz,
                                         uint2 pkey, uint32_t row, uint64_t offset, const Cut& c) {
  using KT = KeyTraits<T>;
  float best_s = -INFINITY;
  int best_i = 0x7FFFFFFF;
  float nxt[4];
  load4(x, (int)threadIdx.x, V, vec, nxt);
  for (int g = threadIdx.x; g * 4 < V; g += kSampleThreads) {
    float raw[4];
#pragma unroll
    for (int j = 0; j < 4; ++j) raw[j] = nxt[j];
    load4(x, g + kSampleThreads, V, vec, nxt);     // next group's logits are in flight while this one is scored
    float lq[4];                                   // log2 of the Exp(1) noise
    if (nz) {
#pragma unroll
      for (int j = 0; j < 4; ++j) lq[j] = g * 4 + j < V ? fast_log2(nz[g * 4 + j]) : 0.f;
    } else {
      const uint4 rnd = philox4x32_10(make_uint4((uint32_t)g, row, (uint32_t)offset, (uint32_t)(offset >> 32)), pkey);
      const uint32_t w[4] = {rnd.x, rnd.y, rnd.z, rnd.w};
#pragma unroll
      for (int j = 0; j < 4; ++j) lq[j] = fast_log2(-kLn2 * fast_log2(word_to_uniform(w[j])));   // q = -ln u
    }
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      const int i = g * 4 + j;
      float s = fmaf(raw[j], inv_t, -kLn2 * lq[j]);
      if (TRUNC) {
        const uint32_t key = KT::key(i < V ? x[i] : x[0]);
        bool keep = key >= c.kth_key && (key > c.cut_key || (key == c.cut_key && i >= c.cut_idx));
        keep = keep && !(fast_exp2((raw[j] - c.maxv) * c.escale) < c.min_p);
        s = keep ? s : -INFINITY;
      }
      if (i < V && (s > best_s || best_i == 0x7FFFFFFF)) { best_s = s; best_i = i; }
    }
  }
  if (best_i == 0x7FFFFFFF) return 0;
  return ((uint64_t)KeyTraits<float>::key(best_s) << 32) | (uint32_t)(0xFFFFFFFFu - (uint32_t)best_i);
}

template <typename T>
__global__ void __launch_bounds__(kSampleThreads, 1) sample_kernel(SampleParams p) {
  using KT = KeyTraits<T>;
  __shared__ SelectSmem sm;
  extern __shared__ uint32_t hist[];
  const int row = blockIdx.x;
  const int V = p.V;
  const T* __restrict__ x = reinterpret_cast<const T*>(p.logits) + (int64_t)row * p.row_stride;
  const bool vec = (reinterpret_cast<uintptr_t>(x) & (4 * sizeof(T) - 1)) == 0;

  const bool any_param = p.temperatures || p.top_ks || p.top_ps || p.min_ps;
  const float temp = p.temperatures ? p.temperatures[row] : 1.0f;
  const bool greedy = !any_param || !(temp > 0.0f);
  int64_t k = p.top_ks ? p.top_ks[row] : (int64_t)V;
  k = k < 1 ? 1 : k;
  const float top_p = p.top_ps ? p.top_ps[row] : 1.0f;
  const float min_p = p.min_ps ? p.min_ps[row] : 0.0f;
  const bool need_k = k < (int64_t)V, need_p = top_p < 1.0f, need_minp = min_p > 0.0f;
  const float inv_t = 1.0f / temp;

  // row maximum (key, lowest index): the greedy answer, and the softmax shift of top-p / min-p
  uint32_t max_key = 0;
  if (greedy || need_p || need_minp) {
    uint32_t bk = 0;
    int bi = 0;
    for (int g = threadIdx.x; g * 4 < V; g += kSampleThreads) {
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        const int i = g * 4 + j;
        if (i < V) {
          const uint32_t key = KT::key(x[i]);
          if (key > bk || (g == (int)threadIdx.x && j == 0)) { bk = key; bi = i; }
        }
      }
    }
    uint64_t best = threadIdx.x * 4 < V ? (((uint64_t)bk << 32) | (uint32_t)(0xFFFFFFFFu - (uint32_t)bi)) : 0ull;
    block_max_pair(best, sm.red);
    if (greedy) {                                           // sampler.py:28-36
      if (threadIdx.x == 0) p.out[row] = (int64_t)(0xFFFFFFFFu - (uint32_t)best);
      return;
    }
    max_key = (uint32_t)(best >> 32);
  }
  Cut c{0u, 0u, 0, need_minp ? min_p : -1.0f, KT::value(max_key), inv_t * kLog2e};
  // probability mass of element i relative to the row maximum, 2^-40 fixed point
  auto mass = [&](int i) { return (uint64_t)__float2ull_rd(fast_exp2((to_float(x[i]) - c.maxv) * c.escale) * p.fix_scale); };

  // top-k: the k-th largest key = the k-th smallest inverted key (sampler.py:67-73)
  if (need_k) {
    constexpr uint32_t kMask = KT::kBits == 32 ? 0xFFFFFFFFu : ((1u << (KT::kBits & 31)) - 1u);
    auto elem = [&](int i, uint32_t& key, uint64_t& w) { key = ~KT::key(x[i]) & kMask; w = 1; return true; };
    auto tgt = [&](uint64_t) { return (uint64_t)(k - 1); };
    c.kth_key = ~radix_select<KT::kBits, false>(V, 0, elem, tgt, sm, hist).key & kMask;
  }

  // top-p over what top-k kept (sampler.py:75-83)
  if (need_p) {
    // removed prefix = mass <= (1 - p) * Z, Z = mass of the top-k survivors = the first pass's grand total
    uint64_t target = 0;
    auto tgt = [&](uint64_t z) {
      const double t = (double)(1.0f - top_p) * (double)z;
      target = t <= 0.0 ? 0ull : (uint64_t)t;
      return target;
    };
    auto elem = [&](int i, uint32_t& key, uint64_t& w) {
      key = KT::key(x[i]);
      if (key < c.kth_key) return false;
      w = mass(i);
      return true;
    };
    const SelectResult r = radix_select<KT::kBits, true>(V, p.limb_bits, elem, tgt, sm, hist);
    c.cut_key = r.key;
    const uint64_t e_v = (uint64_t)__float2ull_rd(fast_exp2((KT::value(r.key) - c.maxv) * c.escale) * p.fix_scale);
    uint64_t removed_ties = 0;
    if (e_v > 0) {
      const uint64_t ties = r.equal / e_v;
      if (r.below + r.equal > target) removed_ties = (target - r.below) / e_v;   // < ties by construction
      else removed_ties = ties > 0 ? ties - 1 : 0;                                // sampler.py:82 keeps the last one
    }
    if (removed_ties > 0) {
      // ties at the cut leave lowest index first: the index of the (removed_ties + 1)-th tie is the first kept
      auto tie = [&](int i, uint32_t& key, uint64_t& w) { key = (uint32_t)i; w = 1; return KT::key(x[i]) == r.key; };
      auto tgt2 = [&](uint64_t) { return removed_ties; };
      c.cut_idx = (int)radix_select<24, false>(V, 0, tie, tgt2, sm, hist).key;
    }
  }

  const float* __restrict__ nz = p.noise ? p.noise + (int64_t)row * p.noise_row_stride : nullptr;
  const uint2 pkey = make_uint2((uint32_t)p.seed, (uint32_t)(p.seed >> 32));
  uint64_t win = (need_k || need_p || need_minp)
                     ? draw<T, true>(x, V, vec, inv_t, nz, pkey, (uint32_t)row, p.offset, c)
                     : draw<T, false>(x, V, vec, inv_t, nz, pkey, (uint32_t)row, p.offset, c);
  block_max_pair(win, sm.red);
  if (threadIdx.x == 0) p.out[row] = (int64_t)(0xFFFFFFFFu - (uint32_t)win);
}

}  // namespace
}  // namespace b200

extern "C" int b200_sample(const void* logits, int64_t row_stride, int32_t logits_dtype, int32_t B, int32_t V,
                           const float* temperatures, const int64_t* top_ks, const float* top_ps, const float* min_ps,
                           const float* noise, int64_t noise_row_stride, uint64_t seed, uint64_t offset,
                           int64_t* out_ids, b200_stream_t stream) {
  using namespace b200;
  if (B < 0 || V < 1 || V > (1 << 24)) return fail("b200_sample: bad shape B=%d V=%d (V <= 2^24)", B, V);
  if (B == 0) return 0;
  if (!logits || !out_ids) return fail("b200_sample: null pointer");
  if (row_stride < V || (noise && noise_row_stride < V)) return fail("b200_sample: row stride smaller than V");
  if (logits_dtype != B200_DTYPE_BF16 && logits_dtype != B200_DTYPE_FP32)
    return fail("b200_sample: logits dtype %d not supported (bf16, fp32)", logits_dtype);
  int per_cell = (V + 31) / 32, lg = 0;
  while ((1 << lg) < per_cell) ++lg;
  const int limb_bits = 32 - lg < 20 ? 32 - lg : 20;
  SampleParams p{logits, row_stride, B, V, temperatures, top_ks, top_ps, min_ps, noise, noise_row_stride, seed, offset, out_ids,
                 limb_bits, ldexpf(1.0f, 2 * limb_bits - 1)};
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  const size_t smem = kHistWords * sizeof(uint32_t);
  {   // per device, cheap, not a stream operation (legal during graph capture)
    if (cudaFuncSetAttribute(sample_kernel<__nv_bfloat16>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem) != cudaSuccess ||
        cudaFuncSetAttribute(sample_kernel<float>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem) != cudaSuccess) {
      (void)cudaGetLastError();
      return fail("b200_sample: cannot opt in to %zu bytes of shared memory", smem);
    }
  }
  if (logits_dtype == B200_DTYPE_BF16) sample_kernel<__nv_bfloat16><<<B, kSampleThreads, smem, st>>>(p);
  else sample_kernel<float><<<B, kSampleThreads, smem, st>>>(p);
  return check_launch("b200_sample");
}
