// (4) Ragged paged prefill attention — warp-level mma.sync formulation.
// Replaces layers/attention.py:360-364 (FlashInfer BatchPrefillWithPagedKVCacheWrapper).
//
// This is the first, simple-and-correct prefill path (round 1 step "a first correct CUDA path"):
// one CTA = 64 query rows of one (sequence, query head); 4 warps x 16 rows; K/V tiles of 64
// tokens gathered through the page table with 16-byte cp.async into XOR-swizzled shared memory,
// double buffered; q.K^T and p.V on mma.sync m16n8k16 with fp32 online softmax; bottom-right
// aligned causal mask (query j sees kv positions <= Lkv - Lq + j).  The tcgen05/TMEM kernel
// (prefill_sm100.cu) supersedes it for the shapes it covers.

#include "common.cuh"
#include "prefill_params.cuh"

namespace b200 {

constexpr int kPD = 128;
constexpr int kPRowBytes = kPD * 2;
constexpr int kQTile = 64;
constexpr int kKVTile = 64;
constexpr int kPWarps = 4;
constexpr int kQBytes = kQTile * kPRowBytes;        // 16 KiB
constexpr int kKVBytes = kKVTile * kPRowBytes;      // 16 KiB per K or V tile
constexpr int kPStages = 2;
constexpr int kPrefillSmem = kQBytes + kPStages * 2 * kKVBytes;  // 80 KiB


__global__ void __launch_bounds__(kPWarps * 32, 2) prefill_mma_kernel(const PrefillParams p) {
  extern __shared__ __align__(1024) uint8_t smem[];
  __shared__ int s_scan[kPWarps + 1];
  __shared__ int s_loc[2];
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const int hq = blockIdx.y, kv_head = hq / p.group;

  // ---- locate (sequence, q tile) for blockIdx.x -------------------------------------------------
  {
    const int target = blockIdx.x;
    int carry = 0;
    bool found = false;
    for (int base = 0; base < p.B; base += kPWarps * 32) {
      const int i = base + tid;
      int t = 0;
      if (i < p.B) t = (p.qo_indptr[i + 1] - p.qo_indptr[i] + kQTile - 1) / kQTile;
      int blk;
      const int incl = block_inclusive_scan<kPWarps * 32>(t, s_scan, &blk) + carry;
      const int excl = incl - t;
      if (t > 0 && excl <= target && target < incl) {
        s_loc[0] = i;
        s_loc[1] = target - excl;
      }
      carry += blk;
      if (carry > target) {
        found = true;
        break;
      }
    }
    __syncthreads();
    if (!found) return;
  }
  const int s = s_loc[0];
  const int q0 = p.qo_indptr[s];
  const int Lq = p.qo_indptr[s + 1] - q0;
  const int j0 = s_loc[1] * kQTile;  // first query row of this tile inside the sequence
  const int Lkv = seq_kv_len(p.kv_indptr, p.last_page_len, s, p.log2_ps);
  const int page0 = p.kv_indptr[s];
  const int ps_mask = (1 << p.log2_ps) - 1;
  const int rows_here = min(kQTile, Lq - j0);

  const uint32_t q_smem = smem_u32(smem);
  const uint32_t kv_smem = q_smem + kQBytes;

  // kv positions needed by this q tile
  int kv_hi = Lkv;
  if (p.causal) kv_hi = min(Lkv, Lkv - Lq + j0 + rows_here);
  if (kv_hi < 0) kv_hi = 0;
  const int n_tiles = (kv_hi + kKVTile - 1) / kKVTile;

  // ---- async loads ------------------------------------------------------------------------------
  const int ld_chunk = tid & 15, ld_row0 = tid >> 4;  // 8 rows per pass, 8 passes per 64-row tile
  {
#pragma unroll
    for (int i = 0; i < kQTile / 8; ++i) {
      const int r = ld_row0 + 8 * i;
      const bool ok = r < rows_here;
      const __nv_bfloat16* src = p.q + (int64_t)(q0 + j0 + (ok ? r : 0)) * p.q_row_stride + (int64_t)hq * kPD + ld_chunk * 8;
      cp_async_16(q_smem + r * kPRowBytes + ((ld_chunk ^ (r & 7)) << 4), src, ok ? 16u : 0u);
    }
  }
  auto issue_kv = [&](int tile) {
    const uint32_t st = kv_smem + (tile % kPStages) * 2 * kKVBytes;
#pragma unroll
    for (int i = 0; i < kKVTile / 8; ++i) {
      const int r = ld_row0 + 8 * i;
      const int tok = tile * kKVTile + r;
      const bool ok = tok < kv_hi;
      int64_t off = 0;
      if (ok) {
        const int page = __ldg(p.kv_indices + page0 + (tok >> p.log2_ps));
        off = ((int64_t)(page << p.log2_ps) + (tok & ps_mask)) * p.kv_slot_stride;
      }
      off += (int64_t)kv_head * kPD + ld_chunk * 8;
      const uint32_t dst = st + r * kPRowBytes + ((ld_chunk ^ (r & 7)) << 4);
      cp_async_16(dst, p.k_cache + off, ok ? 16u : 0u);
      cp_async_16(dst + kKVBytes, p.v_cache + off, ok ? 16u : 0u);
    }
  };
  if (n_tiles > 0) issue_kv(0);
  cp_async_commit();  // group 0: Q + KV tile 0

  const int lm = lane >> 3, lr = lane & 7;
  const int a_row = warp * 16 + (lm & 1) * 8 + lr, a_csel = lm >> 1;       // Q (A operand)
  const int k_tok = (lm >> 1) * 8 + lr, k_csel = lm & 1;                  // K (B operand)
  const int v_tok = (lm & 1) * 8 + lr, v_csel = lm >> 1;                  // V (B operand, trans)
  const int quad = lane & 3;
  const int r_lo = warp * 16 + (lane >> 2), r_hi = r_lo + 8;              // rows inside the q tile
  const int lim_lo = p.causal ? (Lkv - Lq + j0 + r_lo) : (Lkv - 1);       // last visible kv position
  const int lim_hi = p.causal ? (Lkv - Lq + j0 + r_hi) : (Lkv - 1);

  uint32_t qa[kPD / 16][4];
  float o[kPD / 8][4];
#pragma unroll
  for (int j = 0; j < kPD / 8; ++j) o[j][0] = o[j][1] = o[j][2] = o[j][3] = 0.f;
  float m_lo = -INFINITY, m_hi = -INFINITY, l_lo = 0.f, l_hi = 0.f;

  for (int tile = 0; tile < n_tiles; ++tile) {
    if (tile + 1 < n_tiles) issue_kv(tile + 1);
    cp_async_commit();
    cp_async_wait<1>();
    __syncthreads();
    if (tile == 0) {
#pragma unroll
      for (int ks = 0; ks < kPD / 16; ++ks)
        ldmatrix_x4(qa[ks], q_smem + a_row * kPRowBytes + (((2 * ks + a_csel) ^ (a_row & 7)) << 4));
    }
    const uint32_t kst = kv_smem + (tile % kPStages) * 2 * kKVBytes;
    const uint32_t vst = kst + kKVBytes;

    float sc[kKVTile / 8][4];
#pragma unroll
    for (int n = 0; n < kKVTile / 8; ++n) sc[n][0] = sc[n][1] = sc[n][2] = sc[n][3] = 0.f;
#pragma unroll
    for (int ks = 0; ks < kPD / 16; ++ks) {
#pragma unroll
      for (int nb = 0; nb < kKVTile / 16; ++nb) {
        uint32_t b[4];
        const int tok = nb * 16 + k_tok;
        ldmatrix_x4(b, kst + tok * kPRowBytes + (((2 * ks + k_csel) ^ (tok & 7)) << 4));
        mma_bf16_16816(sc[2 * nb], qa[ks], b[0], b[1]);
        mma_bf16_16816(sc[2 * nb + 1], qa[ks], b[2], b[3]);
      }
    }
    // scale, mask, row max
    const int kv0 = tile * kKVTile;
    const bool need_mask = (kv0 + kKVTile - 1 > lim_lo) || (kv0 + kKVTile > kv_hi);  // lim_lo <= lim_hi
    float mx_lo = -INFINITY, mx_hi = -INFINITY;
#pragma unroll
    for (int n = 0; n < kKVTile / 8; ++n) {
#pragma unroll
      for (int e = 0; e < 4; ++e) {
        float x = sc[n][e] * p.scale_log2;
        if (need_mask) {
          const int kvp = kv0 + n * 8 + quad * 2 + (e & 1);
          const int lim = (e < 2) ? lim_lo : lim_hi;
          if (kvp > lim || kvp >= kv_hi) x = -INFINITY;
        }
        sc[n][e] = x;
        if (e < 2) mx_lo = fmaxf(mx_lo, x); else mx_hi = fmaxf(mx_hi, x);
      }
    }
    mx_lo = fmaxf(mx_lo, __shfl_xor_sync(0xffffffffu, mx_lo, 1));
    mx_lo = fmaxf(mx_lo, __shfl_xor_sync(0xffffffffu, mx_lo, 2));
    mx_hi = fmaxf(mx_hi, __shfl_xor_sync(0xffffffffu, mx_hi, 1));
    mx_hi = fmaxf(mx_hi, __shfl_xor_sync(0xffffffffu, mx_hi, 2));
    const float mn_lo = fmaxf(m_lo, mx_lo), mn_hi = fmaxf(m_hi, mx_hi);
    // a row with nothing visible so far keeps m = -inf; use 0 as the subtrahend to avoid inf-inf
    const float sub_lo = mn_lo == -INFINITY ? 0.f : mn_lo, sub_hi = mn_hi == -INFINITY ? 0.f : mn_hi;
    const float al_lo = fast_exp2(m_lo - sub_lo), al_hi = fast_exp2(m_hi - sub_hi);
    float sum_lo = 0.f, sum_hi = 0.f;
    uint32_t pa[kKVTile / 16][4];
#pragma unroll
    for (int n = 0; n < kKVTile / 8; ++n) {
      const float p0 = fast_exp2(sc[n][0] - sub_lo), p1 = fast_exp2(sc[n][1] - sub_lo);
      const float p2 = fast_exp2(sc[n][2] - sub_hi), p3 = fast_exp2(sc[n][3] - sub_hi);
      sum_lo += p0 + p1;
      sum_hi += p2 + p3;
      // A fragment of P for k-block n/2: a0,a1 = k 0..7 (n even), a2,a3 = k 8..15 (n odd)
      pa[n >> 1][(n & 1) * 2 + 0] = pack_bf16x2(p0, p1);
      pa[n >> 1][(n & 1) * 2 + 1] = pack_bf16x2(p2, p3);
    }
    l_lo = l_lo * al_lo + sum_lo;
    l_hi = l_hi * al_hi + sum_hi;
    m_lo = mn_lo;
    m_hi = mn_hi;
#pragma unroll
    for (int j = 0; j < kPD / 8; ++j) {
      o[j][0] *= al_lo; o[j][1] *= al_lo; o[j][2] *= al_hi; o[j][3] *= al_hi;
    }
#pragma unroll
    for (int kb = 0; kb < kKVTile / 16; ++kb) {
#pragma unroll
      for (int j = 0; j < kPD / 16; ++j) {
        uint32_t b[4];
        const int tok = kb * 16 + v_tok;
        ldmatrix_x4_trans(b, vst + tok * kPRowBytes + (((2 * j + v_csel) ^ (tok & 7)) << 4));
        mma_bf16_16816(o[2 * j], pa[kb], b[0], b[1]);
        mma_bf16_16816(o[2 * j + 1], pa[kb], b[2], b[3]);
      }
    }
    __syncthreads();  // everyone done with this stage before it is refilled
  }
  cp_async_wait<0>();

  l_lo += __shfl_xor_sync(0xffffffffu, l_lo, 1);
  l_lo += __shfl_xor_sync(0xffffffffu, l_lo, 2);
  l_hi += __shfl_xor_sync(0xffffffffu, l_hi, 1);
  l_hi += __shfl_xor_sync(0xffffffffu, l_hi, 2);
  const float inv_lo = l_lo > 0.f ? 1.f / l_lo : 0.f, inv_hi = l_hi > 0.f ? 1.f / l_hi : 0.f;
  if (r_lo < rows_here) {
    __nv_bfloat16* dst = p.out + (int64_t)(q0 + j0 + r_lo) * p.out_row_stride + (int64_t)hq * kPD + quad * 2;
#pragma unroll
    for (int j = 0; j < kPD / 8; ++j)
      *reinterpret_cast<uint32_t*>(dst + j * 8) = pack_bf16x2(o[j][0] * inv_lo, o[j][1] * inv_lo);
  }
  if (r_hi < rows_here) {
    __nv_bfloat16* dst = p.out + (int64_t)(q0 + j0 + r_hi) * p.out_row_stride + (int64_t)hq * kPD + quad * 2;
#pragma unroll
    for (int j = 0; j < kPD / 8; ++j)
      *reinterpret_cast<uint32_t*>(dst + j * 8) = pack_bf16x2(o[j][2] * inv_hi, o[j][3] * inv_hi);
  }
}

static int ilog2_exact_p(int x) {
  if (x <= 0 || (x & (x - 1))) return -1;
  int l = 0;
  while ((1 << l) < x) ++l;
  return l;
}

int launch_prefill_mma(const PrefillParams& p, int total_q, cudaStream_t st) {
  static std::atomic<bool> attr_done[64];
  if (cudaError_t e = ensure_dyn_smem(prefill_mma_kernel, kPrefillSmem, attr_done); e != cudaSuccess)
    return fail("b200_prefill: cudaFuncSetAttribute: %s", cudaGetErrorString(e));
  const int max_q_tiles = total_q / kQTile + p.B;  // >= sum_i ceil(Lq_i / 64)
  prefill_mma_kernel<<<dim3(max_q_tiles, p.Hq), kPWarps * 32, kPrefillSmem, st>>>(p);
  return check_launch("b200_prefill(mma)");
}

}  // namespace b200

extern "C" size_t b200_prefill_workspace_bytes(int32_t B, int32_t total_q, int32_t Hq, int32_t Hkv, int32_t D) {
  (void)D;
  if (B <= 0 || total_q <= 0 || Hq <= 0 || Hkv <= 0 || Hq % Hkv) return 256;
  // work list of the persistent tcgen05 kernel (the mma.sync fallback needs none); never zero so that
  // callers can always pass a pointer
  const size_t need = b200::prefill_sm100_supports(Hq / Hkv) ? b200::prefill_sm100_workspace_bytes(B, total_q, Hq / Hkv) : 0;
  return (need + 255) & ~(size_t)255 ? ((need + 255) & ~(size_t)255) : 256;
}

extern "C" int b200_prefill(const void* q, int64_t q_row_stride, const void* k_cache, const void* v_cache,
                            int64_t kv_slot_stride, const int32_t* qo_indptr, const int32_t* kv_indptr,
                            const int32_t* kv_indices, const int32_t* last_page_len, void* out,
                            int64_t out_row_stride, void* workspace, size_t workspace_bytes, int32_t B,
                            int32_t total_q, int32_t Hq, int32_t Hkv, int32_t D, int32_t page_size, float sm_scale,
                            int32_t causal, int32_t dtype, b200_stream_t stream) {
  using namespace b200;
  if (B < 0 || total_q < 0) return fail("b200_prefill: negative size");
  if (B == 0 || total_q == 0) return 0;
  if (D != kPD) return fail("b200_prefill: head_dim %d not built (round 1 builds D=128)", D);
  if (dtype != B200_DTYPE_BF16) return fail("b200_prefill: dtype %d not built (bf16 only)", dtype);
  if (Hkv <= 0 || Hq % Hkv) return fail("b200_prefill: Hq %d not a multiple of Hkv %d", Hq, Hkv);
  const int log2_ps = ilog2_exact_p(page_size);
  if (log2_ps < 0) return fail("b200_prefill: page_size %d is not a power of two", page_size);
  if (!q || !k_cache || !v_cache || !qo_indptr || !kv_indptr || !kv_indices || !last_page_len || !out)
    return fail("b200_prefill: null pointer");
  if (((uintptr_t)q | (uintptr_t)k_cache | (uintptr_t)v_cache) & 15 || (kv_slot_stride & 7) || (q_row_stride & 7))
    return fail("b200_prefill: q / k_cache / v_cache must be 16-byte aligned");
  if (((uintptr_t)out & 3) || (out_row_stride & 1)) return fail("b200_prefill: out must be 4-byte aligned");
  PrefillParams p;
  p.q = (const __nv_bfloat16*)q;
  p.k_cache = (const __nv_bfloat16*)k_cache;
  p.v_cache = (const __nv_bfloat16*)v_cache;
  p.qo_indptr = qo_indptr;
  p.kv_indptr = kv_indptr;
  p.kv_indices = kv_indices;
  p.last_page_len = last_page_len;
  p.out = (__nv_bfloat16*)out;
  p.q_row_stride = q_row_stride;
  p.out_row_stride = out_row_stride;
  p.kv_slot_stride = kv_slot_stride;
  p.B = B; p.Hq = Hq; p.Hkv = Hkv; p.group = Hq / Hkv; p.log2_ps = log2_ps; p.causal = causal ? 1 : 0;
  p.scale_log2 = sm_scale * kLog2e;
  // One kernel per shape class, decided by the shape alone (no environment switches): the tcgen05/TMEM kernel for
  // every GQA group that packs into a 128-row tile (1, 2, 4, 8, ...: all Qwen3 models but the 14B), the mma.sync
  // kernel for the others (group 3, 5 (Qwen3-14B: 40/8), 6, 7).
  if (prefill_sm100_supports(p.group)) return launch_prefill_sm100(p, total_q, workspace, workspace_bytes, (cudaStream_t)stream);
  return launch_prefill_mma(p, total_q, (cudaStream_t)stream);
}
