// (4) Ragged paged prefill attention on the 5th-gen tensor cores: tcgen05.mma with TMEM
// accumulators.  Replaces layers/attention.py:360-364 (FlashInfer prefill, which resolves to the
// sm_80 mma.sync "fa2" template on sm_100: flashinfer/utils.py:509-518).
//
// Tile: one CTA = 128 packed query rows of one KV head.  Row r is (token j0 + r / G, query head
// kvh*G + r % G) with G = Hq/Hkv, so all G query heads that share a KV head sit in the M dimension
// of ONE UMMA tile and K/V are gathered once per KV head, not once per query head.
//   S[128 x 128] = Q[128 x 128d] . K_tile^T      UMMA 128x128x16, A/B K-major, SWIZZLE_128B
//   softmax rows in registers (thread == row == TMEM lane: no shuffles), exp2 domain, fp32
//   O[128 x 128d] += P[128 x 128tok] . V_tile      UMMA 128x128x16, A K-major (P, bf16 in smem),
//                                                  B MN-major (V rows are [token][d])
// K/V tiles (128 tokens) are gathered through the page table with 16-byte cp.async into the
// canonical SWIZZLE_128B layout (two [128 x 64] blocks of 128-byte rows, chunk ^= row & 7) — the
// same bytes TMA would produce, but page_size 1 needs a per-row gather TMA tiles cannot express.
// Double-buffered: tile t+1 streams in while tile t is on the tensor cores.
// O lives in TMEM across the KV loop; a row is rescaled in place (tcgen05.ld/st) only when its
// running max moves.
#include "common.cuh"
#include "prefill_params.cuh"
#include "sm100.cuh"

namespace b200 {

using namespace sm100;

constexpr int kTD = 128;                       // head dim
constexpr int kTM = 128;                       // packed query rows per CTA
constexpr int kTN = 128;                       // KV tokens per tile
constexpr int kBlkBytes = 128 * 128;           // one [128 rows x 64 elem] swizzled block = 16 KiB
constexpr int kTileBytes2 = 2 * kBlkBytes;     // a full [128 x 128] bf16 operand tile = 32 KiB
constexpr int kOffQ = 0;
constexpr int kOffP = kOffQ + kTileBytes2;
constexpr int kOffK = kOffP + kTileBytes2;     // 2 stages
constexpr int kOffV = kOffK + 2 * kTileBytes2; // 2 stages
constexpr int kSm100Smem = kOffV + 2 * kTileBytes2 + 1024;  // + alignment slack = 193 KiB
constexpr int kTmemCols = 256;                 // S: [0,128)  O: [128,256)


__global__ void __launch_bounds__(128, 1) prefill_sm100_kernel(const PrefillParams p, const int log2_g) {
  extern __shared__ uint8_t smem_raw[];
  __shared__ int s_scan[5];
  __shared__ int s_loc[2];
  __shared__ uint32_t s_tmem;
  __shared__ __align__(8) uint64_t s_bar[2];

  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const int kvh = blockIdx.y;
  const int G = 1 << log2_g;
  const int tok_per_tile = kTM >> log2_g;

  // ---- locate (sequence, q tile) -------------------------------------------------------------------
  {
    const int target = blockIdx.x;
    int carry = 0;
    bool found = false;
    for (int base = 0; base < p.B; base += 128) {
      const int i = base + tid;
      int t = 0;
      if (i < p.B) t = (p.qo_indptr[i + 1] - p.qo_indptr[i] + tok_per_tile - 1) / tok_per_tile;
      int blk;
      const int incl = block_inclusive_scan<128>(t, s_scan, &blk) + carry;
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
  const int j0 = s_loc[1] * tok_per_tile;
  const int toks_here = min(tok_per_tile, Lq - j0);
  const int rows_here = toks_here << log2_g;
  const int Lkv = seq_kv_len(p.kv_indptr, p.last_page_len, s, p.log2_ps);
  const int page0 = p.kv_indptr[s];
  const int ps_mask = (1 << p.log2_ps) - 1;
  int kv_hi = Lkv;
  if (p.causal) kv_hi = min(Lkv, Lkv - Lq + j0 + toks_here);
  if (kv_hi < 0) kv_hi = 0;
  const int n_tiles = (kv_hi + kTN - 1) / kTN;

  const uint32_t smem = (smem_u32(smem_raw) + 1023u) & ~1023u;   // SWIZZLE_128B atoms need 1024-byte alignment
  const uint32_t bar_s = smem_u32(&s_bar[0]), bar_o = smem_u32(&s_bar[1]);

  // ---- one-time setup: barriers + TMEM ---------------------------------------------------------------
  if (tid == 0) {
    mbar_init(bar_s, 1);
    mbar_init(bar_o, 1);
    fence_mbar_init();
  }
  if (warp == 0) {
    tmem_alloc(smem_u32(&s_tmem), kTmemCols);
    tmem_relinquish();
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem = s_tmem;
  const uint32_t tmem_lane = tmem + ((uint32_t)(warp * 32) << 16);   // this warp's 32-lane quadrant
  const uint32_t tS = tmem_lane, tO = tmem_lane + kTN;

  // ---- loaders (all 128 threads; half-warp = one 256-byte row) --------------------------------------
  const int gc = tid & 15, lrow0 = tid >> 4;           // global 16-byte chunk 0..15; rows lrow0 + 8*i
  const uint32_t chunk_off = (uint32_t)(gc >> 3) * kBlkBytes;
  const int c8 = gc & 7;
  {
#pragma unroll
    for (int i = 0; i < kTM / 8; ++i) {
      const int r = lrow0 + 8 * i;
      const bool ok = r < rows_here;
      const int tok = j0 + (r >> log2_g), head = kvh * G + (r & (G - 1));
      const __nv_bfloat16* src = p.q + (int64_t)(q0 + (ok ? tok : j0)) * p.q_row_stride + (int64_t)(ok ? head : kvh * G) * kTD + gc * 8;
      cp_async_16(smem + kOffQ + chunk_off + r * 128 + ((c8 ^ (r & 7)) << 4), src, ok ? 16u : 0u);
    }
  }
  auto issue_kv = [&](int tile) {
    const uint32_t kst = smem + kOffK + (tile & 1) * kTileBytes2 + chunk_off;
    const uint32_t vst = smem + kOffV + (tile & 1) * kTileBytes2 + chunk_off;
#pragma unroll
    for (int i = 0; i < kTN / 8; ++i) {
      const int r = lrow0 + 8 * i;
      const int tok = tile * kTN + r;
      const bool ok = tok < kv_hi;
      int64_t off = (int64_t)kvh * kTD + gc * 8;
      if (ok) {
        const int page = __ldg(p.kv_indices + page0 + (tok >> p.log2_ps));
        off += ((int64_t)(page << p.log2_ps) + (tok & ps_mask)) * p.kv_slot_stride;
      }
      const uint32_t o = r * 128 + ((c8 ^ (r & 7)) << 4);
      cp_async_16(kst + o, p.k_cache + off, ok ? 16u : 0u);
      cp_async_16(vst + o, p.v_cache + off, ok ? 16u : 0u);
    }
  };
  if (n_tiles > 0) issue_kv(0);
  cp_async_commit();

  // ---- per-row state ------------------------------------------------------------------------------------
  const int r = tid;                                              // my packed row == my TMEM lane
  const int my_tok = j0 + (r >> log2_g);
  const int lim = p.causal ? (Lkv - Lq + my_tok) : (Lkv - 1);     // last visible kv position
  const int lim0 = p.causal ? (Lkv - Lq + j0) : (Lkv - 1);        // smallest limit in the tile
  float m_run = -INFINITY, l_run = 0.f;

  constexpr uint32_t idesc_qk = make_idesc_bf16_f32(kTM, kTN, 0, 0);   // A K-major, B K-major
  constexpr uint32_t idesc_pv = make_idesc_bf16_f32(kTM, kTD, 0, 1);   // A K-major, B MN-major (V)
  uint32_t phase_s = 0, phase_o = 0;

  for (int tile = 0; tile < n_tiles; ++tile) {
    if (tile + 1 < n_tiles) issue_kv(tile + 1);
    cp_async_commit();
    cp_async_wait<1>();            // my chunks of tile `tile` (and Q) have landed
    fence_proxy_async_smem();      // ... and are visible to the tensor-core (async) proxy
    __syncthreads();

    const uint32_t kst = smem + kOffK + (tile & 1) * kTileBytes2;
    const uint32_t vst = smem + kOffV + (tile & 1) * kTileBytes2;

    // ---- S = Q K^T ---------------------------------------------------------------------------------------
    if (tid == 0) {
      tc_fence_after();
#pragma unroll
      for (int ks = 0; ks < kTD / 16; ++ks) {
        const uint32_t koff = (uint32_t)(ks >> 2) * kBlkBytes + (uint32_t)(ks & 3) * 32;
        umma_ss(tmem, make_smem_desc(smem + kOffQ + koff, 16, 1024), make_smem_desc(kst + koff, 16, 1024), idesc_qk,
                ks > 0);
      }
      umma_commit(bar_s);
    }
    __syncwarp();
    mbar_wait(bar_s, phase_s);
    phase_s ^= 1;
    tc_fence_after();

    // ---- softmax on my row --------------------------------------------------------------------------------
    const int kv0 = tile * kTN;
    const bool need_mask = (kv0 + kTN - 1 > lim0) || (kv0 + kTN > kv_hi);
    uint32_t sv[4][32];
#pragma unroll
    for (int c = 0; c < 4; ++c) tmem_ld_32x32b_x32(tS + c * 32, sv[c]);
    tmem_wait_ld();
    float mx = -INFINITY;
#pragma unroll
    for (int c = 0; c < 4; ++c) {
#pragma unroll
      for (int j = 0; j < 32; ++j) {
        float x = __uint_as_float(sv[c][j]) * p.scale_log2;
        if (need_mask) {
          const int kvp = kv0 + c * 32 + j;
          if (kvp > lim || kvp >= kv_hi) x = -INFINITY;
        }
        sv[c][j] = __float_as_uint(x);
        mx = fmaxf(mx, x);
      }
    }
    const float m_new = fmaxf(m_run, mx);
    const float sub = (m_new == -INFINITY) ? 0.f : m_new;
    const float alpha = fast_exp2(m_run - sub);
    float sum = 0.f;
    // P -> bf16 -> swizzled K-major smem tile (A operand of the PV MMA)
#pragma unroll
    for (int c = 0; c < 4; ++c) {
#pragma unroll
      for (int j8 = 0; j8 < 4; ++j8) {
        uint32_t w[4];
#pragma unroll
        for (int e = 0; e < 4; ++e) {
          const float a = fast_exp2(__uint_as_float(sv[c][j8 * 8 + 2 * e]) - sub);
          const float b = fast_exp2(__uint_as_float(sv[c][j8 * 8 + 2 * e + 1]) - sub);
          sum += a + b;
          w[e] = pack_bf16x2(a, b);
        }
        const int chunk = c * 4 + j8;                   // 16-byte chunk 0..15 along the token (K) axis
        const uint32_t dst = smem + kOffP + (uint32_t)(chunk >> 3) * kBlkBytes + r * 128 + (((chunk & 7) ^ (r & 7)) << 4);
        asm volatile("st.shared.v4.b32 [%0], {%1,%2,%3,%4};\n" ::"r"(dst), "r"(w[0]), "r"(w[1]), "r"(w[2]), "r"(w[3]) : "memory");
      }
    }
    l_run = l_run * alpha + sum;
    // ---- rescale my O row in TMEM if the running max moved (PV of the previous tile has completed) ----------
    if (tile > 0 && __any_sync(0xffffffffu, m_new != m_run)) {
#pragma unroll
      for (int c = 0; c < 4; ++c) {
        uint32_t ov[32];
        tmem_ld_32x32b_x32(tO + c * 32, ov);
        tmem_wait_ld();
#pragma unroll
        for (int j = 0; j < 32; ++j) ov[j] = __float_as_uint(__uint_as_float(ov[j]) * alpha);
        tmem_st_32x32b_x32(tO + c * 32, ov);
      }
      tmem_wait_st();
    }
    m_run = m_new;
    fence_proxy_async_smem();      // P (generic-proxy stores) -> async proxy
    tc_fence_before();
    __syncthreads();

    // ---- O += P V --------------------------------------------------------------------------------------------
    if (tid == 0) {
      tc_fence_after();
#pragma unroll
      for (int ks = 0; ks < kTN / 16; ++ks) {
        const uint32_t poff = (uint32_t)(ks >> 2) * kBlkBytes + (uint32_t)(ks & 3) * 32;   // 16 tokens along K of P
        const uint32_t voff = (uint32_t)ks * 16 * 128;                                      // 16 token rows of V
        umma_ss(tmem + kTN, make_smem_desc(smem + kOffP + poff, 16, 1024), make_smem_desc(vst + voff, kBlkBytes, 1024),
                idesc_pv, (tile > 0 || ks > 0));
      }
      umma_commit(bar_o);
    }
    __syncwarp();
    mbar_wait(bar_o, phase_o);     // K/V stage, P tile and O are free/complete again
    phase_o ^= 1;
    tc_fence_after();
  }
  cp_async_wait<0>();

  // ---- epilogue: O / l -> bf16 -> global ------------------------------------------------------------------------
  const float inv_l = l_run > 0.f ? 1.f / l_run : 0.f;
  const bool row_ok = r < rows_here;
  __nv_bfloat16* dst = p.out + (int64_t)(q0 + (row_ok ? my_tok : j0)) * p.out_row_stride + (int64_t)(kvh * G + (r & (G - 1))) * kTD;
#pragma unroll
  for (int c = 0; c < 4; ++c) {
    uint32_t ov[32];
    if (n_tiles > 0) {
      tmem_ld_32x32b_x32(tO + c * 32, ov);
      tmem_wait_ld();
    } else {
#pragma unroll
      for (int j = 0; j < 32; ++j) ov[j] = 0u;
    }
    if (row_ok) {
#pragma unroll
      for (int j8 = 0; j8 < 4; ++j8) {
        uint4 w;
        w.x = pack_bf16x2(__uint_as_float(ov[j8 * 8 + 0]) * inv_l, __uint_as_float(ov[j8 * 8 + 1]) * inv_l);
        w.y = pack_bf16x2(__uint_as_float(ov[j8 * 8 + 2]) * inv_l, __uint_as_float(ov[j8 * 8 + 3]) * inv_l);
        w.z = pack_bf16x2(__uint_as_float(ov[j8 * 8 + 4]) * inv_l, __uint_as_float(ov[j8 * 8 + 5]) * inv_l);
        w.w = pack_bf16x2(__uint_as_float(ov[j8 * 8 + 6]) * inv_l, __uint_as_float(ov[j8 * 8 + 7]) * inv_l);
        *reinterpret_cast<uint4*>(dst + c * 32 + j8 * 8) = w;
      }
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 0) tmem_dealloc(tmem, kTmemCols);
}

int launch_prefill_sm100(const PrefillParams& p, int total_q, cudaStream_t st) {
  int log2_g = 0;
  while ((1 << log2_g) < p.group) ++log2_g;
  static bool attr_set = false;
  if (!attr_set) {
    cudaError_t e = cudaFuncSetAttribute(prefill_sm100_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, kSm100Smem);
    if (e != cudaSuccess) return fail("b200_prefill(sm100): cudaFuncSetAttribute: %s", cudaGetErrorString(e));
    attr_set = true;
  }
  const int tok_per_tile = kTM >> log2_g;
  const int max_q_tiles = total_q / tok_per_tile + p.B;   // >= sum_i ceil(Lq_i / tok_per_tile)
  prefill_sm100_kernel<<<dim3(max_q_tiles, p.Hkv), 128, kSm100Smem, st>>>(p, log2_g);
  return check_launch("b200_prefill(sm100)");
}

bool prefill_sm100_supports(int group) { return group >= 1 && group <= 64 && (group & (group - 1)) == 0; }

}  // namespace b200
