// (3) Paged GQA decode attention for sm_100a — split-KV streaming kernel + merge pass.
// Replaces layers/attention.py:365-369 (FlashInfer BatchDecodeWithPagedKVCacheWrapper, CUDA-core
// cp.async kernel decode.cuh:613 + host plan()).  Own design, not a port:
//
//  * Schedule on the device, balanced by construction.  All sequences' KV is viewed as one
//    flat list of 16-token tiles; the grid is ONE CTA PER SM.  The list is cut into G*subs
//    contiguous "streams" (subs = warps per KV head inside a CTA: 1 at Hkv >= 8 ... 8 at Hkv = 1);
//    stream k = sub*G + cta owns tiles [k*per, (k+1)*per).  A stream may cover the tail of one
//    sequence and the head of the next; every (stream, sequence) intersection is a "segment"
//    that yields one partial (o, lse) in slot k + s (unique: streams and sequences are both
//    ordered).  No host plan, no sync, grid independent of the batch => graph-safe.
//  * One warp = one KV head of one stream (all `group` query heads that share the KV head form
//    the M rows of the mma tile), so K/V are read from HBM exactly once per KV head, and the
//    warps of a stream walk the SAME tokens of different heads => whole contiguous token rows
//    (page_size 1: 2 x Hkv*256 B per token; page_size 16: whole pages).
//  * Each warp runs a private 3-stage cp.async (LDGSTS, 16 B, L1-bypass) pipeline into
//    XOR-swizzled shared memory that NEVER drains inside a stream: the producer cursor (slot
//    ids one tile ahead of the copies, copies two tiles ahead of the math, the next sequence's
//    q rows fetched when the producer crosses into it) runs across segment boundaries, so short
//    sequences (the engine's decode batches: B 256, L <= 2048) do not pay a pipeline refill per
//    sequence.  No __syncthreads after the prologue, no combine inside the CTA.
//  * Prologue: ONE block scan of the batch's tile counts (inclusive prefix kept in shared
//    memory), then every warp binary-searches the sequence its stream starts in.
//  * q.K^T and p.V use mma.sync m16n8k16 (bf16 in, fp32 accumulate) fed by ldmatrix; decode is
//    HBM-bound, tensor throughput is irrelevant here, the point is ~10x fewer issue slots than
//    an FMA formulation.  Online softmax in fp32 with quad shuffles (exp2 domain).
//  * Segments that cover a whole sequence write bf16 output directly; the rest write fp32
//    partials + log-sum-exp to the workspace and a small merge kernel (PDL) combines them.
//
// Algorithmic bytes per call: sum_i 2*Lkv_i*Hkv*D*2 + 2*B*Hq*D*2 + index bytes.
#include "common.cuh"

namespace b200 {

constexpr int kD = 128;             // head dim built in round 1
constexpr int kTile = 16;           // tokens per pipeline tile
constexpr int kWarps = 8;           // warps per CTA
constexpr int kStages = 3;          // cp.async stages per warp
constexpr int kRowBytes = kD * 2;   // 256 B per (token, head) row
constexpr int kTileBytes = kTile * kRowBytes;          // 4 KiB (K or V)
constexpr int kStageBytes = 2 * kTileBytes;            // K + V
constexpr int kWarpSmem = kStages * kStageBytes;       // 24 KiB
constexpr int kDecodeSmem = kWarps * kWarpSmem;        // 192 KiB
constexpr int kMinTilesPerStream = 2;  // do not split finer than 32 tokens per stream
constexpr int kMaxSmemSeqs = 4096;     // batches up to this size keep their tile prefix in shared memory

struct DecodeParams {
  const __nv_bfloat16* q;
  const __nv_bfloat16* k_cache;
  const __nv_bfloat16* v_cache;
  const int32_t* kv_indptr;
  const int32_t* kv_indices;
  const int32_t* last_page_len;
  __nv_bfloat16* out;
  float* part_o;      // [G*subs+B][Hq][D]
  float* part_lse;    // [G*subs+B][Hq]
  int32_t* seg_first; // [B]
  int32_t* seg_last;  // [B]
  int64_t q_row_stride, out_row_stride, kv_slot_stride;
  int B, Hq, Hkv, group, log2_ps;
  int heads_per_cta;  // KV heads handled by one CTA (<= 8)
  int subs;           // streams per CTA = warps per KV head
  float scale_log2;   // sm_scale * log2(e)
};

__device__ __forceinline__ int tiles_of(const DecodeParams& p, int i) {
  return (seq_kv_len(p.kv_indptr, p.last_page_len, i, p.log2_ps) + kTile - 1) / kTile;
}

__global__ void __launch_bounds__(kWarps * 32, 1) decode_kernel(const DecodeParams p) {
  extern __shared__ __align__(1024) uint8_t smem[];
  __shared__ int s_scan[kWarps + 1];
  __shared__ int s_incl[kMaxSmemSeqs];

  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const int G = gridDim.x;
  const int head_group = blockIdx.y;  // which block of heads_per_cta KV heads
  griddep_launch_dependents();        // the merge grid may become resident now; it waits for this grid to finish

  // ---- device-side schedule: one scan of the batch's tile counts ------------------------------
  const bool in_smem = p.B <= kMaxSmemSeqs;
  int total = 0;
  for (int base = 0; base < p.B; base += kWarps * 32) {
    const int i = base + tid;
    const int t = i < p.B ? tiles_of(p, i) : 0;
    int blk;
    const int incl = block_inclusive_scan<kWarps * 32>(t, s_scan, &blk) + total;
    if (in_smem && i < p.B) s_incl[i] = incl;
    total += blk;
  }
  __syncthreads();   // the last block-wide barrier of this kernel

  const int hh = warp % p.heads_per_cta;           // KV head inside the CTA's head block
  const int sub = warp / p.heads_per_cta;          // stream of this CTA
  if (sub >= p.subs) return;
  const int n_streams = G * p.subs;
  const int k_stream = sub * G + blockIdx.x;       // consecutive streams sit on different SMs
  int per = (total + n_streams - 1) / n_streams;
  if (per < kMinTilesPerStream) per = kMinTilesPerStream;
  const int t_begin = k_stream * per;
  if (t_begin >= total) return;
  const int t_end = min(t_begin + per, total);

  // ---- the sequence this stream starts in: smallest s with incl(s) > t_begin ----------------------
  int s0, pre0;
  if (in_smem) {
    int lo = 0, hi = p.B - 1;
    while (lo < hi) {
      const int mid = (lo + hi) >> 1;
      if (s_incl[mid] > t_begin) hi = mid; else lo = mid + 1;
    }
    s0 = lo;
    pre0 = lo > 0 ? s_incl[lo - 1] : 0;
  } else {
    int carry = 0;
    s0 = pre0 = 0;
    for (int base = 0; base < p.B; base += 32) {
      const int i = base + lane;
      const int t = i < p.B ? tiles_of(p, i) : 0;
      int x = t;
#pragma unroll
      for (int o = 1; o < 32; o <<= 1) {
        const int y = __shfl_up_sync(0xffffffffu, x, o);
        if (lane >= o) x += y;
      }
      x += carry;
      const unsigned m = __ballot_sync(0xffffffffu, x > t_begin);
      if (m) {
        const int l = __ffs(m) - 1;
        s0 = base + l;
        pre0 = __shfl_sync(0xffffffffu, x, l) - __shfl_sync(0xffffffffu, t, l);
        break;
      }
      carry = __shfl_sync(0xffffffffu, x, 31);
    }
  }

  const int kv_head = head_group * p.heads_per_cta + hh;
  const int row = lane >> 2;                        // mma row = query head inside the group
  const int quad = lane & 3;
  const bool row_real = row < p.group;
  const uint32_t warp_smem = smem_u32(smem) + warp * kWarpSmem;

  // per-lane constants of the loaders
  const int ld_chunk = lane & 15;                   // 16-byte chunk inside the 256-byte row
  const int ld_rsel = lane >> 4;                    // rows 2i + ld_rsel
  const __nv_bfloat16* kbase = p.k_cache + (int64_t)kv_head * kD + ld_chunk * 8;
  const __nv_bfloat16* vbase = p.v_cache + (int64_t)kv_head * kD + ld_chunk * 8;
  const int ps_mask = (1 << p.log2_ps) - 1;

  // ldmatrix lane addresses (byte offsets inside a 4 KiB tile, before the per-step chunk xor)
  //   K (non-trans): matrix m = lane/8: token = (m>>1)*8 + lane%8, chunk = 2*ks + (m&1)
  //   V (trans)    : token = (m&1)*8 + lane%8,  chunk = 2*j + (m>>1)
  const int lm = lane >> 3, lr = lane & 7;
  const int k_tok = (lm >> 1) * 8 + lr, k_csel = lm & 1;
  const int v_tok = (lm & 1) * 8 + lr, v_csel = lm >> 1;

  // q rows of one sequence as mma A fragments (the two words of a row this lane owns per 16-column step)
  auto load_q = [&](int s, uint32_t (&dst)[kD / 16][2]) {
    const __nv_bfloat16* qrow = p.q + (int64_t)s * p.q_row_stride + (int64_t)(kv_head * p.group + row) * kD;
#pragma unroll
    for (int ks = 0; ks < kD / 16; ++ks) {
      uint32_t lo = 0, hi = 0;
      if (row_real) {
        lo = *reinterpret_cast<const uint32_t*>(qrow + ks * 16 + quad * 2);
        hi = *reinterpret_cast<const uint32_t*>(qrow + ks * 16 + 8 + quad * 2);
      }
      dst[ks][0] = lo;
      dst[ks][1] = hi;
    }
  };

  // ---- producer cursor: the sequence that holds the tile whose slot ids are fetched next -------------
  int ps = s0, ppre = pre0;
  int pL = seq_kv_len(p.kv_indptr, p.last_page_len, ps, p.log2_ps);
  int pnt = (pL + kTile - 1) / kTile;
  int ppage0 = p.kv_indptr[ps];
  uint32_t qn[kD / 16][2];          // q rows of the sequence the producer is in
  int qn_s = ps;
  load_q(ps, qn);
  // slot id of row (lane & 15) of flat tile ti, fetched one tile ahead of its copies
  auto fetch_slot = [&](int ti) -> int {
    if (ti >= ppre + pnt) {
      do {
        ppre += pnt;
        ++ps;
        pL = seq_kv_len(p.kv_indptr, p.last_page_len, ps, p.log2_ps);
        pnt = (pL + kTile - 1) / kTile;
      } while (ti >= ppre + pnt);
      ppage0 = p.kv_indptr[ps];
      load_q(ps, qn);
      qn_s = ps;
    }
    const int tok = (ti - ppre) * kTile + (lane & 15);
    if (tok >= pL) return -1;
    const int page = __ldg(p.kv_indices + ppage0 + (tok >> p.log2_ps));
    return (page << p.log2_ps) + (tok & ps_mask);
  };
  auto issue = [&](int it, int my_slot) {
    const uint32_t st = warp_smem + (it % kStages) * kStageBytes;
#pragma unroll
    for (int i = 0; i < kTile / 2; ++i) {
      const int r = 2 * i + ld_rsel;
      const int slot = __shfl_sync(0xffffffffu, my_slot, r);
      const bool ok = slot >= 0;
      const int64_t off = ok ? (int64_t)slot * p.kv_slot_stride : 0;
      const uint32_t dst = st + r * kRowBytes + ((ld_chunk ^ (r & 7)) << 4);
      cp_async_16(dst, kbase + off, ok ? 16u : 0u);
      cp_async_16(dst + kTileBytes, vbase + off, ok ? 16u : 0u);
    }
  };

  const int n_my = t_end - t_begin;
  int next_slot = fetch_slot(t_begin);
#pragma unroll
  for (int it = 0; it < kStages - 1; ++it) {
    if (it < n_my) {
      const int cur = next_slot;
      next_slot = (it + 1 < n_my) ? fetch_slot(t_begin + it + 1) : -1;
      issue(it, cur);
    }
    cp_async_commit();
  }

  // ---- consumer cursor ---------------------------------------------------------------------------
  int cs = s0, cpre = pre0;
  int cL = seq_kv_len(p.kv_indptr, p.last_page_len, cs, p.log2_ps);
  int cnt = (cL + kTile - 1) / kTile;
  uint32_t qa[kD / 16][2];
  float o[kD / 8][2];
  float m_run = -INFINITY, l_run = 0.f;
  bool seg_start = true;

  for (int it = 0; it < n_my; ++it) {
    {
      const int nx = it + kStages - 1;
      if (nx < n_my) {
        const int cur = next_slot;
        next_slot = (nx + 1 < n_my) ? fetch_slot(t_begin + nx + 1) : -1;
        issue(nx, cur);
      }
      cp_async_commit();
    }
    const int ti = t_begin + it;
    if (ti >= cpre + cnt) {
      do {
        cpre += cnt;
        ++cs;
        cL = seq_kv_len(p.kv_indptr, p.last_page_len, cs, p.log2_ps);   // L1 hit: the producer cursor was here
        cnt = (cL + kTile - 1) / kTile;
      } while (ti >= cpre + cnt);
      seg_start = true;
    }
    if (seg_start) {
      seg_start = false;
      if (qn_s == cs) {
#pragma unroll
        for (int ks = 0; ks < kD / 16; ++ks) { qa[ks][0] = qn[ks][0]; qa[ks][1] = qn[ks][1]; }
      } else {
        load_q(cs, qa);       // the producer is already more than one sequence ahead (sequences of 1-2 tiles)
      }
#pragma unroll
      for (int j = 0; j < kD / 8; ++j) o[j][0] = o[j][1] = 0.f;
      m_run = -INFINITY;
      l_run = 0.f;
    }
    cp_async_wait<kStages - 1>();
    __syncwarp();

    const uint32_t kst = warp_smem + (it % kStages) * kStageBytes;
    const uint32_t vst = kst + kTileBytes;
    const int tile = ti - cpre;

    // ---- S = Q K^T (16 x 16 tokens) -------------------------------------------------------
    float sc[2][4];
#pragma unroll
    for (int n = 0; n < 2; ++n) sc[n][0] = sc[n][1] = sc[n][2] = sc[n][3] = 0.f;
#pragma unroll
    for (int ks = 0; ks < kD / 16; ++ks) {
      uint32_t b[4];
      ldmatrix_x4(b, kst + k_tok * kRowBytes + (((2 * ks + k_csel) ^ (k_tok & 7)) << 4));
      const uint32_t a[4] = {qa[ks][0], 0u, qa[ks][1], 0u};     // rows 8..15 of the A tile are zero
      mma_bf16_16816(sc[0], a, b[0], b[1]);
      mma_bf16_16816(sc[1], a, b[2], b[3]);
    }
    // ---- online softmax on row `row`, columns tok0 + {0,1,8,9} ---------------------------
    const int tok0 = tile * kTile + quad * 2;
    float x0 = sc[0][0] * p.scale_log2, x1 = sc[0][1] * p.scale_log2;
    float x2 = sc[1][0] * p.scale_log2, x3 = sc[1][1] * p.scale_log2;
    if (tile * kTile + kTile > cL) {
      if (tok0 >= cL) x0 = -INFINITY;
      if (tok0 + 1 >= cL) x1 = -INFINITY;
      if (tok0 + 8 >= cL) x2 = -INFINITY;
      if (tok0 + 9 >= cL) x3 = -INFINITY;
    }
    float mx = fmaxf(fmaxf(x0, x1), fmaxf(x2, x3));
    mx = fmaxf(mx, __shfl_xor_sync(0xffffffffu, mx, 1));
    mx = fmaxf(mx, __shfl_xor_sync(0xffffffffu, mx, 2));
    const float m_new = fmaxf(m_run, mx);  // finite: every tile holds >= 1 valid token
    const float alpha = fast_exp2(m_run - m_new);
    const float p0 = fast_exp2(x0 - m_new), p1 = fast_exp2(x1 - m_new);
    const float p2 = fast_exp2(x2 - m_new), p3 = fast_exp2(x3 - m_new);
    l_run = l_run * alpha + (p0 + p1) + (p2 + p3);
    if (__any_sync(0xffffffffu, m_new != m_run)) {
#pragma unroll
      for (int j = 0; j < kD / 8; ++j) {
        o[j][0] *= alpha;
        o[j][1] *= alpha;
      }
    }
    m_run = m_new;
    const uint32_t pa[4] = {pack_bf16x2(p0, p1), 0u, pack_bf16x2(p2, p3), 0u};
    // ---- O += P V ---------------------------------------------------------------------------
#pragma unroll
    for (int j = 0; j < kD / 16; ++j) {
      uint32_t b[4];
      ldmatrix_x4_trans(b, vst + v_tok * kRowBytes + (((2 * j + v_csel) ^ (v_tok & 7)) << 4));
      float c0[4] = {o[2 * j][0], o[2 * j][1], 0.f, 0.f};
      float c1[4] = {o[2 * j + 1][0], o[2 * j + 1][1], 0.f, 0.f};
      mma_bf16_16816(c0, pa, b[0], b[1]);
      mma_bf16_16816(c1, pa, b[2], b[3]);
      o[2 * j][0] = c0[0]; o[2 * j][1] = c0[1];
      o[2 * j + 1][0] = c1[0]; o[2 * j + 1][1] = c1[1];
    }
    __syncwarp();

    // ---- segment epilogue: last tile of this sequence inside this stream -----------------------------
    const int seg_e = min(t_end, cpre + cnt) - cpre;
    if (tile + 1 == seg_e) {
      const int seg_b = max(t_begin, cpre) - cpre;
      const int slot_id = k_stream + cs;     // unique per (stream, sequence) segment, see header
      const bool whole = (seg_b == 0) && (seg_e == cnt);
      if (lane == 0 && hh == 0 && head_group == 0) {
        if (seg_b == 0) p.seg_first[cs] = slot_id;
        if (seg_e == cnt) p.seg_last[cs] = slot_id;
      }
      float l_tot = l_run + __shfl_xor_sync(0xffffffffu, l_run, 1);
      l_tot += __shfl_xor_sync(0xffffffffu, l_tot, 2);
      if (row_real) {
        const float inv_l = l_tot > 0.f ? 1.f / l_tot : 0.f;
        const int hq = kv_head * p.group + row;
        if (whole) {
          __nv_bfloat16* dst = p.out + (int64_t)cs * p.out_row_stride + (int64_t)hq * kD + quad * 2;
#pragma unroll
          for (int j = 0; j < kD / 8; ++j)
            *reinterpret_cast<uint32_t*>(dst + j * 8) = pack_bf16x2(o[j][0] * inv_l, o[j][1] * inv_l);
        } else {
          float* dst = p.part_o + ((int64_t)slot_id * p.Hq + hq) * kD + quad * 2;
#pragma unroll
          for (int j = 0; j < kD / 8; ++j)
            *reinterpret_cast<float2*>(dst + j * 8) = make_float2(o[j][0] * inv_l, o[j][1] * inv_l);
          if (quad == 0) p.part_lse[(int64_t)slot_id * p.Hq + hq] = l_tot > 0.f ? m_run + fast_log2(l_tot) : -INFINITY;
        }
      }
    }
  }
  cp_async_wait<0>();
}

// Merge pass: one CTA per (sequence, query head); 4 warps split the partial list, a lane owns 4 output
// dims (float4 loads: one coalesced 512-byte row per partial per warp), warps combine through smem.
__global__ void __launch_bounds__(128) decode_merge_kernel(const DecodeParams p) {
  const int s = blockIdx.x, hq = blockIdx.y, lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  __nv_bfloat16* dst = p.out + (int64_t)s * p.out_row_stride + (int64_t)hq * kD;
  griddep_wait();                     // launched with PDL: partials / segment maps come from decode_kernel
  const int L = seq_kv_len(p.kv_indptr, p.last_page_len, s, p.log2_ps);
  if (L == 0) {
    dst[threadIdx.x] = __float2bfloat16(0.f);
    return;
  }
  const int first = p.seg_first[s], last = p.seg_last[s];
  if (first == last) return;  // one segment: written directly by the streaming kernel
  const int n = last - first + 1;
  const float* lse = p.part_lse + (int64_t)first * p.Hq + hq;
  const float4* po = reinterpret_cast<const float4*>(p.part_o + ((int64_t)first * p.Hq + hq) * kD) + lane;
  const int64_t po_stride = (int64_t)p.Hq * kD / 4;
  __shared__ float s_m[4], s_w[4];
  __shared__ float4 s_acc[4][32];
  float mx = -INFINITY;
  for (int i = warp; i < n; i += 4) mx = fmaxf(mx, lse[(int64_t)i * p.Hq]);
  float4 acc = make_float4(0.f, 0.f, 0.f, 0.f);
  float wsum = 0.f;
#pragma unroll 4
  for (int i = warp; i < n; i += 4) {
    const float l = lse[(int64_t)i * p.Hq];
    const float w = l == -INFINITY ? 0.f : fast_exp2(l - mx);
    const float4 v = po[(int64_t)i * po_stride];
    wsum += w;
    acc.x += w * v.x; acc.y += w * v.y; acc.z += w * v.z; acc.w += w * v.w;
  }
  if (lane == 0) { s_m[warp] = mx; s_w[warp] = wsum; }
  s_acc[warp][lane] = acc;
  __syncthreads();
  if (warp == 0) {
    const float M = fmaxf(fmaxf(s_m[0], s_m[1]), fmaxf(s_m[2], s_m[3]));
    float4 r = make_float4(0.f, 0.f, 0.f, 0.f);
    float W = 0.f;
#pragma unroll
    for (int w = 0; w < 4; ++w) {
      const float f = s_m[w] == -INFINITY ? 0.f : fast_exp2(s_m[w] - M);
      const float4 a = s_acc[w][lane];
      W += f * s_w[w];
      r.x += f * a.x; r.y += f * a.y; r.z += f * a.z; r.w += f * a.w;
    }
    const float inv = W > 0.f ? 1.f / W : 0.f;
    uint2 o2;
    o2.x = pack_bf16x2(r.x * inv, r.y * inv);
    o2.y = pack_bf16x2(r.z * inv, r.w * inv);
    *reinterpret_cast<uint2*>(dst + lane * 4) = o2;
  }
}

static int ilog2_exact_d(int x) {
  if (x <= 0 || (x & (x - 1))) return -1;
  int l = 0;
  while ((1 << l) < x) ++l;
  return l;
}

struct DecodeCfg {
  int heads_per_cta, subs, head_groups;
};
static bool decode_cfg(int Hkv, DecodeCfg* c) {
  if (Hkv <= 0) return false;
  if (Hkv >= kWarps) {
    if (Hkv % kWarps) return false;
    *c = {kWarps, 1, Hkv / kWarps};
  } else {
    *c = {Hkv, kWarps / Hkv, 1};  // Hkv in 1..7: subs = floor(8/Hkv) warps per head
  }
  return true;
}

}  // namespace b200

extern "C" size_t b200_decode_workspace_bytes(int32_t B, int32_t Hq, int32_t Hkv, int32_t D) {
  using namespace b200;
  DecodeCfg c;
  if (B <= 0 || Hq <= 0 || D <= 0 || !decode_cfg(Hkv, &c)) return 0;
  const int G = cached_num_sms();
  const size_t parts = (size_t)G * c.subs + (size_t)B;
  size_t bytes = parts * Hq * D * sizeof(float);       // part_o
  bytes += parts * Hq * sizeof(float);                 // part_lse
  bytes += 2 * (size_t)B * sizeof(int32_t);            // seg_first / seg_last
  return (bytes + 255) & ~(size_t)255;
}

extern "C" int b200_decode(const void* q, int64_t q_row_stride, const void* k_cache, const void* v_cache,
                           int64_t kv_slot_stride, const int32_t* kv_indptr, const int32_t* kv_indices,
                           const int32_t* last_page_len, void* out, int64_t out_row_stride, void* workspace,
                           size_t workspace_bytes, int32_t B, int32_t Hq, int32_t Hkv, int32_t D, int32_t page_size,
                           float sm_scale, int32_t dtype, b200_stream_t stream) {
  using namespace b200;
  if (B < 0) return fail("b200_decode: B < 0");
  if (B == 0) return 0;
  if (D != kD) return fail("b200_decode: head_dim %d not built (round 1 builds D=128)", D);
  if (dtype != B200_DTYPE_BF16) return fail("b200_decode: dtype %d not built (bf16 only)", dtype);
  DecodeCfg c;
  if (!decode_cfg(Hkv, &c)) return fail("b200_decode: unsupported num_kv_heads %d", Hkv);
  if (Hq % Hkv) return fail("b200_decode: Hq %d not a multiple of Hkv %d", Hq, Hkv);
  const int group = Hq / Hkv;
  if (group > 8) return fail("b200_decode: GQA group %d > 8 not built", group);
  const int log2_ps = ilog2_exact_d(page_size);
  if (log2_ps < 0) return fail("b200_decode: page_size %d is not a power of two", page_size);
  if (!q || !k_cache || !v_cache || !kv_indptr || !kv_indices || !last_page_len || !out || !workspace)
    return fail("b200_decode: null pointer");
  if (((uintptr_t)k_cache | (uintptr_t)v_cache | (uintptr_t)workspace) & 15 || (kv_slot_stride & 7))
    return fail("b200_decode: k_cache / v_cache / workspace must be 16-byte aligned");
  if (((uintptr_t)q & 3) || (q_row_stride & 1)) return fail("b200_decode: q must be 4-byte aligned");
  if (((uintptr_t)out & 7) || (out_row_stride & 3)) return fail("b200_decode: out must be 8-byte aligned");
  const size_t need = b200_decode_workspace_bytes(B, Hq, Hkv, D);
  if (workspace_bytes < need) return fail("b200_decode: workspace %zu < %zu bytes", workspace_bytes, need);

  const int G = cached_num_sms();
  DecodeParams p;
  p.q = (const __nv_bfloat16*)q;
  p.k_cache = (const __nv_bfloat16*)k_cache;
  p.v_cache = (const __nv_bfloat16*)v_cache;
  p.kv_indptr = kv_indptr;
  p.kv_indices = kv_indices;
  p.last_page_len = last_page_len;
  p.out = (__nv_bfloat16*)out;
  const size_t parts = (size_t)G * c.subs + (size_t)B;
  p.part_o = (float*)workspace;
  p.part_lse = p.part_o + parts * Hq * D;
  p.seg_first = (int32_t*)(p.part_lse + parts * Hq);
  p.seg_last = p.seg_first + B;
  p.q_row_stride = q_row_stride;
  p.out_row_stride = out_row_stride;
  p.kv_slot_stride = kv_slot_stride;
  p.B = B; p.Hq = Hq; p.Hkv = Hkv; p.group = group; p.log2_ps = log2_ps;
  p.heads_per_cta = c.heads_per_cta;
  p.subs = c.subs;
  p.scale_log2 = sm_scale * kLog2e;

  static std::atomic<bool> attr_done[64];
  if (cudaError_t e = ensure_dyn_smem(decode_kernel, kDecodeSmem, attr_done); e != cudaSuccess)
    return fail("b200_decode: cudaFuncSetAttribute: %s", cudaGetErrorString(e));
  cudaStream_t st = (cudaStream_t)stream;
  decode_kernel<<<dim3(G, c.head_groups), kWarps * 32, kDecodeSmem, st>>>(p);
  if (int rc = check_launch("b200_decode(stream)")) return rc;
  cudaError_t e = launch_pdl(decode_merge_kernel, dim3(B, Hq), dim3(128), 0, st, p);
  if (e != cudaSuccess) return fail("b200_decode(merge): %s", cudaGetErrorString(e));
  return check_launch("b200_decode(merge)");
}
