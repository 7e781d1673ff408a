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
// (Design "A".  Round 2 also built and measured design "B": S pulled into registers so that QK(j+1) runs under
// softmax(j), P through shared memory, 3-slot ring -- git 7f19014.  Same box, 16 x 4096 Hq32/Hkv8: B 0.63 of bf16 peak
// with 48 % tensor pipe, A 0.66, A with the packed-f32x2 softmax below 0.69 with 55 %: profiles/r2_s9_*, r2_s10_*.
// B's SS-mode PV reads P and V from shared memory and saturates the tensor cores' shared-memory port; in both
// designs the time goes into the ~700 instructions a softmax warp runs per 128 x 128 tile.)
// Feed: Q always by TMA (tiled 3-D boxes); K/V tiles (128 tokens) by TMA tile::gather4 through the page table when
// page_size >= 16 and by 16-byte cp.async gathers for the reference's token-granular pools (page_size 1: gather4
// there is 64 single-row instructions per tile, each a ~90 clk elect loop on the uniform datapath: measured 0.69 ->
// 0.46 of bf16 peak, profiles/r2_s4_*).  Either way the bytes land in the canonical SWIZZLE_128B layout (two
// [128 x 64] blocks of 128-byte rows, chunk ^= row & 7).
// O lives in TMEM across the KV loop; a row is rescaled in place (tcgen05.ld/st) only when its
// running max moves.
#include "common.cuh"
#include "prefill_params.cuh"
#include "sm100.cuh"
#include "tma.cuh"

namespace b200 {

using namespace sm100;

constexpr int kTD = 128;                       // head dim
constexpr int kTM = 128;                       // packed query rows per CTA
constexpr int kTN = 128;                       // KV tokens per tile
constexpr int kBlkBytes = 128 * 128;           // one [128 rows x 64 elem] swizzled block = 16 KiB
constexpr int kTileBytes2 = 2 * kBlkBytes;     // a full [128 x 128] bf16 operand tile = 32 KiB

// =====================================================================================================
// Warp-specialised pipeline, two query tiles per CTA in ping-pong.
//
//   warps 0-3   softmax group 0  (rows   0..127 of the CTA's 256 packed rows; thread == row == TMEM lane)
//   warps 4-7   softmax group 1  (rows 128..255)
//   warp  8     MMA issuer (one thread): S_t = Q_t K^T, O_t += P_t V, t = 0,1
//   warps 9-10  producers: TMA for Q (and for K/V at page_size >= 16), cp.async gathers for K/V at page_size < 16,
//               into a 4-slot ring of SWIZZLE_128B tiles (K slots 0/2, V slots 1/3), released by tcgen05.commit
//   warp  11    scheduler of the persistent work list
//
// TMEM (512 columns): S0 [0,128) S1 [128,256) O0 [256,384) O1 [384,512).  P_t (bf16) is written by
// tcgen05.st over the first 64 columns of S_t and consumed as the TMEM A operand of the PV MMA, so P
// never touches shared memory.  Issue order per KV tile j:  PV0(j) QK0(j+1) PV1(j) QK1(j+1): while
// softmax group 0 works on S0(j+1) the tensor core runs PV1(j)/QK1(j+1) and vice versa.  The two
// query tiles share every K/V tile, which halves L2->smem traffic per FLOP (a 128-row tile alone is
// L2-bandwidth bound at ~40 B/clk/SM).
// O rows are rescaled lazily: only when the row max grew by more than 2^8 since the scale in use
// (exact: O and l carry the same stale scale until the final division).
// =====================================================================================================
constexpr int kWsThreads = 384;   // 3 warpgroups: softmax0, softmax1, {MMA, 2 loaders, 1 spare}
constexpr int kRing = 4;
constexpr int kWsOffQ = 0;                           // 2 x 32 KiB
constexpr int kWsOffRing = 2 * kTileBytes2;          // kRing x 32 KiB
constexpr float kLazyTau = 8.0f;

// ---- packed f32x2 helpers (FFMA2 / FADD2 on sm_100).  The arithmetic forms are `asm volatile`: together with the
// bar.warp.sync fences they pin the instruction ORDER of the softmax software pipeline; the mov.b64 pack / unpack forms
// are plain register-pair aliases and cost no instruction.
#ifndef B200_PREFILL_PACKED
#define B200_PREFILL_PACKED 1
#endif
constexpr bool kPackedSoftmax = B200_PREFILL_PACKED != 0;
#ifndef B200_PREFILL_UNIMASK
#define B200_PREFILL_UNIMASK 1
#endif
constexpr bool kUnifiedMask = B200_PREFILL_UNIMASK != 0;   // masked tiles share the unmasked tile body (see softmax_tile)
// kLeanEpilogue: output-row addresses as one 64-bit base per item + 32-bit deltas per store, 1/l scaling with packed
// FMUL2.  On bench-shaped prompts (1..8 KV tiles per item) the epilogue was 28 % of the softmax warps' time, most of it
// 64-bit address arithmetic of the 16 stores per thread (profiles/r2s22_tree_prefill_short_roles.txt).
#ifndef B200_PREFILL_LEANEPI
#define B200_PREFILL_LEANEPI 1
#endif
constexpr bool kLeanEpilogue = B200_PREFILL_LEANEPI != 0;
__device__ __forceinline__ uint64_t pack2(uint32_t lo, uint32_t hi) {
  uint64_t r;
  asm("mov.b64 %0, {%1, %2};\n" : "=l"(r) : "r"(lo), "r"(hi));
  return r;
}
__device__ __forceinline__ void unpack2(uint64_t v, uint32_t& lo, uint32_t& hi) {
  asm("mov.b64 {%0, %1}, %2;\n" : "=r"(lo), "=r"(hi) : "l"(v));
}
__device__ __forceinline__ uint64_t fma2_ordered(uint64_t a, uint64_t b, uint64_t c) {
  uint64_t r;
  asm volatile("fma.rn.f32x2 %0, %1, %2, %3;\n" : "=l"(r) : "l"(a), "l"(b), "l"(c));
  return r;
}
__device__ __forceinline__ uint64_t mul2(uint64_t a, uint64_t b) {
  uint64_t r;
  asm("mul.rn.f32x2 %0, %1, %2;\n" : "=l"(r) : "l"(a), "l"(b));
  return r;
}
__device__ __forceinline__ uint64_t add2_ordered(uint64_t a, uint64_t b) {
  uint64_t r;
  asm volatile("add.rn.f32x2 %0, %1, %2;\n" : "=l"(r) : "l"(a), "l"(b));
  return r;
}

// One softmax step of one row (thread) over a 128-column S tile held in TMEM.
// MASK is a template parameter so the (rare) masked tiles do not tax the common path with
// predicated-off instructions (nvcc if-converts a runtime `if` around the element loop).
template <bool MASK, int NCH = 4>
__device__ __forceinline__ void softmax_tile(const uint32_t tS, const uint32_t tO, const int nvalid, const bool first_pv_done,
                                             const float scale_log2, float& m_used, float& l_run, const uint32_t bar_p0,
                                             const uint32_t bar_p1, const bool runtime_mask = false) {
  // NCH (masked tiles only, tile-uniform): number of 32-column chunks of S that hold any valid column of this
  // query tile.  The MMA issuer trims QK to N = 32*NCH and PV to K = 32*NCH, so chunks >= NCH are neither
  // loaded, exponentiated nor stored (their S/P columns are stale and never read).  nvalid <= 32*NCH.
  static_assert(MASK || NCH == 4, "only masked tiles are trimmed");
  uint32_t sv[NCH][32];
  float mx0 = -INFINITY, mx1 = -INFINITY, mx2 = -INFINITY, mx3 = -INFINITY;
  // invalid columns -> -inf.  MASK: compile-time masked instance.  runtime_mask (kUnifiedMask): masked tiles (causal
  // diagonal, ragged sequence end) mask here and then run the SAME packed pipeline as the unmasked ones (exp2(-inf) = 0):
  // one copy of the 450-instruction tile body instead of two (the kernel lives on instruction fetch: `no_inst` is 8-19 %
  // of the softmax warps' stalls).  The branch is tile-uniform; the warp barriers inside keep ptxas from if-converting
  // the compare / select pairs into predicated-off instructions of the common path.
  auto mask_chunks = [&](int c0, int c1) {
    if (MASK) {
#pragma unroll
      for (int c = c0; c < c1; ++c)
#pragma unroll
        for (int e = 0; e < 32; ++e)
          if (c * 32 + e >= nvalid) sv[c][e] = 0xff800000u;   // -inf
    } else if (runtime_mask) {
      sched_fence();
#pragma unroll
      for (int c = c0; c < c1; ++c)
#pragma unroll
        for (int e = 0; e < 32; ++e)
          if (c * 32 + e >= nvalid) sv[c][e] = 0xff800000u;   // -inf
      sched_fence();
    }
  };
  auto max_chunks = [&](int c0, int c1) {
#pragma unroll
    for (int c = c0; c < c1; ++c) {
#pragma unroll
      for (int e = 0; e < 32; e += 4) {
        mx0 = fmaxf(mx0, __uint_as_float(sv[c][e]));
        mx1 = fmaxf(mx1, __uint_as_float(sv[c][e + 1]));
        mx2 = fmaxf(mx2, __uint_as_float(sv[c][e + 2]));
        mx3 = fmaxf(mx3, __uint_as_float(sv[c][e + 3]));
      }
    }
  };
  // (loading the chunks in two halves so that the maximum of the first runs under the loads of the second was
  //  measured: -2..3 % on every shape, profiles/r2_s22_prefill_splitload_ab.txt)
#pragma unroll
  for (int c = 0; c < NCH; ++c) tmem_ld_32x32b_x32(tS + c * 32, sv[c]);
  tmem_wait_ld();
  mask_chunks(0, NCH);
  max_chunks(0, NCH);
  const float mx = fmaxf(fmaxf(mx0, mx1), fmaxf(mx2, mx3)) * scale_log2;   // scale > 0: max commutes
  float alpha = 1.f;
  bool grow = false;
  if (l_run == 0.f) {                      // nothing accumulated yet: adopt this tile's max
    if (mx != -INFINITY) m_used = mx;
  } else if (mx > m_used + kLazyTau) {
    grow = true;
    alpha = fast_exp2(m_used - mx);
    m_used = mx;
  }
  const bool rescale = first_pv_done && __any_sync(0xffffffffu, grow);   // warp-uniform, rare (lazy threshold 2^8)
  const float neg_m = -m_used;
  float s0 = 0.f, s1 = 0.f;
  // The O rescale sits AFTER the exponentials in its own (rare) branch: with the rescale ahead of them ptxas
  // hoists all 128 ex2 above it and spills 15 registers around the 32-register O chunk on the hot path.
  if constexpr (!MASK && kPackedSoftmax) {
    // Unmasked tiles (the common case): explicit software pipeline over the 64 element PAIRS of the row in groups of
    // 4 pairs, with packed f32x2 arithmetic (FFMA2 / FADD2: half the FMA-pipe instructions of the scalar form).
    // Per group two instruction queues are interleaved one to one:
    //   M  the 8 MUFU.EX2 of the group
    //   F  the scale FFMA2 of the NEXT group, then row sum (FADD2) and bf16 pack of the PREVIOUS group
    // so every MUFU (one per 8 clk per scheduler) has independent FMA / ALU work beside it.
    const uint64_t sc2 = pack2(__float_as_uint(scale_log2), __float_as_uint(scale_log2));
    const uint64_t nm2 = pack2(__float_as_uint(neg_m), __float_as_uint(neg_m));
    uint64_t sp[64];
#pragma unroll
    for (int q = 0; q < 64; ++q) sp[q] = pack2(sv[q >> 4][(2 * q) & 31], sv[q >> 4][(2 * q + 1) & 31]);
    uint64_t acc_a = pack2(0u, 0u), acc_b = pack2(0u, 0u);
    uint32_t pw[16];       // packed P of one 32-column chunk
#pragma unroll
    for (int q = 0; q < 4; ++q) sp[q] = fma2_ordered(sp[q], sc2, nm2);
#pragma unroll
    for (int g = 0; g <= 16; ++g) {
      sched_fence();
#pragma unroll
      for (int k = 0; k < 12; ++k) {
        if (g < 16 && k < 8) {                                           // ---- M
          const int q = 4 * g + (k >> 1);
          uint32_t lo, hi;
          unpack2(sp[q], lo, hi);
          if (k & 1) hi = __float_as_uint(exp2_ordered(__uint_as_float(hi)));
          else lo = __float_as_uint(exp2_ordered(__uint_as_float(lo)));
          sp[q] = pack2(lo, hi);
        }
        if (k < 4) {                                                     // ---- F
          if (g + 1 < 16) sp[4 * (g + 1) + k] = fma2_ordered(sp[4 * (g + 1) + k], sc2, nm2);
        } else if (k < 8) {
          if (g >= 1) {
            const int q = 4 * (g - 1) + (k - 4);
            if (k & 1) acc_b = add2_ordered(acc_b, sp[q]);
            else acc_a = add2_ordered(acc_a, sp[q]);
          }
        } else if (g >= 1) {
          const int q = 4 * (g - 1) + (k - 8);
          uint32_t lo, hi;
          unpack2(sp[q], lo, hi);
          pw[q & 15] = pack_bf16x2_ordered(__uint_as_float(lo), __uint_as_float(hi));
        }
      }
      if (g >= 1 && ((g - 1) & 3) == 3) {
        const int c = (g - 1) >> 2;
        tmem_st_32x32b_x16(tS + c * 16, pw);          // P: packed bf16x2 over S
        if (!rescale && (c == 1 || c == 3)) {         // hand P over in two halves: PV starts on the first
          tmem_wait_st();
          tc_fence_before();
          mbar_arrive(c == 1 ? bar_p0 : bar_p1);
        }
      }
    }
    {
      uint32_t lo, hi;
      unpack2(add2_ordered(acc_a, acc_b), lo, hi);
      s0 = __uint_as_float(lo);
      s1 = __uint_as_float(hi);
    }
  } else if constexpr (!MASK) {
    // Unmasked tiles (the common case): explicit software pipeline over the 128 elements.  The scale FMA runs
    // kAhead elements ahead of its MUFU.EX2, the row-sum add and the bf16 pack kAhead elements behind it, so every
    // MUFU (one per 8 clk per scheduler) has independent FMA-pipe work beside it.
    constexpr int kAhead = 12;
    uint32_t pw[64];
#pragma unroll
    for (int e = 0; e < kAhead; ++e)
      sv[e >> 5][e & 31] = __float_as_uint(fma_ordered(__uint_as_float(sv[e >> 5][e & 31]), scale_log2, neg_m));
#pragma unroll
    for (int e = 0; e < 128 + kAhead; ++e) {
      if ((e & 7) == 0) sched_fence();
      if (e < 128) sv[e >> 5][e & 31] = __float_as_uint(exp2_ordered(__uint_as_float(sv[e >> 5][e & 31])));
      if (e + kAhead < 128) {
        const int f = e + kAhead;
        sv[f >> 5][f & 31] = __float_as_uint(fma_ordered(__uint_as_float(sv[f >> 5][f & 31]), scale_log2, neg_m));
      }
      if (e >= kAhead) {
        const int d = e - kAhead;
        if (d & 1) {
          s1 = add_ordered(s1, __uint_as_float(sv[d >> 5][d & 31]));
          pw[d >> 1] = pack_bf16x2_ordered(__uint_as_float(sv[(d - 1) >> 5][(d - 1) & 31]), __uint_as_float(sv[d >> 5][d & 31]));
          if ((d & 31) == 31) {
            const int c = d >> 5;
            tmem_st_32x32b_x16(tS + c * 16, *reinterpret_cast<uint32_t(*)[16]>(&pw[c * 16]));   // P: packed bf16x2 over S
            if (!rescale && (c == 1 || c == 3)) {     // hand P over in two halves: PV starts on the first
              tmem_wait_st();
              tc_fence_before();
              mbar_arrive(c == 1 ? bar_p0 : bar_p1);
            }
          }
        } else {
          s0 = add_ordered(s0, __uint_as_float(sv[d >> 5][d & 31]));
        }
      }
    }
  } else {
#pragma unroll
  for (int c = 0; c < NCH; ++c) {
    uint32_t pw[16];
#pragma unroll
    for (int e = 0; e < 16; ++e) {
      const float a = exp2_fma(__uint_as_float(sv[c][2 * e]), scale_log2, neg_m);
      const float b = exp2_fma(__uint_as_float(sv[c][2 * e + 1]), scale_log2, neg_m);
      s0 += a;
      s1 += b;
      pw[e] = pack_bf16x2(a, b);
    }
    tmem_st_32x32b_x16(tS + c * 16, pw);            // P: 64 packed bf16x2 columns over S
    constexpr int kFirst = NCH > 1 ? 1 : 0;         // hand P over in two halves: PV starts on the first
    if (!rescale && (c == kFirst || c == NCH - 1)) { // while the second is still in the exp unit
      tmem_wait_st();
      tc_fence_before();
      if (c == kFirst) mbar_arrive(bar_p0);
      if (c == NCH - 1) mbar_arrive(bar_p1);
    }
  }
  }
  if (rescale) {                                    // O must carry the new scale before PV(j) accumulates into it
#pragma unroll
    for (int c = 0; c < 4; ++c) {
      uint32_t ov[32];
      tmem_ld_32x32b_x32(tO + c * 32, ov);
      tmem_wait_ld();
#pragma unroll
      for (int e = 0; e < 32; ++e) ov[e] = __float_as_uint(__uint_as_float(ov[e]) * alpha);
      tmem_st_32x32b_x32(tO + c * 32, ov);
    }
    tmem_wait_st();
    tc_fence_before();
    mbar_arrive(bar_p0);
    mbar_arrive(bar_p1);
  }
  l_run = l_run * alpha + (s0 + s1);
}


// =====================================================================================================
// The pipeline is PERSISTENT.
//
// One CTA per SM walks a device-built work list instead of paying launch prologue + pipeline ramp
// per 256-row tile (bench-shaped prompts are 100..1024 tokens: a tile owns only 1..8 KV tiles, so the
// per-CTA fixed cost was ~2/3 of the run time).
//   * prefill_plan_kernel (1 CTA) lists every (sequence, 256-row query tile) pair, counting-sorted
//     by its number of KV tiles, heaviest first, and zeroes the work counter.
//   * warp 11 of every CTA is the scheduler: atomicAdd on the counter, (pair, kv head) -> geometry,
//     published one item AHEAD through a 4-entry shared-memory ring (mbarrier full/empty), so
//     the heaviest items start first and the tail is made of the lightest (LPT, dynamic).
//   * loaders, MMA issuer and softmax groups each run their own item loop with running mbarrier
//     phases; the loaders prefetch the next item's Q and K/V while the softmax groups are still in
//     the current item's last tile and epilogue (Q buffer hand-over: tcgen05.commit -> bar_qfree).
//   * the softmax warps stage output rows through a 7th 32 KiB tile so that stores are full 128-byte lines; the
//     loaders fill Q with 32-bit offsets (~8 instructions per cp.async) and load the next tile's page ids ahead
//     of the current tile's fills (session 38: +8..10 % on bench-shaped prompts, +3 % on 16 x 4096).
//   * TMEM hazards across items need no extra barrier: QK_t(0) of the next item is issued after
//     PV_t(last) on the in-order MMA pipe, and PV_t(0) waits for P_t(0), which group t only
//     produces after its epilogue has read O_t out (tcgen05.wait::ld precedes the arrive).
// =====================================================================================================
constexpr int kItemRing = 4;
constexpr int kPlanThreads = 1024;
constexpr int kPlanBuckets = 1024;
constexpr int kWorkHeaderInts = 16;     // [0] work counter, [1] number of (sequence, tile) pairs
// Dynamic shared memory: 2 Q tiles + 4-slot K/V ring + one 32 KiB tile the softmax warps stage output rows in.
constexpr int kWsOffStage = kWsOffRing + kRing * kTileBytes2;   // persistent kernel only
constexpr int kPersistSmem = kWsOffStage + kTileBytes2 + 1024;

struct ItemDesc {
  int s, kvh, j0, Lq, Lkv, q0, page0, valid;
};

struct ItemGeo {
  int toks0, toks1, kvhi0, kvhi1, nt0, nt1, J, kv_hi_all;
};

__device__ __forceinline__ ItemGeo item_geometry(const ItemDesc& d, const int tok_per_tile, const int causal) {
  ItemGeo g;
  g.toks0 = max(0, min(tok_per_tile, d.Lq - d.j0));
  g.toks1 = max(0, min(tok_per_tile, d.Lq - d.j0 - tok_per_tile));
  g.kvhi0 = causal ? min(d.Lkv, d.Lkv - d.Lq + d.j0 + g.toks0) : d.Lkv;
  g.kvhi1 = causal ? min(d.Lkv, d.Lkv - d.Lq + d.j0 + tok_per_tile + g.toks1) : d.Lkv;
  if (g.kvhi0 < 0 || g.toks0 == 0) g.kvhi0 = 0;
  if (g.kvhi1 < 0 || g.toks1 == 0) g.kvhi1 = 0;
  g.nt0 = (g.kvhi0 + kTN - 1) / kTN;
  g.nt1 = (g.kvhi1 + kTN - 1) / kTN;
  g.J = max(g.nt0, g.nt1);
  g.kv_hi_all = max(g.kvhi0, g.kvhi1);
  return g;
}

// Work list for the persistent kernel.  Sequences are counting-sorted by KV length (longest first); the
// (sequence, query tile) pairs of one sequence stay CONTIGUOUS, latest (heaviest) query tile first.
// Consecutive work items are therefore the KV heads of neighbouring query tiles of ONE sequence: the ~148
// items in flight share that sequence's K/V through L2 (a list sorted purely by weight interleaves all
// sequences and turns the kernel DRAM-bound: measured 0.60 -> 0.53 of peak on 16 x 4096), while heavy work
// still starts first and the tail of the run is made of the shortest sequences' lightest tiles.
// Order among sequences of equal length bucket is arbitrary (atomics): it only affects scheduling, every
// item is computed independently, so results do not depend on it.
__global__ void __launch_bounds__(kPlanThreads, 1) prefill_plan_kernel(const PrefillParams p, const int tok_per_cta,
                                                                        int* __restrict__ work, int2* __restrict__ items,
                                                                        const int max_pairs) {
  __shared__ int s_hist[kPlanBuckets];
  __shared__ int s_scan[kPlanThreads / 32 + 1];
  const int tid = threadIdx.x;
  griddep_launch_dependents();   // the persistent grid may set up (barriers, TMEM) while the list is built
  s_hist[tid] = 0;
  __syncthreads();
  auto bucket_of = [&](int Lkv) { return min((Lkv + kTN - 1) / kTN, kPlanBuckets - 1); };
  for (int i = tid; i < p.B; i += kPlanThreads) {
    const int Lq = p.qo_indptr[i + 1] - p.qo_indptr[i];
    const int nq = (Lq + tok_per_cta - 1) / tok_per_cta;
    if (nq > 0) atomicAdd(&s_hist[bucket_of(seq_kv_len(p.kv_indptr, p.last_page_len, i, p.log2_ps))], nq);
  }
  __syncthreads();
  // thread t owns bucket (kPlanBuckets - 1 - t): an inclusive scan in thread order is a scan longest-first
  const int mine = s_hist[kPlanBuckets - 1 - tid];
  int total;
  const int incl = block_inclusive_scan<kPlanThreads>(mine, s_scan, &total);
  __syncthreads();
  s_hist[kPlanBuckets - 1 - tid] = incl - mine;   // start offset of my bucket
  __syncthreads();
  for (int i = tid; i < p.B; i += kPlanThreads) {
    const int Lq = p.qo_indptr[i + 1] - p.qo_indptr[i];
    const int nq = (Lq + tok_per_cta - 1) / tok_per_cta;
    if (nq <= 0) continue;
    const int pos = atomicAdd(&s_hist[bucket_of(seq_kv_len(p.kv_indptr, p.last_page_len, i, p.log2_ps))], nq);
    for (int k = 0; k < nq; ++k)
      if (pos + k < max_pairs) items[pos + k] = make_int2(i, nq - 1 - k);
  }
  if (tid == 0) {
    work[0] = 0;
    work[1] = min(total, max_pairs);
  }
}

__global__ void __launch_bounds__(kWsThreads, 1) prefill_sm100_persist_kernel(const PrefillParams p, const int log2_g,
                                                                               int* __restrict__ work,
                                                                               const int2* __restrict__ items,
                                                                               const __grid_constant__ CUtensorMap tm_q,
                                                                               const __grid_constant__ CUtensorMap tm_k,
                                                                               const __grid_constant__ CUtensorMap tm_v) {
  extern __shared__ uint8_t smem_raw[];
  __shared__ uint32_t s_tmem;
  __shared__ __align__(8) uint64_t s_bar[2 + 2 * kRing + 8 + 2 * kItemRing];
  __shared__ __align__(16) ItemDesc s_items[kItemRing];
  __shared__ __align__(16) int s_slots[2][2][2][32];   // cp.async producers: [tile parity][warp][row parity][i]

  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const bool tma_kv = p.log2_ps >= 4;                  // K/V tiles by TMA (page_size >= 16) or by cp.async gathers
  const int G = 1 << log2_g;
  const int tok_per_tile = kTM >> log2_g;
  const int tok_per_cta = 2 * tok_per_tile;
  // Trimmed last tiles: 32-column chunks of KV tile j that hold columns below the query tile's kv_hi.  Bench-shaped
  // prompts (100..1024 tokens) own 1..8 KV tiles per item and the last one of each query tile is only
  // partly valid (causal diagonal of a 32/64-token packed tile, ragged sequence end): QK, softmax and PV
  // run on 32*nch columns instead of 128 (8-12 % of all chunk steps on the bench prompts).
  // Compile-time switches, set from a same-box A/B on B200 of a build that had them as run-time flags
  // (profiles/r1_s38_prefill_flags_ab.txt; bench prompts Hq32 / uniform 16x4096 Hq32, fraction of bf16 peak):
  //   none 0.303 / 0.633   trim 0.298 / 0.640   staged 0.319 / 0.650   lean_q 0.326 / 0.648   all three 0.291 / 0.670
  // Trimming costs more than it saves on short prompts (four masked-softmax instances; the kernel is ~7000 SASS
  // instructions and every extra hot path shows), so it stays off; dead branches are removed by the compiler.
  constexpr bool trim = false, staged = true, lean_q = true;
  auto chunks_of = [&](int kv_hi, int j) { return trim ? min(4, (kv_hi - j * kTN + 31) >> 5) : 4; };

  const uint32_t smem = (smem_u32(smem_raw) + 1023u) & ~1023u;
  const uint32_t bar0 = smem_u32(&s_bar[0]);
  const uint32_t bar_q = bar0, bar_qfree = bar0 + 8u;
  auto bar_full = [&](int slot) { return bar0 + 8u * (2 + slot); };
  auto bar_empty = [&](int slot) { return bar0 + 8u * (2 + kRing + slot); };
  auto bar_s = [&](int t) { return bar0 + 8u * (2 + 2 * kRing + t); };
  auto bar_p = [&](int t, int half) { return bar0 + 8u * (2 + 2 * kRing + 2 + 2 * t + half); };
  auto bar_o = [&](int t) { return bar0 + 8u * (2 + 2 * kRing + 6 + t); };
  auto bar_ifull = [&](int slot) { return bar0 + 8u * (2 + 2 * kRing + 8 + slot); };
  auto bar_iempty = [&](int slot) { return bar0 + 8u * (2 + 2 * kRing + 8 + kItemRing + slot); };
  constexpr int kItemConsumers = 11;   // 8 softmax warps + MMA warp + 2 TMA warps (lane 0 of each arrives)

  if (tid == 0) {
    mbar_init(bar_q, 1);              // one arrive.expect_tx by the TMA warp + the transaction bytes
    mbar_init(bar_qfree, 1);
    for (int i = 0; i < kRing; ++i) {
      mbar_init(bar_full(i), tma_kv ? 1 : 64);   // TMA: one arrive.expect_tx + bytes; cp.async: the 64 producer threads
      mbar_init(bar_empty(i), 1);
    }
    for (int t = 0; t < 2; ++t) {
      mbar_init(bar_s(t), 1);
      mbar_init(bar_p(t, 0), 128);
      mbar_init(bar_p(t, 1), 128);
      mbar_init(bar_o(t), 1);
    }
    for (int i = 0; i < kItemRing; ++i) {
      mbar_init(bar_ifull(i), 1);
      mbar_init(bar_iempty(i), kItemConsumers);
    }
    fence_mbar_init();
  }
  if (warp == 0) {
    tmem_alloc(smem_u32(&s_tmem), 512);
    tmem_relinquish();
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem = s_tmem;

  // every consumer warp: wait for item `it`, copy it to registers, release the ring entry
  auto take_item = [&](int it) -> ItemDesc {
    const int islot = it % kItemRing;
    mbar_wait(bar_ifull(islot), (it / kItemRing) & 1);
    const ItemDesc d = s_items[islot];
    __syncwarp();
    if (lane == 0) mbar_arrive(bar_iempty(islot));
    return d;
  };

  if (warp == 11) {
    // =================================== scheduler =========================================================
    asm volatile("setmaxnreg.dec.sync.aligned.u32 96;\n");
    griddep_wait();                // launched with PDL: the work list comes from prefill_plan_kernel
    const int n_total = work[1] * p.Hkv;
    for (int it = 0;; ++it) {
      const int islot = it % kItemRing;
      mbar_wait(bar_iempty(islot), ((it / kItemRing) & 1) ^ 1);
      int valid = 0;
      if (lane == 0) {
        ItemDesc d;
        const int w = atomicAdd(&work[0], 1);
        d.valid = w < n_total ? 1 : 0;
        d.s = d.kvh = d.j0 = d.Lq = d.Lkv = d.q0 = d.page0 = 0;
        if (d.valid) {
          const int pair = w / p.Hkv;
          const int2 sq = items[pair];
          d.s = sq.x;
          d.kvh = w - pair * p.Hkv;
          d.j0 = sq.y * tok_per_cta;
          d.q0 = p.qo_indptr[sq.x];
          d.Lq = p.qo_indptr[sq.x + 1] - d.q0;
          d.Lkv = seq_kv_len(p.kv_indptr, p.last_page_len, sq.x, p.log2_ps);
          d.page0 = p.kv_indptr[sq.x];
        }
        s_items[islot] = d;
        mbar_arrive(bar_ifull(islot));
        valid = d.valid;
      }
      valid = __shfl_sync(0xffffffffu, valid, 0);
      if (!valid) break;
    }
  } else if (warp >= 9) {
    // =================================== producers (warps 9, 10) ==========================================
    // Q, always TMA: four tiled 3-D boxes per item ([64 d] x [G heads] x [128/G tokens]: two query tiles x two
    // 64-column halves), issued by one thread; rows past the sequence / past T are other tokens or zero fill and their
    // outputs are never stored.
    // K/V ring order: K(0), then K(j+1) ahead of V(j) -- QK(j+1) is issued a whole tile period before PV(j).
    //   * page_size >= 16: TMA.  Warp 9 stages the K tiles, warp 10 the V tiles, tile::gather4 through the page table
    //     (lane l owns tile rows 4l..4l+3: 2 instructions per lane per tile; completion counted in bytes on the ring's
    //     mbarriers).  Rows past the end of the sequence re-read its last valid slot (finite data; their S columns are
    //     masked to -inf, their P is exactly 0).
    //   * page_size < 16 (the reference's token-granular mode, page_size 1): 16-byte cp.async gathers by both warps,
    //     zero fill past the end.  MEASURED: gather4 needs 64 instructions per tile, each a 14-instruction elect loop
    //     trip on the uniform datapath (~90 clk); two producer warps were 85 % busy issuing and the kernel fell from
    //     0.69 to 0.46 of bf16 peak on 16 x 4096 (profiles/r2_s4_*gather4*).  2048 LDGSTS per tile cost ~8 clk per warp
    //     instruction, a twentieth of that, so a per-token gather stays on the LSU path; the bytes are laid out exactly as
    //     TMA SWIZZLE_128B would write them.
    asm volatile("setmaxnreg.dec.sync.aligned.u32 96;\n");
    const int ps_mask = (1 << p.log2_ps) - 1;
    const int lw = warp - 9;
    const uint32_t smem_q = smem + kWsOffQ, smem_ring = smem + kWsOffRing;
    if (lane == 0) {
      if (lw == 0) tma::prefetch_desc(&tm_q);
      if (tma_kv) tma::prefetch_desc(lw ? &tm_v : &tm_k);
    }
    int qn = 0;
    if (tma_kv) {
      const CUtensorMap* tm = lw ? &tm_v : &tm_k;
      int fill = lw;                                   // running ring-fill index of this warp's next tile
      for (int it = 0;; ++it) {
        const ItemDesc d = take_item(it);
        if (!d.valid) break;
        const ItemGeo g = item_geometry(d, tok_per_tile, p.causal);
        if (g.J == 0) continue;
        const int last_tok = g.kv_hi_all - 1;
        const int last_slot = (__ldg(p.kv_indices + d.page0 + (last_tok >> p.log2_ps)) << p.log2_ps) + (last_tok & ps_mask);
        int sl[4], sn[4];
        auto ldg_slots = [&](int tile, int (&dst)[4]) {
#pragma unroll
          for (int i = 0; i < 4; ++i) {
            const int tok = tile * kTN + 4 * lane + i;
            dst[i] = tok <= last_tok ? (__ldg(p.kv_indices + d.page0 + (tok >> p.log2_ps)) << p.log2_ps) + (tok & ps_mask) : last_slot;
          }
        };
        ldg_slots(0, sl);
        if (lw == 0 && lane == 0) {
          mbar_wait(bar_qfree, (qn & 1) ^ 1);
          tma::mbar_expect_tx(bar_q, 2 * kTileBytes2);
#pragma unroll
          for (int t = 0; t < 2; ++t)
#pragma unroll
            for (int h = 0; h < 2; ++h)
              tma::load_3d(smem_q + t * kTileBytes2 + h * kBlkBytes, &tm_q, bar_q, h * 64, d.kvh * G, d.q0 + d.j0 + t * tok_per_tile);
        }
        ++qn;
        const int col = d.kvh * kTD;
        for (int tile = 0; tile < g.J; ++tile, fill += 2) {
          const int slot = fill % kRing, use = fill / kRing;
          if (tile + 1 < g.J) ldg_slots(tile + 1, sn);      // in flight behind this tile's issue loop
          mbar_wait(bar_empty(slot), (use & 1) ^ 1);
          const uint32_t full = bar_full(slot);
          if (lane == 0) tma::mbar_expect_tx(full, kTileBytes2);
          __syncwarp();
          const uint32_t dst = smem_ring + slot * kTileBytes2 + lane * 512;
          tma::gather4(dst, tm, full, col, sl[0], sl[1], sl[2], sl[3]);
          tma::gather4(dst + kBlkBytes, tm, full, col + 64, sl[0], sl[1], sl[2], sl[3]);
#pragma unroll
          for (int i = 0; i < 4; ++i) sl[i] = sn[i];
        }
      }
    } else {
      const int lt = tid - 9 * 32;                    // 0..63
      const int gc = lt & 15, lrow0 = lt >> 4;        // 16-byte column gc of rows lrow0 + 4*i
      const uint32_t chunk_off = (uint32_t)(gc >> 3) * kBlkBytes;
      const int c8 = gc & 7;
      const int hrow = lane >> 4;
      const uint32_t stride_b = (uint32_t)(p.kv_slot_stride * 2);   // bytes between token slots (< 4 GiB)
      const uint32_t dst_even = chunk_off + lrow0 * 128 + ((c8 ^ lrow0) << 4);
      const uint32_t dst_odd = chunk_off + lrow0 * 128 + ((c8 ^ (lrow0 + 4)) << 4);
      // One cp.async group per fill.  A fill's mbarrier is signalled when the NEXT fill has been issued
      // (wait_group 1), so one fill is always in flight; before blocking on anything that may depend on the
      // consumers, the pending fill is flushed.
      uint32_t pend_bar = 0;
      auto rotate = [&](uint32_t this_bar) {
        cp_async_commit();
        cp_async_wait<1>();
        fence_proxy_async_smem();
        if (pend_bar) mbar_arrive(pend_bar);
        pend_bar = this_bar;
      };
      auto flush = [&]() {
        cp_async_wait<0>();
        fence_proxy_async_smem();
        if (pend_bar) mbar_arrive(pend_bar);
        pend_bar = 0;
      };
      auto wait_or_flush = [&](uint32_t bar, uint32_t parity) {
        if (!mbar_test_wait(bar, parity)) {
          flush();
          mbar_wait(bar, parity);
        }
      };
      int kc = 0, vc = 0;                              // K / V tiles filled so far (whole kernel)
      for (int it = 0;; ++it) {
        wait_or_flush(bar_ifull(it % kItemRing), (it / kItemRing) & 1);
        const ItemDesc d = take_item(it);
        if (!d.valid) break;
        const ItemGeo g = item_geometry(d, tok_per_tile, p.causal);
        if (g.J == 0) continue;
        // slot ids of a tile: this thread resolves tokens 4*lane + 2*lw (+1) and parks them in shared memory, where
        // the 32 rows it copies (lrow0 + 4i) find theirs; two tiles are resident (K runs one tile ahead of V)
        int pg_a = -1, pg_b = -1;
        auto ldg_pages = [&](int tile) {
          const int tok = tile * kTN + 4 * lane + 2 * lw;
          pg_a = pg_b = -1;
          if (tok < g.kv_hi_all) pg_a = __ldg(p.kv_indices + d.page0 + (tok >> p.log2_ps));
          if (tok + 1 < g.kv_hi_all) pg_b = __ldg(p.kv_indices + d.page0 + ((tok + 1) >> p.log2_ps));
        };
        auto put_slots = [&](int tile) {
          const int tok = tile * kTN + 4 * lane + 2 * lw;
          s_slots[tile & 1][lw][0][lane] = pg_a >= 0 ? (pg_a << p.log2_ps) + (tok & ps_mask) : -1;
          s_slots[tile & 1][lw][1][lane] = pg_b >= 0 ? (pg_b << p.log2_ps) + ((tok + 1) & ps_mask) : -1;
        };
        ldg_pages(0);
        wait_or_flush(bar_qfree, (qn & 1) ^ 1);
        if (lt == 0) {
          tma::mbar_expect_tx(bar_q, 2 * kTileBytes2);
#pragma unroll
          for (int t = 0; t < 2; ++t)
#pragma unroll
            for (int h = 0; h < 2; ++h)
              tma::load_3d(smem_q + t * kTileBytes2 + h * kBlkBytes, &tm_q, bar_q, h * 64, d.kvh * G, d.q0 + d.j0 + t * tok_per_tile);
        }
        ++qn;
        __syncwarp();                                    // every lane is done reading the previous item's slot ids
        put_slots(0);
        if (g.J > 1) {
          ldg_pages(1);
          put_slots(1);
        }
        const char* kbase = reinterpret_cast<const char*>(p.k_cache + (int64_t)d.kvh * kTD + gc * 8);
        const char* vbase = reinterpret_cast<const char*>(p.v_cache + (int64_t)d.kvh * kTD + gc * 8);
        auto fill_tile = [&](const char* base, int tile, int slot, int use) {
          __syncwarp();
          int sl[32];
          const int4* src = reinterpret_cast<const int4*>(&s_slots[tile & 1][lw][hrow][0]);
#pragma unroll
          for (int i = 0; i < 8; ++i) {
            const int4 v = src[i];
            sl[4 * i] = v.x; sl[4 * i + 1] = v.y; sl[4 * i + 2] = v.z; sl[4 * i + 3] = v.w;
          }
          wait_or_flush(bar_empty(slot), (use & 1) ^ 1);
          const uint32_t dst0 = smem_ring + slot * kTileBytes2;
          if ((tile + 1) * kTN <= g.kv_hi_all) {            // full tile: no zero-fill predicates
#pragma unroll
            for (int i = 0; i < 32; ++i)
              cp_async_16(dst0 + ((i & 1) ? dst_odd : dst_even) + i * 512, base + (uint64_t)(uint32_t)sl[i] * stride_b, 16u);
          } else {
#pragma unroll
            for (int i = 0; i < 32; ++i) {
              const bool ok = sl[i] >= 0;
              cp_async_16(dst0 + ((i & 1) ? dst_odd : dst_even) + i * 512, base + (uint64_t)(uint32_t)(ok ? sl[i] : 0) * stride_b,
                          ok ? 16u : 0u);
            }
          }
          rotate(bar_full(slot));
        };
        fill_tile(kbase, 0, 2 * (kc & 1), kc >> 1);
        ++kc;
        for (int j = 0; j < g.J; ++j) {
          if (j + 2 < g.J) ldg_pages(j + 2);               // consumed after V(j) has been issued
          if (j + 1 < g.J) {
            fill_tile(kbase, j + 1, 2 * (kc & 1), kc >> 1);
            ++kc;
          }
          fill_tile(vbase, j, 2 * (vc & 1) + 1, vc >> 1);
          ++vc;
          if (j + 2 < g.J) {
            __syncwarp();                                  // V(j) has read tile j's slot ids: its parity is free
            put_slots(j + 2);
          }
        }
      }
      flush();
    }
  } else if (warp == 8) {
    // =================================== MMA issuer =======================================================
    asm volatile("setmaxnreg.dec.sync.aligned.u32 96;\n");
    constexpr uint32_t idesc_qk0 = make_idesc_bf16_f32(kTM, 0, 0, 0);             // N field filled per tile
    constexpr uint32_t idesc_pv = make_idesc_bf16_f32(kTM, kTD, 0, 1);
    const uint64_t qdesc0 = make_smem_desc(smem + kWsOffQ, 16, 1024);
    const uint64_t kdesc0 = make_smem_desc(smem + kWsOffRing, 16, 1024);          // K tiles: K-major
    const uint64_t vdesc0 = make_smem_desc(smem + kWsOffRing, kBlkBytes, 1024);   // V tiles: MN-major
    // f = running ring-fill index of the K tile; nch = 32-column chunks of S to produce (N = 32 * nch: the
    // first N token rows of the K-major tile)
    auto issue_qk = [&](int t, int f, int nch) {
      const uint64_t qd = qdesc0 + (uint64_t)(t * (kTileBytes2 >> 4));
      const uint64_t kd = kdesc0 + (uint64_t)((f % kRing) * (kTileBytes2 >> 4));
      const uint32_t idesc_qk = idesc_qk0 | ((uint32_t)(nch << 2) << 17);
#pragma unroll
      for (int ks = 0; ks < kTD / 16; ++ks) {
        const uint32_t koff = ((uint32_t)(ks >> 2) * kBlkBytes + (uint32_t)(ks & 3) * 32) >> 4;
        umma_ss_elect(tmem + t * kTN, qd + koff, kd + koff, idesc_qk, ks > 0);
      }
      umma_commit_elect(bar_s(t));
    };
    // f = running ring-fill index of the V tile; nks = 16-token K steps holding valid P columns (2 per chunk)
    auto issue_pv = [&](int t, int f, int half, bool acc, int nks) {
      const uint64_t vd = vdesc0 + (uint64_t)((f % kRing) * (kTileBytes2 >> 4));
#pragma unroll
      for (int ks = half * 4; ks < half * 4 + 4; ++ks)
        if (ks < nks)
          umma_ts_elect(tmem + 2 * kTN + t * kTD, tmem + t * kTN + ks * 8, vd + (uint64_t)(ks * (2048 >> 4)), idesc_pv,
                        (acc || ks > 0));
    };
    int fillc = 0, qn = 0;
    int pc[2] = {0, 0};
    for (int it = 0;; ++it) {
      const ItemDesc d = take_item(it);
      if (!d.valid) break;
      const ItemGeo g = item_geometry(d, tok_per_tile, p.causal);
      if (g.J == 0) continue;
      const int base = fillc;
      mbar_wait(bar_q, qn & 1);
      mbar_wait(bar_full(base % kRing), (base / kRing) & 1);
      tc_fence_after();
      if (g.nt0 > 0) issue_qk(0, base, chunks_of(g.kvhi0, 0));
      if (g.nt1 > 0) issue_qk(1, base, chunks_of(g.kvhi1, 0));
      umma_commit_elect(bar_empty(base % kRing));                        // K(0) free once both QK(0) are done
      if (g.J == 1) umma_commit_elect(bar_qfree);                        // last QK of the item: Q may be refilled
      for (int j = 0; j < g.J; ++j) {
        const int fv = base + 2 * j + 1, fk = base + 2 * j + 2;
        mbar_wait(bar_full(fv % kRing), (fv / kRing) & 1);               // V(j)
        bool k_ready = false;
#pragma unroll
        for (int t = 0; t < 2; ++t) {
          const int ntt = t ? g.nt1 : g.nt0;
          const int kvhit = t ? g.kvhi1 : g.kvhi0;
          if (j < ntt) {
            const int nks = 2 * chunks_of(kvhit, j);
            mbar_wait(bar_p(t, 0), pc[t] & 1);                           // first 64 columns of P_t(j) are in TMEM
            tc_fence_after();
            issue_pv(t, fv, 0, j > 0, nks);
            mbar_wait(bar_p(t, 1), pc[t] & 1);                           // the rest
            tc_fence_after();
            issue_pv(t, fv, 1, true, nks);
            ++pc[t];
            if (j == ntt - 1) umma_commit_elect(bar_o(t));
          }
          if (j + 1 < ntt) {
            if (!k_ready) {
              mbar_wait(bar_full(fk % kRing), (fk / kRing) & 1);         // K(j+1)
              tc_fence_after();
              k_ready = true;
            }
            issue_qk(t, fk, chunks_of(kvhit, j + 1));
          }
        }
        umma_commit_elect(bar_empty(fv % kRing));
        if (j + 1 < g.J) umma_commit_elect(bar_empty(fk % kRing));
        if (j + 2 == g.J) umma_commit_elect(bar_qfree);
      }
      fillc += 2 * g.J;
      ++qn;
    }
  } else {
    // =================================== softmax groups ===================================================
    asm volatile("setmaxnreg.inc.sync.aligned.u32 200;\n");
    const int t = warp >> 2;                          // query tile 0/1
    const int r = tid & 127;                          // row in the tile == TMEM lane
    const uint32_t lane_base = (uint32_t)((warp & 3) * 32) << 16;
    const uint32_t tS = tmem + lane_base + t * kTN;
    const uint32_t tO = tmem + lane_base + 2 * kTN + t * kTD;
    int sc = 0, oc = 0;
    for (int it = 0;; ++it) {
      const ItemDesc d = take_item(it);
      if (!d.valid) break;
      const ItemGeo g = item_geometry(d, tok_per_tile, p.causal);
      const int my_tok = d.j0 + t * tok_per_tile + (r >> log2_g);
      const int lim = p.causal ? (d.Lkv - d.Lq + my_tok) : (d.Lkv - 1);
      const int lim0 = p.causal ? (d.Lkv - d.Lq + d.j0 + t * tok_per_tile) : (d.Lkv - 1);
      const int kv_hi = t ? g.kvhi1 : g.kvhi0;
      const int ntt = t ? g.nt1 : g.nt0;
      const int tokst = t ? g.toks1 : g.toks0;
      float m_used = 0.f, l_run = 0.f;
      for (int j = 0; j < ntt; ++j, ++sc) {
        mbar_wait(bar_s(t), sc & 1);
        tc_fence_after();
        const int kv0 = j * kTN;
        const bool need_mask = (kv0 + kTN - 1 > lim0) || (kv0 + kTN > kv_hi);
        const int nvalid = need_mask ? min(lim, kv_hi - 1) - kv0 + 1 : kTN;   // valid columns of my row in this tile
        if constexpr (kUnifiedMask) {
          softmax_tile<false>(tS, tO, nvalid, j > 0, p.scale_log2, m_used, l_run, bar_p(t, 0), bar_p(t, 1), need_mask);
        } else if (need_mask) {
          softmax_tile<true, 4>(tS, tO, nvalid, j > 0, p.scale_log2, m_used, l_run, bar_p(t, 0), bar_p(t, 1));
        } else {
          softmax_tile<false>(tS, tO, kTN, j > 0, p.scale_log2, m_used, l_run, bar_p(t, 0), bar_p(t, 1));
        }
      }
      // ---- epilogue -------------------------------------------------------------------------------------------
      const bool has = ntt > 0;
      if (has) {
        mbar_wait(bar_o(t), oc & 1);
        ++oc;
        tc_fence_after();
      }
      const float inv_l = l_run > 0.f ? 1.f / l_run : 0.f;
      // Output rows go through shared memory so that every STG.128 of a warp writes four full 128-byte lines
      // (8 lanes per row); straight from the TMEM layout (thread = row) a warp store touches 32 lines with 16 bytes
      // each and the 8 softmax warps spent ~4000 LSU wavefront-cycles per item in the epilogue (14 % of their
      // time on bench-shaped prompts).  Staging area, 4 KiB per warp (32 rows x 64 columns bf16, 16-byte pieces
      // XOR-swizzled by row): the 7th operand tile, touched by nobody else.
      const uint32_t stage = smem + kWsOffStage + (uint32_t)warp * 4096u;
      const int rows_ok = tokst << log2_g;
      const int rsub = lane >> 3, piece = lane & 7;
      if (!has || !staged) {
        // thread = row, straight from the TMEM layout.  Rows without any visible KV (Lkv == 0, or causal with
        // Lkv < Lq) are zeros; with data this is the unstaged form kept for A/B measurements.
        if (r < rows_ok) {
          uint4* dst = reinterpret_cast<uint4*>(p.out + (int64_t)(d.q0 + my_tok) * p.out_row_stride +
                                                (int64_t)(d.kvh * G + (r & (G - 1))) * kTD);
          if (!has) {
#pragma unroll
            for (int e = 0; e < 16; ++e) dst[e] = make_uint4(0u, 0u, 0u, 0u);
          }
        }
        if (has) {
          uint4* dst = reinterpret_cast<uint4*>(p.out + (int64_t)(d.q0 + (r < rows_ok ? my_tok : d.j0)) * p.out_row_stride +
                                                (int64_t)(d.kvh * G + (r & (G - 1))) * kTD);
#pragma unroll
          for (int c = 0; c < 4; ++c) {
            uint32_t ov[32];
            tmem_ld_32x32b_x32(tO + c * 32, ov);
            tmem_wait_ld();
            if (r < rows_ok) {
#pragma unroll
              for (int j8 = 0; j8 < 4; ++j8) {
                uint4 w;
                w.x = pack_bf16x2(__uint_as_float(ov[j8 * 8 + 0]) * inv_l, __uint_as_float(ov[j8 * 8 + 1]) * inv_l);
                w.y = pack_bf16x2(__uint_as_float(ov[j8 * 8 + 2]) * inv_l, __uint_as_float(ov[j8 * 8 + 3]) * inv_l);
                w.z = pack_bf16x2(__uint_as_float(ov[j8 * 8 + 4]) * inv_l, __uint_as_float(ov[j8 * 8 + 5]) * inv_l);
                w.w = pack_bf16x2(__uint_as_float(ov[j8 * 8 + 6]) * inv_l, __uint_as_float(ov[j8 * 8 + 7]) * inv_l);
                dst[c * 4 + j8] = w;
              }
            }
          }
        }
      }
      if (kLeanEpilogue && has && staged) {
        // one 64-bit base per item (first token of this query tile, first head of the KV head's group, this lane's
        // 16-byte piece); per store a 32-bit element delta from the packed row index R = (token, head-in-group)
        __nv_bfloat16* const out0 = p.out + (int64_t)(d.q0 + d.j0 + t * tok_per_tile) * p.out_row_stride +
                                    (int64_t)(d.kvh * G) * kTD + piece * 8;
        const uint32_t stride32 = (uint32_t)p.out_row_stride;
        const int R0 = (warp & 3) * 32 + rsub;
        const uint64_t inv2 = pack2(__float_as_uint(inv_l), __float_as_uint(inv_l));
#pragma unroll
        for (int half = 0; half < 2; ++half) {
#pragma unroll
          for (int cc = 0; cc < 2; ++cc) {
            uint32_t ov[32];
            tmem_ld_32x32b_x32(tO + (half * 2 + cc) * 32, ov);
            tmem_wait_ld();
#pragma unroll
            for (int j8 = 0; j8 < 4; ++j8) {
              uint32_t w[4];
#pragma unroll
              for (int q = 0; q < 4; ++q) {
                uint32_t lo, hi;
                unpack2(mul2(pack2(ov[j8 * 8 + 2 * q], ov[j8 * 8 + 2 * q + 1]), inv2), lo, hi);
                w[q] = pack_bf16x2(__uint_as_float(lo), __uint_as_float(hi));
              }
              const uint32_t dst_s = stage + (uint32_t)lane * 128u + (uint32_t)(((cc * 4 + j8) ^ (lane & 7)) << 4);
              asm volatile("st.shared.v4.b32 [%0], {%1,%2,%3,%4};\n" ::"r"(dst_s), "r"(w[0]), "r"(w[1]), "r"(w[2]), "r"(w[3]) : "memory");
            }
          }
          __syncwarp();
#pragma unroll
          for (int k = 0; k < 8; ++k) {
            const int row = 4 * k + rsub;                  // row of this warp's 32
            uint4 w;
            const uint32_t src_s = stage + (uint32_t)row * 128u + (uint32_t)((piece ^ (row & 7)) << 4);
            asm volatile("ld.shared.v4.b32 {%0,%1,%2,%3}, [%4];\n" : "=r"(w.x), "=r"(w.y), "=r"(w.z), "=r"(w.w) : "r"(src_s) : "memory");
            const int R = R0 + 4 * k;                      // packed row of tile t
            if (R < rows_ok) {
              const uint32_t delta = (uint32_t)(R >> log2_g) * stride32 + (uint32_t)(R & (G - 1)) * kTD + half * 64;
              *reinterpret_cast<uint4*>(out0 + delta) = w;
            }
          }
          __syncwarp();                                    // the next half (or item) overwrites the staging rows
        }
      }
#pragma unroll
      for (int half = 0; half < 2 && has && staged && !kLeanEpilogue; ++half) {
#pragma unroll
        for (int cc = 0; cc < 2; ++cc) {
          uint32_t ov[32];
          tmem_ld_32x32b_x32(tO + (half * 2 + cc) * 32, ov);
          tmem_wait_ld();
#pragma unroll
          for (int j8 = 0; j8 < 4; ++j8) {
            const uint32_t w0 = pack_bf16x2(__uint_as_float(ov[j8 * 8 + 0]) * inv_l, __uint_as_float(ov[j8 * 8 + 1]) * inv_l);
            const uint32_t w1 = pack_bf16x2(__uint_as_float(ov[j8 * 8 + 2]) * inv_l, __uint_as_float(ov[j8 * 8 + 3]) * inv_l);
            const uint32_t w2 = pack_bf16x2(__uint_as_float(ov[j8 * 8 + 4]) * inv_l, __uint_as_float(ov[j8 * 8 + 5]) * inv_l);
            const uint32_t w3 = pack_bf16x2(__uint_as_float(ov[j8 * 8 + 6]) * inv_l, __uint_as_float(ov[j8 * 8 + 7]) * inv_l);
            const uint32_t dst_s = stage + (uint32_t)lane * 128u + (uint32_t)(((cc * 4 + j8) ^ (lane & 7)) << 4);
            asm volatile("st.shared.v4.b32 [%0], {%1,%2,%3,%4};\n" ::"r"(dst_s), "r"(w0), "r"(w1), "r"(w2), "r"(w3) : "memory");
          }
        }
        __syncwarp();
#pragma unroll
        for (int k = 0; k < 8; ++k) {
          const int row = 4 * k + rsub;                  // row of this warp's 32
          uint4 w;
          const uint32_t src_s = stage + (uint32_t)row * 128u + (uint32_t)((piece ^ (row & 7)) << 4);
          asm volatile("ld.shared.v4.b32 {%0,%1,%2,%3}, [%4];\n" : "=r"(w.x), "=r"(w.y), "=r"(w.z), "=r"(w.w) : "r"(src_s) : "memory");
          const int R = (warp & 3) * 32 + row;           // packed row of tile t
          if (R < rows_ok) {
            const int tok = d.j0 + t * tok_per_tile + (R >> log2_g);
            __nv_bfloat16* dst = p.out + (int64_t)(d.q0 + tok) * p.out_row_stride +
                                 (int64_t)(d.kvh * G + (R & (G - 1))) * kTD + half * 64 + piece * 8;
            *reinterpret_cast<uint4*>(dst) = w;
          }
        }
        __syncwarp();                                    // the next half (or item) overwrites the staging rows
      }
      tc_fence_before();   // O_t reads are complete before this group's next P arrive lets PV_t(0) overwrite O_t
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 0) tmem_dealloc(tmem, 512);
}

size_t prefill_sm100_workspace_bytes(int B, int total_q, int group) {
  int log2_g = 0;
  while ((1 << log2_g) < group) ++log2_g;
  const int tok_per_cta = 2 * (kTM >> log2_g);
  const size_t max_pairs = (size_t)total_q / tok_per_cta + (size_t)B;
  return kWorkHeaderInts * sizeof(int) + max_pairs * sizeof(int2);
}

int launch_prefill_sm100(const PrefillParams& p, int total_q, void* workspace, size_t workspace_bytes, cudaStream_t st) {
  int log2_g = 0;
  while ((1 << log2_g) < p.group) ++log2_g;
  const int tok_per_tile = kTM >> log2_g;
  // device-built work list + persistent kernel, one CTA per SM
  const size_t need = prefill_sm100_workspace_bytes(p.B, total_q, p.group);
  if (!workspace || workspace_bytes < need) return fail("b200_prefill: workspace %zu < %zu bytes", workspace_bytes, need);
  if ((uintptr_t)workspace & 15) return fail("b200_prefill: workspace must be 16-byte aligned");
  if (p.out_row_stride < 0 || p.out_row_stride > (1 << 22)) return fail("b200_prefill: out row stride %lld not in [0, 2^22]", (long long)p.out_row_stride);
  static std::atomic<bool> attr_done[64];
  if (cudaError_t e = ensure_dyn_smem(prefill_sm100_persist_kernel, kPersistSmem, attr_done); e != cudaSuccess)
    return fail("b200_prefill(sm100 persist): cudaFuncSetAttribute: %s", cudaGetErrorString(e));
  int* work = reinterpret_cast<int*>(workspace);
  int2* items = reinterpret_cast<int2*>(work + kWorkHeaderInts);
  const int max_pairs = (int)((workspace_bytes - kWorkHeaderInts * sizeof(int)) / sizeof(int2));
  prefill_plan_kernel<<<1, kPlanThreads, 0, st>>>(p, 2 * tok_per_tile, work, items, max_pairs);
  if (int rc = check_launch("b200_prefill(plan)")) return rc;
  // Tensor maps (host-encoded per call: q and the layer's pool pointers change between calls; the maps travel as
  // __grid_constant__ kernel parameters, so the launch stays CUDA-graph capturable).
  //   Q  : [T tokens][Hq heads][128 d], box [tok_per_tile][G][64]  -> one box = one 64-column half of a 128-row packed tile
  //   K/V: [slots][Hkv * 128], box [1][64], issued as tile::gather4 (four token slots per instruction).  The pool's
  //        slot count is not part of the C ABI; the extent is an upper bound (rows are only touched through valid slot
  //        ids, and out-of-range tile rows re-read a valid slot, so no out-of-bounds fill is relied on).
  CUtensorMap tm_q, tm_k, tm_v;
  {
    const int G = 1 << log2_g;
    const uint64_t qd[3] = {(uint64_t)kTD, (uint64_t)p.Hq, (uint64_t)total_q};
    const uint64_t qs[2] = {(uint64_t)kTD * 2, (uint64_t)p.q_row_stride * 2};
    const uint32_t qb[3] = {64u, (uint32_t)G, (uint32_t)tok_per_tile};
    const uint64_t kd[2] = {(uint64_t)p.Hkv * kTD, (uint64_t)1 << 30};
    const uint64_t ks[1] = {(uint64_t)p.kv_slot_stride * 2};
    const uint32_t kb[2] = {64u, 1u};
    if (!tma::encode_bf16(&tm_q, p.q, 3, qd, qs, qb) || !tma::encode_bf16(&tm_k, p.k_cache, 2, kd, ks, kb) ||
        !tma::encode_bf16(&tm_v, p.v_cache, 2, kd, ks, kb))
      return fail("b200_prefill: cuTensorMapEncodeTiled failed (q stride %lld, kv slot stride %lld)", (long long)p.q_row_stride,
                  (long long)p.kv_slot_stride);
  }
  cudaError_t e = launch_pdl(prefill_sm100_persist_kernel, dim3(cached_num_sms()), dim3(kWsThreads), kPersistSmem, st, p, log2_g, work,
                             (const int2*)items, tm_q, tm_k, tm_v);
  if (e != cudaSuccess) return fail("b200_prefill(sm100 persist): %s", cudaGetErrorString(e));
  return check_launch("b200_prefill(sm100 persist)");
}

bool prefill_sm100_supports(int group) { return group >= 1 && group <= 64 && (group & (group - 1)) == 0; }

}  // namespace b200
