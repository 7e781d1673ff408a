/*
 * Plain-C restatement of the attention / KV-cache hot path of nano-vllm-x.
 *
 * TEST INFRASTRUCTURE ONLY: linked/loaded by tests/ and by bench.py's cpu_baseline leg, never by
 * the product package.  Mirrors oracle/paged_attention_oracle.py function by function (that file
 * carries the pinning notes); citations are into /root/reference/nanovllmx/.
 *
 * Build: make -C oracle   (gcc -O3 -pthread -shared -> oracle/_build/liboracle.so)
 */
#include <math.h>
#include <pthread.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>

static inline float bf16_to_f32(uint16_t h) {
  uint32_t u = (uint32_t)h << 16;
  float f;
  memcpy(&f, &u, 4);
  return f;
}
static inline uint16_t f32_to_bf16_rn(float f) {
  uint32_t u;
  memcpy(&u, &f, 4);
  if ((u & 0x7fffffffu) > 0x7f800000u) return (uint16_t)((u >> 16) | 0x40); /* NaN */
  u += 0x7fffu + ((u >> 16) & 1u);
  return (uint16_t)(u >> 16);
}

/* a-1: core/kv_cache.py:12-44.  Rows are raw 2-byte elements; slot -1 is skipped (:25-26). */
void oracle_store_kvcache(const uint16_t* k, int64_t k_row_stride, const uint16_t* v, int64_t v_row_stride,
                          uint16_t* k_cache, uint16_t* v_cache, const int64_t* slots, int T, int row_elems) {
  for (int t = 0; t < T; ++t) {
    const int64_t s = slots[t];
    if (s == -1) continue;
    memcpy(k_cache + s * row_elems, k + t * k_row_stride, (size_t)row_elems * 2);
    memcpy(v_cache + s * row_elems, v + t * v_row_stride, (size_t)row_elems * 2);
  }
}

/* a-3: layers/attention.py:62-101, 134-141.  Flat inputs: kv_flat = concat of every sequence's
 * kv_indices.  Outputs must be preallocated: qo_indptr/kv_indptr [B+1], last_page_len [B],
 * kv_indices [sum pages], out_slots [sum new].  Returns 0, or -1 if page_size>1 slots are not
 * page aligned (semantics defined by this build, see the .py header). */
int oracle_build_metadata(const int64_t* num_token_ids, const int64_t* cached_kv_len, const int64_t* kv_lens,
                          const int64_t* kv_flat, int B, int page_size, int32_t* qo_indptr, int32_t* kv_indptr,
                          int32_t* kv_indices, int32_t* last_page_len, int64_t* out_slots) {
  int64_t q = 0, pg = 0, nn = 0, off = 0;
  qo_indptr[0] = 0;
  kv_indptr[0] = 0;
  for (int i = 0; i < B; ++i) {
    const int64_t L = kv_lens[i];
    q += num_token_ids[i] - cached_kv_len[i];
    qo_indptr[i + 1] = (int32_t)q;
    const int64_t pages = L / page_size + (L % page_size != 0);
    for (int64_t p = 0; p < pages; ++p) {
      const int64_t first = kv_flat[off + p * page_size];
      if (page_size > 1 && first % page_size) return -1;
      kv_indices[pg + p] = (int32_t)(first / page_size);
    }
    pg += pages;
    kv_indptr[i + 1] = (int32_t)pg;
    const int64_t r = L % page_size;
    last_page_len[i] = (int32_t)(r == 0 ? page_size : r); /* L == 0 -> page_size, :88-90 */
    for (int64_t j = cached_kv_len[i]; j < L; ++j) out_slots[nn++] = kv_flat[off + j];
    off += L;
  }
  return 0;
}

/* a-4: layers/attention.py:219-275 — refill of the named static buffers for one replayed step. */
void oracle_replay_graph_buffers(const int64_t* cached_kv_len, const int64_t* kv_lens, const int64_t* kv_flat, int bs,
                                 int padded_bs, int32_t* indptr_buf, int32_t* indices_buf, int64_t indices_cap,
                                 int64_t* out_buf, int64_t out_cap) {
  int64_t acc = 0, off = 0;
  indptr_buf[0] = 0;
  for (int i = 0; i < padded_bs; ++i) {
    acc += i < bs ? kv_lens[i] : 1; /* padding sequences: kv_len 1 (:239-244) */
    indptr_buf[i + 1] = (int32_t)acc;
  }
  memset(indices_buf, 0, (size_t)indices_cap * sizeof(int32_t)); /* fill_(0) (:258) */
  for (int i = 0; i < bs; ++i)
    for (int64_t j = 0; j < kv_lens[i]; ++j, ++off) indices_buf[off] = (int32_t)kv_flat[off];
  for (int64_t j = 0; j < out_cap; ++j) out_buf[j] = -1; /* fill_(-1) (:265) */
  off = 0;
  int64_t n = 0;
  for (int i = 0; i < bs; ++i) {
    for (int64_t j = cached_kv_len[i]; j < kv_lens[i]; ++j) out_buf[n++] = kv_flat[off + j];
    off += kv_lens[i];
  }
}

/* a-6 / a-7: layers/attention.py:332-373 semantics in fp32: gather, GQA head map h -> h/(Hq/Hkv),
 * softmax(scale * q.k) v, bottom-right causal (kv_pos <= Lkv - Lq + j).  bf16 in, fp32 out.
 * Work items (sequence, KV head) are handed to `threads` pthreads through an atomic counter (this image's
 * gcc ships without libgomp) — all host threads, as the reported CPU baseline. */
typedef struct {
  const uint16_t *q, *k_cache, *v_cache;
  const int32_t *qo_indptr, *kv_indptr, *kv_indices, *last_page_len;
  int B, Hq, Hkv, D, page_size, causal;
  float sm_scale;
  float* out;
  int next; /* atomic work counter */
} attn_job;

static void* attn_worker(void* arg) {
  attn_job* J = (attn_job*)arg;
  const int group = J->Hq / J->Hkv, D = J->D, ps = J->page_size;
  float* s = NULL; /* [group][lkv] scores, then weights */
  float* qf = (float*)malloc((size_t)group * D * sizeof(float));
  float* rowf = (float*)malloc((size_t)D * sizeof(float));
  float* mx = (float*)malloc((size_t)group * sizeof(float));
  size_t cap = 0;
  for (;;) {
    /* one item = (sequence, KV head): the `group` query heads that share it walk its K/V rows together, so every
     * row is fetched and widened to fp32 once per group (GQA as the GPU kernels exploit it); the arithmetic of
     * each head is unchanged: 8 partial sums in a fixed order for q.k, tokens in order for softmax and P.V */
    const int item = __atomic_fetch_add(&J->next, 1, __ATOMIC_RELAXED);
    if (item >= J->B * J->Hkv) break;
    const int i = item / J->Hkv, kvh = item % J->Hkv;
    const int q0 = J->qo_indptr[i], lq = J->qo_indptr[i + 1] - q0;
    const int p0 = J->kv_indptr[i], np = J->kv_indptr[i + 1] - p0;
    const int lkv = np > 0 ? (np - 1) * ps + J->last_page_len[i] : 0;
    if ((size_t)lkv * group > cap) {
      free(s);
      cap = (size_t)lkv * group;
      s = (float*)malloc(cap * sizeof(float));
    }
    for (int j = 0; j < lq; ++j) {
      float* o = J->out + ((int64_t)(q0 + j) * J->Hq + (int64_t)kvh * group) * D; /* group heads are adjacent */
      for (int d = 0; d < group * D; ++d) o[d] = 0.f;
      int vis = J->causal ? lkv - lq + j + 1 : lkv;
      if (vis > lkv) vis = lkv;
      if (vis <= 0) continue;
      const uint16_t* qr = J->q + ((int64_t)(q0 + j) * J->Hq + (int64_t)kvh * group) * D;
      for (int d = 0; d < group * D; ++d) qf[d] = bf16_to_f32(qr[d]);
      for (int g = 0; g < group; ++g) mx[g] = -INFINITY;
      for (int t = 0; t < vis; ++t) {
        const int64_t slot = (int64_t)J->kv_indices[p0 + t / ps] * ps + t % ps;
        const uint16_t* kr = J->k_cache + (slot * J->Hkv + kvh) * D;
        for (int d = 0; d < D; ++d) rowf[d] = bf16_to_f32(kr[d]);
        for (int g = 0; g < group; ++g) {
          const float* qg = qf + (size_t)g * D;
          float a8[8] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
          int d = 0;
          for (; d + 8 <= D; d += 8)
            for (int u = 0; u < 8; ++u) a8[u] += qg[d + u] * rowf[d + u];
          float acc = ((a8[0] + a8[4]) + (a8[1] + a8[5])) + ((a8[2] + a8[6]) + (a8[3] + a8[7]));
          for (; d < D; ++d) acc += qg[d] * rowf[d];
          const float sc = acc * J->sm_scale;
          s[(size_t)g * lkv + t] = sc;
          if (sc > mx[g]) mx[g] = sc;
        }
      }
      for (int g = 0; g < group; ++g) {
        float* sg = s + (size_t)g * lkv;
        float l = 0.f;
        for (int t = 0; t < vis; ++t) {
          sg[t] = expf(sg[t] - mx[g]);
          l += sg[t];
        }
        for (int t = 0; t < vis; ++t) sg[t] = sg[t] / l;
      }
      for (int t = 0; t < vis; ++t) {
        const int64_t slot = (int64_t)J->kv_indices[p0 + t / ps] * ps + t % ps;
        const uint16_t* vr = J->v_cache + (slot * J->Hkv + kvh) * D;
        for (int d = 0; d < D; ++d) rowf[d] = bf16_to_f32(vr[d]);
        for (int g = 0; g < group; ++g) {
          const float w = s[(size_t)g * lkv + t];
          float* og = o + (size_t)g * D;
          for (int d = 0; d < D; ++d) og[d] += w * rowf[d];
        }
      }
    }
  }
  free(s);
  free(qf);
  free(rowf);
  free(mx);
  return NULL;
}

void oracle_paged_attention(const uint16_t* q, const uint16_t* k_cache, const uint16_t* v_cache, const int32_t* qo_indptr,
                            const int32_t* kv_indptr, const int32_t* kv_indices, const int32_t* last_page_len, int B,
                            int Hq, int Hkv, int D, int page_size, float sm_scale, int causal, float* out, int threads) {
  attn_job J = {q, k_cache, v_cache, qo_indptr, kv_indptr, kv_indices, last_page_len, B, Hq, Hkv, D, page_size, causal,
                sm_scale, out, 0};
  if (threads < 1) threads = 1;
  if (threads > 256) threads = 256;
  pthread_t th[256];
  for (int t = 1; t < threads; ++t) pthread_create(&th[t], NULL, attn_worker, &J);
  attn_worker(&J);
  for (int t = 1; t < threads; ++t) pthread_join(th[t], NULL);
}

uint16_t oracle_f32_to_bf16(float f) { return f32_to_bf16_rn(f); }
