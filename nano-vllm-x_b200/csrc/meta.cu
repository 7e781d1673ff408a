// (2) Device-side attention metadata: slot-table maintenance + indptr / page table /
// last-page-len / slot-map construction.  Replaces layers/attention.py:62-101, 134-141 (eager)
// and :219-275 (CUDA-graph replay); see include/b200attn.h for the message layout.
//
// Integer work, bit-exact against the reference arrays.  Three tiny HBM-bound kernels:
//   slot_table_update : scatter this step's new slot ids into their table rows
//   meta_scan         : one CTA, exclusive scans over B sequences (q lens, pages, new slots)
//   meta_gather       : grid over (sequence, chunk): coalesced row copy table -> page table,
//                       plus the tail fills the reference does with fill_(0) / fill_(-1)
#include "common.cuh"

namespace b200 {

__global__ void __launch_bounds__(256) slot_table_update_kernel(int32_t* __restrict__ table, int64_t row_stride,
                                                                const int32_t* __restrict__ msg, int B) {
  // grid.x = sequence, grid.y = chunk of that sequence's new ids
  const int i = blockIdx.x;
  const int row = msg[i];
  const int col = msg[4 * B + i];
  const int cnt = msg[5 * B + i];
  const int off = msg[6 * B + i];
  const int32_t* payload = msg + 7 * (int64_t)B + off;
  int32_t* dst = table + (int64_t)row * row_stride + col;
  for (int j = blockIdx.y * blockDim.x + threadIdx.x; j < cnt; j += gridDim.y * blockDim.x) dst[j] = payload[j];
}

// scratch[0 .. padded_bs] receives the exclusive scan of "new slots per sequence".
__global__ void __launch_bounds__(1024) meta_scan_kernel(const int32_t* __restrict__ msg, int B, int padded_bs,
                                                         int log2_ps, int32_t* __restrict__ qo_indptr,
                                                         int32_t* __restrict__ kv_indptr,
                                                         int32_t* __restrict__ last_page_len,
                                                         int32_t* __restrict__ new_indptr) {
  __shared__ int scratch[3][33];
  const int ps = 1 << log2_ps;
  int carry_q = 0, carry_p = 0, carry_n = 0;
  if (threadIdx.x == 0) {
    if (qo_indptr) qo_indptr[0] = 0;
    kv_indptr[0] = 0;
    new_indptr[0] = 0;
  }
  for (int base = 0; base < padded_bs; base += 1024) {
    const int i = base + threadIdx.x;
    int q = 0, pages = 0, nnew = 0, kv = 0;
    if (i < B) {
      kv = msg[B + i];
      const int cached = msg[2 * B + i];
      q = msg[3 * B + i];
      nnew = kv - cached;
      if (nnew < 0) nnew = 0;
    } else if (i < padded_bs) {
      kv = 1;  // CUDA-graph padding sequence (layers/attention.py:239-244)
    }
    pages = (kv >> log2_ps) + ((kv & (ps - 1)) != 0);
    int tq, tp, tn;
    const int iq = block_inclusive_scan<1024>(q, scratch[0], &tq);
    const int ip = block_inclusive_scan<1024>(pages, scratch[1], &tp);
    const int in = block_inclusive_scan<1024>(nnew, scratch[2], &tn);
    if (i < padded_bs) {
      kv_indptr[i + 1] = carry_p + ip;
      new_indptr[i + 1] = carry_n + in;
      if (last_page_len) {
        const int r = kv & (ps - 1);
        last_page_len[i] = r == 0 ? ps : r;  // kv == 0 -> page_size, as layers/attention.py:88-90
      }
      if (qo_indptr && i < B) qo_indptr[i + 1] = carry_q + iq;
    }
    carry_q += tq;
    carry_p += tp;
    carry_n += tn;
  }
}

// grid.x = padded_bs + 1.  Blocks 0..padded_bs-1 copy one sequence each (grid.y chunks);
// block padded_bs does the tail fills.
__global__ void __launch_bounds__(256) meta_gather_kernel(const int32_t* __restrict__ table, int64_t row_stride,
                                                          const int32_t* __restrict__ msg, int B, int padded_bs,
                                                          int log2_ps, int fill_tail,
                                                          const int32_t* __restrict__ kv_indptr,
                                                          const int32_t* __restrict__ new_indptr,
                                                          int32_t* __restrict__ kv_indices, int64_t kv_cap,
                                                          int64_t* __restrict__ out_slots, int64_t out_cap) {
  const int i = blockIdx.x;
  const int tid = blockIdx.y * blockDim.x + threadIdx.x;
  const int nth = gridDim.y * blockDim.x;
  if (i < padded_bs) {
    const int p0 = kv_indptr[i];
    const int np = kv_indptr[i + 1] - p0;
    if (i < B) {
      const int32_t* row = table + (int64_t)msg[i] * row_stride;
      for (int p = tid; p < np; p += nth) {
        if ((int64_t)p0 + p < kv_cap) kv_indices[p0 + p] = row[(int64_t)p << log2_ps] >> log2_ps;
      }
      const int cached = msg[2 * B + i];
      const int n0 = new_indptr[i];
      const int nn = new_indptr[i + 1] - n0;
      for (int j = tid; j < nn; j += nth) {
        if ((int64_t)n0 + j < out_cap) out_slots[n0 + j] = (int64_t)row[cached + j];
      }
    } else {
      // padding sequence: one page, slot 0 (layers/attention.py:258)
      for (int p = tid; p < np; p += nth)
        if ((int64_t)p0 + p < kv_cap) kv_indices[p0 + p] = 0;
    }
  } else if (fill_tail) {
    // reference: whole-buffer fill_(0) / fill_(-1) before the copies (layers/attention.py:258,265)
    const int64_t total_p = kv_indptr[padded_bs];
    const int64_t total_n = new_indptr[padded_bs];
    const int64_t z = (int64_t)(blockIdx.x - padded_bs);  // tail blocks are grid.x - padded_bs wide
    const int64_t ntail = (int64_t)(gridDim.x - padded_bs) * nth;
    for (int64_t j = total_p + z * nth + tid; j < kv_cap; j += ntail) kv_indices[j] = 0;
    for (int64_t j = total_n + z * nth + tid; j < out_cap; j += ntail) out_slots[j] = -1;
  }
}

static int ilog2_exact(int x) {
  if (x <= 0 || (x & (x - 1))) return -1;
  int l = 0;
  while ((1 << l) < x) ++l;
  return l;
}

}  // namespace b200

extern "C" int b200_slot_table_update(int32_t* slot_table, int64_t table_row_stride, const int32_t* step_msg,
                                      int32_t B, int32_t n_payload, b200_stream_t stream) {
  using namespace b200;
  if (B < 0 || n_payload < 0) return fail("b200_slot_table_update: negative size");
  if (B == 0 || n_payload == 0) return 0;
  if (!slot_table || !step_msg) return fail("b200_slot_table_update: null pointer");
  // chunks per sequence: enough for a full prompt at 256 ids per pass without oversizing decode
  int chunks = (n_payload / B + 1023) / 1024;
  if (chunks < 1) chunks = 1;
  if (chunks > 64) chunks = 64;
  dim3 grid((unsigned)B, (unsigned)chunks);
  slot_table_update_kernel<<<grid, 256, 0, (cudaStream_t)stream>>>(slot_table, table_row_stride, step_msg, B);
  return check_launch("b200_slot_table_update");
}

extern "C" int b200_build_meta(const int32_t* slot_table, int64_t table_row_stride, const int32_t* step_msg,
                               int32_t B, int32_t padded_bs, int32_t page_size, int32_t fill_tail,
                               int32_t* qo_indptr, int32_t* kv_indptr, int32_t* kv_indices, int64_t kv_indices_cap,
                               int32_t* last_page_len, int64_t* out_slots, int64_t out_slots_cap, int32_t* scratch,
                               b200_stream_t stream) {
  using namespace b200;
  const int log2_ps = ilog2_exact(page_size);
  if (log2_ps < 0) return fail("b200_build_meta: page_size %d is not a power of two", page_size);
  if (B < 0 || padded_bs < B) return fail("b200_build_meta: need 0 <= B <= padded_bs (B=%d padded=%d)", B, padded_bs);
  if (padded_bs == 0) return 0;
  if (!slot_table || !step_msg || !kv_indptr || !kv_indices || !out_slots || !scratch)
    return fail("b200_build_meta: null pointer");
  cudaStream_t st = (cudaStream_t)stream;
  meta_scan_kernel<<<1, 1024, 0, st>>>(step_msg, B, padded_bs, log2_ps, qo_indptr, kv_indptr, last_page_len, scratch);
  if (int rc = check_launch("b200_build_meta(scan)")) return rc;
  // y-chunks: a decode row is <= context_len ids; 8 x 256 threads keep each row copy short
  const int tail_blocks = fill_tail ? 8 : 0;
  dim3 grid((unsigned)(padded_bs + tail_blocks), 8);
  meta_gather_kernel<<<grid, 256, 0, st>>>(slot_table, table_row_stride, step_msg, B, padded_bs, log2_ps, fill_tail,
                                           kv_indptr, scratch, kv_indices, kv_indices_cap, out_slots, out_slots_cap);
  return check_launch("b200_build_meta(gather)");
}
