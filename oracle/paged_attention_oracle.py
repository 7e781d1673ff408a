"""CPU oracle for the attention / KV-cache hot path of nano-vllm-x.

TEST INFRASTRUCTURE ONLY.  Nothing in the product package may import this file;
only ``tests/``, ``__graft_entry__.smoke()`` and ``bench.py``'s ``cpu_baseline`` /
``--impl reference`` legs use it, and there only as the checker / reported baseline.

Every function restates (does not copy) the algorithm of one reference function and
cites it.  Paths are relative to ``/root/reference/nanovllmx``.

Pinning status
--------------
* integer half (slot write, indptr / page table / last-page-len / slot map, CUDA-graph
  buffer contents): PINNED against outputs of the reference's own code run on CPU
  (``oracle/gen_golden.py`` imports ``/root/reference`` with the FlashInfer wrappers
  replaced by recording stubs and the Triton kernel run under ``TRITON_INTERPRET=1``);
  fixtures in ``tests/golden/index_golden.npz``.
* floating-point half (decode / prefill attention): the reference has no tests and its
  arithmetic lives in the un-vendored dependency flashinfer-python (pyproject.toml:19,
  ``>=0.2.0``; image has 0.6.11.post2) which is CUDA-only, so on CPU this part is
  "parity unpinned"; it restates the published semantics (softmax(q k^T * scale) v,
  GQA head mapping h -> h // (Hq/Hkv), bottom-right aligned causal mask as in
  flashinfer ``prefill.cuh:1016-1017``) in fp32.  On the GPU box
  ``tests/test_vs_flashinfer.py`` checks the CUDA path against FlashInfer itself,
  constructed exactly as ``layers/attention.py:49-59`` does.
"""
from __future__ import annotations

from dataclasses import dataclass, field
from typing import Sequence as Seq

import numpy as np
import torch

NEG_INF = float("-inf")


# --------------------------------------------------------------------------------------
# a-1  slot write                                     core/kv_cache.py:12-44, 90-99
# --------------------------------------------------------------------------------------
def store_kvcache(k: torch.Tensor, v: torch.Tensor, k_cache: torch.Tensor,
                  v_cache: torch.Tensor, kv_indices: torch.Tensor) -> None:
    """Scatter rows of this step's K/V into the pool; slot ``-1`` means "skip".

    Reference: Triton program ``store_kvcache_kernel`` (core/kv_cache.py:12-33): one program
    per new token, early return on ``kv_index == -1`` (:25-26), copies ``Hkv*D`` elements
    to ``cache + kv_index * Hkv*D`` (:31-33).  Torch form left in a comment at :91-92.
    Duplicated slot ids: programs run in parallel on the GPU so the winner is unspecified;
    the oracle applies rows in order (last wins) and the tests never duplicate a slot.
    """
    T = k.shape[0]
    assert v.shape[0] == T and kv_indices.numel() == T
    idx = kv_indices.reshape(-1).to(torch.int64)
    for t in range(T):
        s = int(idx[t])
        if s == -1:
            continue
        k_cache[s].copy_(k[t])
        v_cache[s].copy_(v[t])


# --------------------------------------------------------------------------------------
# a-3  eager metadata                                  layers/attention.py:62-141
# --------------------------------------------------------------------------------------
@dataclass
class SeqView:
    """The three fields of ``core/common.py:44-54 Sequence`` the path reads."""
    num_token_ids: int
    kv_indices: list
    cached_kv_len: int = 0


@dataclass
class Metadata:
    qo_indptr: np.ndarray             # int32 [B+1]
    paged_kv_indptr: np.ndarray       # int32 [B+1]
    paged_kv_indices: np.ndarray      # int32 [sum pages]
    paged_kv_last_page_len: np.ndarray  # int32 [B]
    output_kv_indices: np.ndarray     # int64 [T]


def build_metadata(seqs: Seq[SeqView], page_size: int = 1) -> Metadata:
    """Index arrays for one eager step.

    Reference (page_size == 1, the only mode it has; layers/attention.py:66):
      q lens  = len(token_ids) - cached_kv_len                       (:67-71)
      kv lens = len(kv_indices)                                      (:72-76)
      qo_indptr        = [0, cumsum(q lens)]   int32                 (:78-81)
      pages per seq    = ceil(kv len / page_size)                    (:83)
      paged_kv_indptr  = [0, cumsum(pages)]    int32                 (:84-87)
      last_page_len    = kv len % page_size, 0 -> page_size          (:88-90)  (kv len 0 -> page_size)
      paged_kv_indices = concat of every seq's kv_indices, int32     (:92-101)
      output_kv_indices= concat of kv_indices[cached_kv_len:], int64 (:134-141)

    page_size > 1 is NOT defined by the reference (its concat would emit token slots while
    the indptr counts pages).  This build defines it as page-aligned
    allocation: token j of a sequence lives in page ``kv_indices[j] // page_size`` at
    offset ``kv_indices[j] % page_size`` and ``kv_indices[p*page_size] % page_size == 0``;
    the page table entry for page p is ``kv_indices[p*page_size] // page_size``.
    """
    B = len(seqs)
    q_lens = np.array([s.num_token_ids - s.cached_kv_len for s in seqs], dtype=np.int64)
    kv_lens = np.array([len(s.kv_indices) for s in seqs], dtype=np.int64)
    qo_indptr = np.zeros(B + 1, dtype=np.int32)
    qo_indptr[1:] = np.cumsum(q_lens).astype(np.int32)
    pages = kv_lens // page_size + (kv_lens % page_size != 0)
    kv_indptr = np.zeros(B + 1, dtype=np.int32)
    kv_indptr[1:] = np.cumsum(pages).astype(np.int32)
    last = (kv_lens % page_size).astype(np.int32)
    last[last == 0] = page_size
    chunks = []
    for s in seqs:
        a = np.asarray(s.kv_indices, dtype=np.int64)
        if page_size == 1:
            chunks.append(a.astype(np.int32))
        else:
            first = a[::page_size]
            assert np.all(first % page_size == 0), "page_size>1 needs page-aligned slots"
            chunks.append((first // page_size).astype(np.int32))
    kv_indices = np.concatenate(chunks) if chunks else np.zeros(0, np.int32)
    outs = [np.asarray(s.kv_indices[s.cached_kv_len:], dtype=np.int64) for s in seqs]
    out_idx = np.concatenate(outs) if outs else np.zeros(0, np.int64)
    return Metadata(qo_indptr, kv_indptr, kv_indices.astype(np.int32), last, out_idx)


# --------------------------------------------------------------------------------------
# a-4  CUDA-graph buffers                              layers/attention.py:151-168, 219-275
# --------------------------------------------------------------------------------------
@dataclass
class GraphBuffers:
    paged_kv_indptr_buffer: np.ndarray         # int32 [max_bs+1]
    paged_kv_indices_buffer: np.ndarray        # int32 [max_bs*context_len]
    paged_kv_last_page_len_buffer: np.ndarray  # int32 [max_bs]
    output_kv_indices_buffer: np.ndarray       # int64 [max_bs]


def prepare_graph_buffers(max_bs: int, context_len: int) -> GraphBuffers:
    """Initial contents of the four named static buffers (layers/attention.py:158-168):
    indptr = arange(max_bs+1)*context_len, indices = 0, last_page_len = 1, slot map = 0."""
    return GraphBuffers(
        (np.arange(max_bs + 1, dtype=np.int64) * context_len).astype(np.int32),
        np.zeros(max_bs * context_len, dtype=np.int32),
        np.ones(max_bs, dtype=np.int32),
        np.zeros(max_bs, dtype=np.int64),
    )


def replay_graph_buffers(buf: GraphBuffers, seqs: Seq[SeqView], padded_bs: int) -> None:
    """In-place refill for one replayed decode step (layers/attention.py:219-275).

    * kv lens of the ``padded_bs - bs`` padding sequences are 1                     (:239-244)
    * indptr[:padded_bs+1] <- [0, cumsum(kv lens)], entries beyond stay untouched   (:245-250)
    * the WHOLE indices buffer is zeroed, then its first sum(len) entries receive the
      concatenated slot ids, so padding sequences read slot 0                       (:258-261)
    * last_page_len buffer is never rewritten (stays all ones at page_size 1)
    * the WHOLE slot-map buffer is set to -1, then [:bs] <- the new-token slots     (:265-275)
    """
    bs = len(seqs)
    lens = np.array([len(s.kv_indices) for s in seqs] + [1] * (padded_bs - bs), dtype=np.int64)
    buf.paged_kv_indptr_buffer[0] = 0
    buf.paged_kv_indptr_buffer[1:padded_bs + 1] = np.cumsum(lens).astype(np.int32)
    buf.paged_kv_indices_buffer[:] = 0
    flat = np.concatenate([np.asarray(s.kv_indices, dtype=np.int64) for s in seqs]).astype(np.int32)
    buf.paged_kv_indices_buffer[:flat.size] = flat
    buf.output_kv_indices_buffer[:] = -1
    new = np.concatenate([np.asarray(s.kv_indices[s.cached_kv_len:], dtype=np.int64) for s in seqs])
    assert new.size == bs, "decode replay writes exactly one new slot per sequence"
    buf.output_kv_indices_buffer[:bs] = new


# --------------------------------------------------------------------------------------
# a-6 / a-7  attention                                 layers/attention.py:332-373
# --------------------------------------------------------------------------------------
def _token_slots(kv_indptr, kv_indices, last_page_len, i: int, page_size: int) -> np.ndarray:
    pages = np.asarray(kv_indices[int(kv_indptr[i]):int(kv_indptr[i + 1])], dtype=np.int64)
    if pages.size == 0:
        return np.zeros(0, np.int64)
    n = (pages.size - 1) * page_size + int(last_page_len[i])
    slots = (pages[:, None] * page_size + np.arange(page_size)[None, :]).reshape(-1)
    return slots[:n]


def paged_attention(q: torch.Tensor, k_cache: torch.Tensor, v_cache: torch.Tensor,
                    qo_indptr, kv_indptr, kv_indices, last_page_len,
                    page_size: int = 1, sm_scale: float | None = None,
                    causal: bool = True, out_dtype: torch.dtype | None = None,
                    return_fp32: bool = False):
    """fp32 gather-then-softmax attention over a paged KV pool.

    q: [T, Hq, D]; k_cache / v_cache: [num_tokens, Hkv, D] (NHD, token granular; a page of
    size ps is ps consecutive token slots).  Sequence i owns query rows
    qo_indptr[i]:qo_indptr[i+1] and the KV tokens named by its pages.

    Reference semantics (layers/attention.py:355-369 -> FlashInfer run):
      o[h] = softmax(scale * q[h] . K[g(h)]^T) V[g(h)],  g(h) = h // (Hq // Hkv),
      scale = head_dim ** -0.5 (:324), query j of a sequence with Lq new and Lkv total
      tokens sees kv positions <= Lkv - Lq + j (bottom-right causal; decode is Lq == 1 so
      nothing is masked).  A query row with no visible key yields zeros.
    """
    T, Hq, D = q.shape
    Hkv = k_cache.shape[-2]
    group = Hq // Hkv
    scale = (D ** -0.5) if sm_scale is None else sm_scale
    out = torch.zeros((T, Hq, D), dtype=torch.float32)
    B = len(qo_indptr) - 1
    kf = k_cache.reshape(-1, Hkv, D)
    vf = v_cache.reshape(-1, Hkv, D)
    for i in range(B):
        q0, q1 = int(qo_indptr[i]), int(qo_indptr[i + 1])
        lq = q1 - q0
        if lq == 0:
            continue
        slots = torch.from_numpy(_token_slots(kv_indptr, kv_indices, last_page_len, i, page_size))
        lkv = slots.numel()
        if lkv == 0:
            continue
        K = kf[slots].to(torch.float32).repeat_interleave(group, dim=1)   # [Lkv, Hq, D]
        V = vf[slots].to(torch.float32).repeat_interleave(group, dim=1)
        Q = q[q0:q1].to(torch.float32)                                    # [Lq, Hq, D]
        s = torch.einsum("qhd,khd->hqk", Q, K) * scale
        if causal:
            qi = torch.arange(lq)[:, None]
            ki = torch.arange(lkv)[None, :]
            s = s.masked_fill((ki > (lkv - lq + qi))[None], NEG_INF)
        m = s.max(dim=-1, keepdim=True).values
        m = torch.where(torch.isinf(m), torch.zeros_like(m), m)
        p = torch.exp(s - m)
        l = p.sum(dim=-1, keepdim=True)
        p = torch.where(l > 0, p / l.clamp_min(1e-38), torch.zeros_like(p))
        out[q0:q1] = torch.einsum("hqk,khd->qhd", p, V)
    if return_fp32:
        return out
    return out.to(out_dtype or q.dtype)


def decode_attention(q, k_cache, v_cache, kv_indptr, kv_indices, last_page_len,
                     page_size: int = 1, sm_scale: float | None = None, **kw):
    """DECODE branch (layers/attention.py:365-369): one query row per sequence, no mask."""
    B = q.shape[0]
    qo_indptr = np.arange(B + 1, dtype=np.int32)
    return paged_attention(q, k_cache, v_cache, qo_indptr, kv_indptr, kv_indices, last_page_len,
                           page_size, sm_scale, causal=False, **kw)


def prefill_attention(q, k_cache, v_cache, qo_indptr, kv_indptr, kv_indices, last_page_len,
                      page_size: int = 1, sm_scale: float | None = None, **kw):
    """PREFILL branch (layers/attention.py:360-364): ragged, causal, prefix + new tokens."""
    return paged_attention(q, k_cache, v_cache, qo_indptr, kv_indptr, kv_indices, last_page_len,
                           page_size, sm_scale, causal=True, **kw)


# --------------------------------------------------------------------------------------
# synthetic workloads; shared by tests and bench so both sides see the same bytes
# --------------------------------------------------------------------------------------
# --------------------------------------------------------------------------------------
# f-3  q/k RMSNorm + RoPE + slot write (the step just ahead of Attention.forward)
#      models/qwen3.py:122-134, layers/layernorm.py:16-36, layers/rotary_embedding.py:6-24, 73-102
# --------------------------------------------------------------------------------------
def qk_norm_rope_store(qkv: torch.Tensor, positions: torch.Tensor, q_weight: torch.Tensor,
                       k_weight: torch.Tensor, eps: float, cos_sin_cache: torch.Tensor,
                       k_cache: torch.Tensor, v_cache: torch.Tensor, slots: torch.Tensor,
                       Hq: int, Hkv: int, D: int = 128):
    """Returns (q, k) after per-head RMSNorm + rotary embedding, and scatters (k, v) into the pool.

    Restates, per token and head (fp32 arithmetic, ONE rounding to the activation dtype per module
    output, which is what the reference's ``@torch.compile``d modules produce on the GPU; run
    eagerly the RoPE products round to bf16 three times, <= 2 bf16 ulp away -- the golden test
    bounds that difference):
      * ``qkv.split([q_size, kv_size, kv_size])`` and the per-head views (models/qwen3.py:123-132);
      * RMSNorm without residual (layers/layernorm.py:24-33): ``x.float()``,
        ``x * rsqrt(mean(x^2) + eps)``, ``(x * weight).to(dtype)``;
      * RotaryEmbedding.forward (layers/rotary_embedding.py:87-101) with ``rotary_dim == head_size``:
        ``cos, sin = cos_sin_cache[pos].chunk(2)``, cast to the activation dtype (:16-17),
        ``x1, x2 = chunk(x, 2)``, ``o1 = x1*cos - x2*sin``, ``o2 = x2*cos + x1*sin`` and the
        INTERLEAVED output ``stack((o1, o2), -1).flatten(-2)`` (:19-24);
      * the slot write of Attention.forward (layers/attention.py:347-352): see store_kvcache.
    """
    T = qkv.shape[0]
    dt = qkv.dtype
    q, k, v = qkv.split([Hq * D, Hkv * D, Hkv * D], dim=-1)

    def norm(x, w):
        xf = x.reshape(T, -1, D).float()
        var = xf.pow(2).mean(dim=-1, keepdim=True)
        xf = xf * torch.rsqrt(var + eps)
        return (xf * w.float()).to(dt)

    cs = cos_sin_cache.index_select(0, positions.reshape(-1).long())
    cos, sin = cs.chunk(2, dim=-1)
    cos = cos.unsqueeze(-2).to(dt).float()
    sin = sin.unsqueeze(-2).to(dt).float()

    def rope(x):
        x1, x2 = torch.chunk(x.float(), 2, dim=-1)
        o1 = x1 * cos - x2 * sin
        o2 = x2 * cos + x1 * sin
        return torch.stack((o1, o2), dim=-1).flatten(-2).to(dt)

    qo = rope(norm(q, q_weight))
    ko = rope(norm(k, k_weight))
    store_kvcache(ko, v.reshape(T, Hkv, D), k_cache, v_cache, slots)
    return qo.reshape(T, Hq * D), ko.reshape(T, Hkv * D)


def bench_workload(num_seqs: int = 256, max_input_len: int = 1024, max_output_len: int = 1024,
                   seed_value: int = 0):
    """Request generator of the reference's bench.py:14-29 (python ``random`` with seed 0:
    prompt ids ``randint(0,10000)`` of length ``randint(100,1024)``, then
    ``max_tokens=randint(100,1024)`` drawn AFTER all prompts).  Known answer
    (README.md:99-100): 256 prompts, sum(max_tokens) == 133966."""
    import random
    rng = random.Random(seed_value)
    prompts = [[rng.randint(0, 10000) for _ in range(rng.randint(100, max_input_len))]
               for _ in range(num_seqs)]
    max_tokens = [rng.randint(100, max_output_len) for _ in range(num_seqs)]
    return prompts, max_tokens
