"""TEST INFRASTRUCTURE ONLY (see oracle/paged_attention_oracle.py header): CPU restatement of the
reference's sampler tail, SURVEY.md section 8 row f-4.  Only tests/, __graft_entry__.smoke() and
bench.py's CPU legs may import this file; the product path never does.

Reference: nanovllmx/layers/sampler.py:19-106 (``Sampler.forward``, ``_apply_top_k_top_p``,
``_apply_min_p``), fed by core/model_runner.py:191-219 (``prepare_sampling_params``: top_k == -1 is
sent as vocab_size, so the reference ALWAYS takes the sort path) and :262-272.

The one step that is not a pure function of its inputs is ``torch.multinomial(probs, 1)``
(sampler.py:53).  For ``num_samples == 1`` ATen implements it as
``argmax(probs / q)`` with ``q = empty_like(probs).exponential_(1)`` drawn from the generator
(aten/src/ATen/native/Distributions.cpp, multinomial_out fast path; checked here against
torch.multinomial on the same seed, tests/test_sampler_oracle.py).  The oracle takes ``q`` as an
explicit input, so the whole tail becomes deterministic and can be pinned.

Pinning: tests/golden/sampler_golden.npz is produced by oracle/gen_golden_sampler.py from the
reference's own ``Sampler`` module run on CPU (TORCHDYNAMO_DISABLE=1) — parity PINNED for this row.
"""
from __future__ import annotations

import numpy as np
import torch


def apply_top_k_top_p(logits: torch.Tensor, p: torch.Tensor | None, k: torch.Tensor | None) -> torch.Tensor:
    """sampler.py:58-90.  ``logits`` fp32 [B, V] (already divided by the temperature)."""
    if p is None and k is None:
        return logits
    logits_sort, logits_idx = logits.sort(dim=-1, descending=False)                       # :65
    if k is not None:
        kth = logits_sort.gather(1, (logits_sort.size(1) - k.to(torch.long)).unsqueeze(1))  # :69-71
        logits_sort = logits_sort.masked_fill(logits_sort < kth, -float("inf"))           # :72-73
    if p is not None:
        probs_sum = logits_sort.softmax(dim=-1).cumsum(dim=-1)                            # :78-79
        mask = probs_sum <= 1 - p.unsqueeze(1)                                            # :80
        mask[:, -1] = False                                                               # :82 at least one
        logits_sort = logits_sort.masked_fill(mask, -float("inf"))                        # :83
    return torch.empty_like(logits_sort).scatter_(-1, logits_idx, logits_sort)            # :86-89


def apply_min_p(logits: torch.Tensor, min_p: torch.Tensor | None) -> torch.Tensor:
    """sampler.py:93-106."""
    if min_p is None:
        return logits
    probs = torch.softmax(logits, dim=-1)
    top = probs.max(dim=-1, keepdim=True).values
    return logits.masked_fill(probs < min_p.unsqueeze(1) * top, -float("inf"))


def masked_logits(logits, temperatures=None, min_ps=None, top_ps=None, top_ks=None) -> torch.Tensor:
    """sampler.py:38-44: fp32 logits / T, then top-k / top-p, then min-p.  Returns fp32 [B, V] with
    -inf at removed tokens."""
    x = logits.to(torch.float).clone()
    if temperatures is not None:
        x.div_(temperatures.to(torch.float).unsqueeze(1))
    x = apply_top_k_top_p(x, top_ps, top_ks)
    return apply_min_p(x, min_ps)


def sample(logits, temperatures=None, min_ps=None, top_ps=None, top_ks=None, q: torch.Tensor | None = None):
    """sampler.py:19-55 with the multinomial draw made explicit: ``q`` is the Exp(1) noise [B, V]
    torch.multinomial would draw.  All-None parameters = the greedy branch (:28-36).
    Returns (ids int64 [B], scores fp32 [B, V]) — scores = probs / q, whose argmax is the id."""
    if temperatures is None and min_ps is None and top_ps is None and top_ks is None:
        return torch.argmax(logits, dim=-1), logits.to(torch.float)
    x = masked_logits(logits, temperatures, min_ps, top_ps, top_ks)
    probs = torch.softmax(x, dim=-1, dtype=torch.float)                                   # :48
    assert q is not None, "the oracle needs the multinomial noise to be deterministic"
    scores = probs / q
    return torch.argmax(scores, dim=-1), scores


# ---- counter-based noise shared with the CUDA kernel's seed mode -------------------------------------
_M0, _M1 = np.uint64(0xD2511F53), np.uint64(0xCD9E8D57)
_W0, _W1 = np.uint32(0x9E3779B9), np.uint32(0xBB67AE85)


def philox4x32_10(ctr: np.ndarray, key: np.ndarray) -> np.ndarray:
    """Philox-4x32-10 (Salmon et al., SC'11).  ctr uint32 [..., 4], key uint32 [..., 2] -> uint32 [..., 4]."""
    c = [ctr[..., i].astype(np.uint32) for i in range(4)]
    k0, k1 = key[..., 0].astype(np.uint32), key[..., 1].astype(np.uint32)
    with np.errstate(over="ignore"):
        for _ in range(10):
            p0 = c[0].astype(np.uint64) * _M0
            p1 = c[2].astype(np.uint64) * _M1
            hi0, lo0 = (p0 >> np.uint64(32)).astype(np.uint32), p0.astype(np.uint32)
            hi1, lo1 = (p1 >> np.uint64(32)).astype(np.uint32), p1.astype(np.uint32)
            c = [hi1 ^ c[1] ^ k0, lo1, hi0 ^ c[3] ^ k1, lo0]
            k0 = (k0 + _W0).astype(np.uint32)
            k1 = (k1 + _W1).astype(np.uint32)
    return np.stack(c, axis=-1)


def seed_noise(seed: int, offset: int, B: int, V: int) -> torch.Tensor:
    """The Exp(1) noise of the kernel's seed mode: element (row b, column v) uses word ``v & 3`` of
    Philox(counter = (v >> 2, b, offset lo, offset hi), key = (seed lo, seed hi));
    u = (word >> 8 + 0.5) * 2^-24 in (0, 1); q = -log(u)."""
    g = np.arange((V + 3) // 4, dtype=np.uint32)
    ctr = np.zeros((B, g.size, 4), dtype=np.uint32)
    ctr[..., 0] = g[None, :]
    ctr[..., 1] = np.arange(B, dtype=np.uint32)[:, None]
    ctr[..., 2] = np.uint32(offset & 0xFFFFFFFF)
    ctr[..., 3] = np.uint32((offset >> 32) & 0xFFFFFFFF)
    key = np.zeros((B, g.size, 2), dtype=np.uint32)
    key[..., 0] = np.uint32(seed & 0xFFFFFFFF)
    key[..., 1] = np.uint32((seed >> 32) & 0xFFFFFFFF)
    w = philox4x32_10(ctr, key).reshape(B, -1)[:, :V]
    u = ((w >> np.uint32(8)).astype(np.float32) + np.float32(0.5)) * np.float32(2.0 ** -24)
    return torch.from_numpy(-np.log(u.astype(np.float32)))
