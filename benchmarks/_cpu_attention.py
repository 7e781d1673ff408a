"""CPU stand-in for nanovllmx/layers/attention.py + KVCachePool built on the oracle (TEST INFRASTRUCTURE: used only
by ``ref_engine.py --backend oracle`` to exercise the harness plumbing around the stock reference engine without a
GPU, and by the CPU tests of that harness).  Same surface as the reference (layers/attention.py:17-373)."""
from __future__ import annotations

from contextlib import contextmanager
from dataclasses import dataclass

import torch
import torch.nn as nn

from oracle import paged_attention_oracle as orc


class KVCachePool:
    def __init__(self, dtype, device, num_tokens, num_layers, num_heads, head_dim, start_layer=None, end_layer=None):
        self.start_layer = start_layer or 0
        self.k_cache = torch.zeros((num_layers, num_tokens, num_heads, head_dim), dtype=dtype)
        self.v_cache = torch.zeros_like(self.k_cache)

    def get_kv_cache(self, layer, index=None):
        return self.k_cache[layer - self.start_layer], self.v_cache[layer - self.start_layer]

    def set_kv_cache(self, layer, index, k, v):
        orc.store_kvcache(k, v, self.k_cache[layer - self.start_layer], self.v_cache[layer - self.start_layer], index)


@dataclass
class AttentionMetadata:
    forward_mode: object
    kv_cache: KVCachePool
    output_kv_indices: torch.Tensor
    wrapper: object


class AttentionBackend:
    def __init__(self, kv_cache, num_qo_heads, num_kv_heads, head_dim, dtype, device):
        self.kv_cache = kv_cache

    def build_metadata(self, batch):
        m = orc.build_metadata([orc.SeqView(len(s.token_ids), s.kv_indices, s.cached_kv_len) for s in batch.seqs])
        return AttentionMetadata(batch.forward_mode, self.kv_cache, torch.from_numpy(m.output_kv_indices), m)


@contextmanager
def attention_kv_cache(model, metadata):
    mods = [m for m in model.modules() if isinstance(m, Attention)]
    for m in mods:
        m.attention_metadata = metadata
    yield
    for m in mods:
        m.attention_metadata = None


class Attention(nn.Module):
    def __init__(self, num_heads, head_dim, scaling, num_kv_heads, layer_id):
        super().__init__()
        self.num_heads, self.head_dim, self.num_kv_heads, self.layer_id = num_heads, head_dim, num_kv_heads, layer_id
        self.scaling = head_dim ** -0.5 if scaling is None else scaling
        self.attention_metadata = None

    def forward(self, q, k, v, save_kv_cache=True):
        md = self.attention_metadata
        if k is not None and save_kv_cache:
            md.kv_cache.set_kv_cache(self.layer_id, md.output_kv_indices, k.reshape(-1, self.num_kv_heads, self.head_dim),
                                     v.reshape(-1, self.num_kv_heads, self.head_dim))
        kc, vc = md.kv_cache.get_kv_cache(self.layer_id)
        m = md.wrapper
        o = orc.paged_attention(q.reshape(-1, self.num_heads, self.head_dim), kc, vc, m.qo_indptr, m.paged_kv_indptr,
                                m.paged_kv_indices, m.paged_kv_last_page_len, 1, self.scaling, causal=True)
        return o.reshape(-1, self.num_heads * self.head_dim)
