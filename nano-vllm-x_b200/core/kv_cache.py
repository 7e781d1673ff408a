"""KV pool + slot write, drop-in for the hot half of core/kv_cache.py (reference :12-99).

``KVCachePool`` keeps the reference's layout contract — two tensors
``[L_local, num_tokens, Hkv_local, D]`` (NHD, token granular; core/kv_cache.py:69-79) — and
method surface (``get_kv_cache`` :81-88, ``set_kv_cache`` :90-99).  ``store_kvcache`` replaces the
Triton launcher (:36-44) with the 16-byte-vector CUDA kernel behind ``b200_store_kv``.
The host-side allocator / radix tree (core/kv_cache.py:101-369) is unchanged reference code
and is not re-implemented here.
"""
from __future__ import annotations

from typing import Optional

import torch

from .. import ops


def store_kvcache(k: torch.Tensor, v: torch.Tensor, k_cache: torch.Tensor, v_cache: torch.Tensor,
                  kv_indices: torch.Tensor) -> None:
    """k_cache[kv_indices[t]] = k[t]; v likewise; index -1 is skipped (core/kv_cache.py:25-26).

    Same argument contract as the reference (asserts :39-42): k, v ``[T, Hkv, D]`` with unit
    inner stride and ``stride(-2) == D`` (any row stride: v is a view into the qkv buffer),
    caches ``[num_tokens, Hkv, D]`` contiguous, ``kv_indices`` int64 ``[T]``.
    Unlike the Triton kernel there is no power-of-two restriction on ``Hkv * D``.
    """
    num_tokens, num_heads, head_dim = k.shape
    block = num_heads * head_dim
    assert k.stride(-1) == 1 and v.stride(-1) == 1
    assert k.stride(-2) == head_dim and v.stride(-2) == head_dim
    assert k_cache.stride(-3) == block and v_cache.stride(-3) == block
    assert kv_indices.numel() == num_tokens
    if kv_indices.dtype != torch.int64:
        kv_indices = kv_indices.to(torch.int64)
    ops.call("store_kv", k, v, k_cache, v_cache, kv_indices.contiguous())


class KVCachePool:
    def __init__(self, dtype: torch.dtype, device: torch.device, num_tokens: int, num_layers: int,
                 num_heads: int, head_dim: int, start_layer: Optional[int] = None,
                 end_layer: Optional[int] = None):
        self.dtype = dtype
        self.device = device
        self.num_tokens = num_tokens
        self.num_layers = num_layers
        self.num_heads = num_heads
        self.head_dim = head_dim
        self.start_layer = start_layer or 0
        self.end_layer = end_layer or num_layers - 1
        self.create_cache_pool()

    def create_cache_pool(self) -> None:
        shape = (self.num_layers, self.num_tokens, self.num_heads, self.head_dim)
        self.k_cache = torch.empty(shape, dtype=self.dtype, device=self.device)
        self.v_cache = torch.empty(shape, dtype=self.dtype, device=self.device)

    def get_kv_cache(self, layer: int, index: Optional[torch.Tensor] = None):
        li = layer - self.start_layer
        if index is None:
            return self.k_cache[li], self.v_cache[li]
        return self.k_cache[li, index], self.v_cache[li, index]

    def set_kv_cache(self, layer: int, index: torch.Tensor, k_cache: torch.Tensor, v_cache: torch.Tensor) -> None:
        li = layer - self.start_layer
        store_kvcache(k_cache, v_cache, self.k_cache[li], self.v_cache[li], index)
