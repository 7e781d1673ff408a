"""nano-vllm-x_b200 — B200-native attention / KV-cache hot path for nano-vllm-x.

The directory name follows the project layout contract and is not a Python identifier; import it
with ``importlib.import_module("nano-vllm-x_b200")`` or through the one-line alias module
``nanovllmx_b200`` at the repo root.  Layout mirrors the two reference files it replaces:

  layers/attention.py  <- nanovllmx/layers/attention.py   (Attention, AttentionBackend, ...)
  core/kv_cache.py     <- nanovllmx/core/kv_cache.py:12-99 (KVCachePool, store_kvcache)
  core/cuda_graph.py   <- nanovllmx/core/cuda_graph.py     (graph registry used by the API)
  layers/sampler.py    <- nanovllmx/layers/sampler.py      (Sampler)
  ops.py / _lib.py        torch.library ops over the C ABI of include/b200attn.h
  csrc/                   CUDA kernels for sm_100a + the extern "C" entry points
"""
from . import _lib            # noqa: F401
from . import ops             # noqa: F401  (registers torch.ops.b200attn.*; loads libb200attn.so lazily)
from .core.common import ForwardBatch, ForwardMode, Sequence       # noqa: F401
from .core.cuda_graph import CUDAGraph                             # noqa: F401
from .core.kv_cache import KVCachePool, store_kvcache              # noqa: F401
from .layers.attention import (Attention, AttentionBackend, AttentionMetadata, AttentionPlan,  # noqa: F401
                               SlotTable, attention_kv_cache)
from .layers.sampler import Sampler                                # noqa: F401



def shard_heads(num_qo_heads: int, num_kv_heads: int, tp_size: int, tp_rank: int):
    """Head ranges owned by one tensor-parallel rank, the partitioning the kernels inherit
    (reference: core/model_runner.py:43-44, models/qwen3.py:73-86; KV heads are replicated when
    tp_size > num_kv_heads, layers/linear.py:333-338).  No collective is needed on the attention path."""
    assert num_qo_heads % tp_size == 0
    hq = num_qo_heads // tp_size
    q_heads = range(tp_rank * hq, (tp_rank + 1) * hq)
    if num_kv_heads >= tp_size:
        assert num_kv_heads % tp_size == 0
        hkv = num_kv_heads // tp_size
        kv_heads = range(tp_rank * hkv, (tp_rank + 1) * hkv)
    else:
        assert tp_size % num_kv_heads == 0
        kv_heads = range(tp_rank // (tp_size // num_kv_heads), tp_rank // (tp_size // num_kv_heads) + 1)
    return q_heads, kv_heads


__all__ = ["shard_heads", "Attention", "AttentionBackend", "AttentionMetadata", "AttentionPlan", "SlotTable",
           "attention_kv_cache", "KVCachePool", "store_kvcache", "CUDAGraph", "ForwardBatch", "ForwardMode",
           "Sequence", "Sampler", "ops"]
