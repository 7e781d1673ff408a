"""ctypes binding of oracle/_build/liboracle.so (the plain-C oracle).  Test infrastructure only."""
import ctypes
import os
import subprocess

import numpy as np
import torch

_HERE = os.path.dirname(os.path.abspath(__file__))
_SO = os.path.join(_HERE, "_build", "liboracle.so")
_lib = None


def load():
    global _lib
    if _lib is None:
        if not os.path.exists(_SO):
            subprocess.run(["make", "-C", _HERE], check=True, capture_output=True)
        _lib = ctypes.CDLL(_SO)
    return _lib


def _p(a):
    return ctypes.c_void_p(a.ctypes.data if isinstance(a, np.ndarray) else a.data_ptr())


def paged_attention(q, k_cache, v_cache, qo_indptr, kv_indptr, kv_indices, last_page_len, page_size=1,
                    sm_scale=None, causal=True, threads=None):
    """bf16 torch tensors + int32 numpy index arrays -> fp32 torch tensor [T, Hq, D]."""
    T, Hq, D = q.shape
    Hkv = k_cache.shape[-2]
    q, k_cache, v_cache = q.contiguous(), k_cache.contiguous(), v_cache.contiguous()
    out = torch.empty((T, Hq, D), dtype=torch.float32)
    arrs = [np.ascontiguousarray(a, dtype=np.int32) for a in (qo_indptr, kv_indptr, kv_indices, last_page_len)]
    load().oracle_paged_attention(_p(q), _p(k_cache), _p(v_cache), _p(arrs[0]), _p(arrs[1]), _p(arrs[2]), _p(arrs[3]),
                                  len(arrs[0]) - 1, Hq, Hkv, D, page_size,
                                  ctypes.c_float(D ** -0.5 if sm_scale is None else sm_scale), int(causal), _p(out),
                                  int(threads or os.cpu_count() or 1))
    return out


def build_metadata(num_token_ids, cached, kv_lens, kv_flat, page_size=1):
    B = len(kv_lens)
    a = [np.ascontiguousarray(x, dtype=np.int64) for x in (num_token_ids, cached, kv_lens, kv_flat)]
    pages = int(sum(-(-int(l) // page_size) for l in kv_lens))
    new = int(sum(max(0, int(l) - int(c)) for l, c in zip(kv_lens, cached)))
    qo, kvp = np.zeros(B + 1, np.int32), np.zeros(B + 1, np.int32)
    idx, last, outs = np.zeros(pages, np.int32), np.zeros(B, np.int32), np.zeros(new, np.int64)
    rc = load().oracle_build_metadata(_p(a[0]), _p(a[1]), _p(a[2]), _p(a[3]), B, page_size, _p(qo), _p(kvp), _p(idx),
                                      _p(last), _p(outs))
    assert rc == 0
    return qo, kvp, idx, last, outs
