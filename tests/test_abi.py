"""CPU: the C-ABI library loads and exports every symbol include/b200attn.h declares."""
import ctypes
import os
import re

from tests.conftest import ROOT


def _declared():
    src = open(os.path.join(ROOT, "include", "b200attn.h")).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(b200_[a-z0-9_]+)\s*\(", src)))


def test_header_symbols_exported(pkg):
    names = _declared()
    assert "b200_decode" in names and "b200_store_kv" in names and len(names) >= 12
    lib = ctypes.CDLL(pkg._lib.LIB_PATH)
    for n in names:
        assert hasattr(lib, n), f"{n} declared in include/b200attn.h but not exported"
    assert sorted(pkg._lib.SIGNATURES) == names, "ctypes signature table out of sync with the header"


def test_loader_and_error_path(pkg):
    lib = pkg._lib.load()
    assert lib.b200_abi_version() == 1
    # argument validation happens before any CUDA call, so it is testable without a GPU
    rc = lib.b200_store_kv(None, 0, None, 0, None, None, None, 4, 256, 2, None)
    assert rc != 0 and b"null" in lib.b200_last_error()
    rc = lib.b200_build_meta(None, 0, None, 1, 1, 3, 0, None, None, None, 0, None, None, 0, None, None)
    assert rc != 0 and b"power of two" in lib.b200_last_error()
    rc = lib.b200_qk_norm_rope_store(None, 0, None, None, None, 1e-6, None, 0, 0, None, 0, None, 0, None, None, None, 4, 16, 8, 64,
                                     0, None)
    assert rc != 0 and b"head_dim 128" in lib.b200_last_error()
    rc = lib.b200_decode(None, 0, None, None, 0, None, None, None, None, 0, None, 0, 1, 8, 8, 64, 1, 1.0, 0, None)
    assert rc != 0 and b"head_dim" in lib.b200_last_error()


def test_ops_registered_and_no_cpu_fallback(pkg):
    import pytest
    import torch
    for name in ("store_kv", "slot_table_update", "build_meta", "step_meta", "decode", "prefill", "qk_norm_rope_store"):
        assert hasattr(torch.ops.b200attn, name)
    k = torch.zeros(2, 1, 8, dtype=torch.bfloat16)
    with pytest.raises(Exception):
        pkg.store_kvcache(k, k, torch.zeros(4, 1, 8, dtype=torch.bfloat16), torch.zeros(4, 1, 8, dtype=torch.bfloat16),
                          torch.tensor([0, 1]))
    with pytest.raises(RuntimeError):
        pkg.AttentionBackend(None, 4, 2, 128, torch.bfloat16, torch.device("cpu"))
