"""GPU: the CUDA path vs the reference's actual arithmetic library — FlashInfer wrappers constructed
exactly as nanovllmx/layers/attention.py:49-59 does ("NHD", backend="auto", decode with
use_tensor_cores=False), planned as :106-132.  Skipped when FlashInfer cannot JIT/load on the box."""
import pytest
import torch

from tests.util import ATOL, dev, make_paged_case

pytestmark = pytest.mark.gpu


def _flashinfer():
    try:
        import flashinfer
        return flashinfer
    except Exception as e:  # pragma: no cover
        pytest.skip(f"flashinfer unavailable: {e}")


@pytest.mark.timeout(900)
@pytest.mark.parametrize("Hq,Hkv", [(16, 8), (32, 8)])
def test_decode_matches_flashinfer(pkg, Hq, Hkv):
    fi = _flashinfer()
    case = make_paged_case([300, 5, 77, 1024, 2000], Hq, Hkv, seed=4)
    d = dev(case)
    try:
        ws = torch.empty(128 * 1024 * 1024, dtype=torch.uint8, device="cuda")
        w = fi.BatchDecodeWithPagedKVCacheWrapper(ws, "NHD", use_tensor_cores=False, backend="auto")
        w.plan(indptr=d["kv_indptr"], indices=d["kv_indices"], last_page_len=d["last_page_len"], num_qo_heads=Hq,
               num_kv_heads=Hkv, head_dim=128, page_size=1, q_data_type=torch.bfloat16, kv_data_type=torch.bfloat16)
        w._sm_scale = 128 ** -0.5
        ref = w.run(d["q"], (d["k_cache"].view(-1, 1, Hkv, 128), d["v_cache"].view(-1, 1, Hkv, 128)))
    except Exception as e:  # JIT toolchain / cubin problems are not this repo's failure
        pytest.skip(f"flashinfer decode could not run here: {type(e).__name__}: {e}")
    mine = pkg.ops.decode(d["q"], d["k_cache"], d["v_cache"], d["kv_indptr"], d["kv_indices"], d["last_page_len"],
                          torch.empty(pkg.ops.decode_workspace_bytes(5, Hq, Hkv, 128), dtype=torch.uint8, device="cuda"),
                          1, 128 ** -0.5)
    assert (mine.float() - ref.float()).abs().max().item() <= ATOL


@pytest.mark.timeout(900)
def test_prefill_matches_flashinfer(pkg):
    fi = _flashinfer()
    Hq, Hkv = 16, 8
    q_lens, prefix = [8, 1, 93, 64, 300], [32, 32, 7, 0, 0]
    case = make_paged_case([a + b for a, b in zip(q_lens, prefix)], Hq, Hkv, q_lens=q_lens, seed=6)
    d = dev(case)
    try:
        ws = torch.empty(128 * 1024 * 1024, dtype=torch.uint8, device="cuda")
        w = fi.BatchPrefillWithPagedKVCacheWrapper(ws, "NHD", backend="auto")
        w.plan(qo_indptr=d["qo_indptr"], paged_kv_indices=d["kv_indices"], paged_kv_indptr=d["kv_indptr"],
               paged_kv_last_page_len=d["last_page_len"], num_qo_heads=Hq, num_kv_heads=Hkv, head_dim_qk=128,
               head_dim_vo=128, page_size=1, causal=True, q_data_type=torch.bfloat16, kv_data_type=torch.bfloat16)
        w._sm_scale = 128 ** -0.5
        ref = w.run(d["q"], (d["k_cache"].view(-1, 1, Hkv, 128), d["v_cache"].view(-1, 1, Hkv, 128)))
    except Exception as e:
        pytest.skip(f"flashinfer prefill could not run here: {type(e).__name__}: {e}")
    mine = pkg.ops.prefill(d["q"], d["k_cache"], d["v_cache"], d["qo_indptr"], d["kv_indptr"], d["kv_indices"],
                           d["last_page_len"], torch.empty(pkg.ops.prefill_workspace_bytes(case["B"], sum(q_lens), Hq, Hkv, 128),
                                                               dtype=torch.uint8, device="cuda"), 1, 128 ** -0.5, True)
    assert (mine.float() - ref.float()).abs().max().item() <= ATOL
