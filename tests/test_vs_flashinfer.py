"""GPU: the CUDA path vs the reference's actual arithmetic library — FlashInfer wrappers constructed
exactly as nanovllmx/layers/attention.py:49-59 does ("NHD", backend="auto", decode with
use_tensor_cores=False), planned as :106-132.

This file PINS the floating-point half of the oracle (the reference has no tests of its own), so on a
CUDA box a FlashInfer that cannot import, JIT or run is a FAILURE here, never a skip: a green run without these
comparisons would leave parity "unpinned" silently."""
import pytest
import torch

from tests.util import ATOL, RTOL, dev, make_paged_case

pytestmark = pytest.mark.gpu


def _flashinfer():
    try:
        import flashinfer
        return flashinfer
    except Exception as e:  # pragma: no cover
        pytest.fail(f"flashinfer cannot be imported on this CUDA box, the FP oracle is unpinned: {type(e).__name__}: {e}")


def _fi_decode(fi, q, k_cache, v_cache, indptr, indices, last, Hq, Hkv, ps):
    try:
        ws = torch.empty(128 * 1024 * 1024, dtype=torch.uint8, device="cuda")
        w = fi.BatchDecodeWithPagedKVCacheWrapper(ws, "NHD", use_tensor_cores=False, backend="auto")
        w.plan(indptr=indptr, indices=indices, last_page_len=last, num_qo_heads=Hq, num_kv_heads=Hkv, head_dim=128,
               page_size=ps, q_data_type=torch.bfloat16, kv_data_type=torch.bfloat16)
        w._sm_scale = 128 ** -0.5
        return w.run(q, (k_cache.view(-1, ps, Hkv, 128), v_cache.view(-1, ps, Hkv, 128)))
    except Exception as e:
        pytest.fail(f"flashinfer decode could not run on this CUDA box, the FP oracle is unpinned: {type(e).__name__}: {e}")


def _close(mine, ref, what):
    d = (mine.float() - ref.float())
    assert d.abs().max().item() <= ATOL, f"{what}: max abs diff {d.abs().max().item():.4g}"
    assert (d.norm() / ref.float().norm()).item() <= RTOL, f"{what}: rel diff {(d.norm() / ref.float().norm()).item():.4g}"


@pytest.mark.timeout(900)
@pytest.mark.parametrize("Hq,Hkv", [(16, 8), (32, 8)])
def test_decode_matches_flashinfer(pkg, Hq, Hkv):
    fi = _flashinfer()
    case = make_paged_case([300, 5, 77, 1024, 2000], Hq, Hkv, seed=4)
    d = dev(case)
    ref = _fi_decode(fi, d["q"], d["k_cache"], d["v_cache"], d["kv_indptr"], d["kv_indices"], d["last_page_len"], Hq, Hkv, 1)
    mine = pkg.ops.decode(d["q"], d["k_cache"], d["v_cache"], d["kv_indptr"], d["kv_indices"], d["last_page_len"],
                          torch.empty(pkg.ops.decode_workspace_bytes(5, Hq, Hkv, 128), dtype=torch.uint8, device="cuda"),
                          1, 128 ** -0.5)
    _close(mine, ref, "decode")


@pytest.mark.timeout(1800)
@pytest.mark.parametrize("ps", [16, 1])
@pytest.mark.parametrize("Hq,Hkv", [(32, 8), (16, 4), (8, 2), (4, 1)])
def test_config5_full_shape_every_row_matches_flashinfer(pkg, Hq, Hkv, ps):
    """BASELINE.json configs[4] at the FULL headline shape (B = 64 x 32 768 tokens, fragmented page table) for every
    TP shape of Qwen3-8B (TP 1/2/4/8: Hq/Hkv = 32/8, 16/4, 8/2, 4/1), page 16 (the bench configuration) and page 1
    (the reference's drop-in mode): EVERY output row against FlashInfer planned with that page size
    (flashinfer/decode.py:1509-1555), within the north_star tolerance."""
    fi = _flashinfer()
    B, L, D = 64, 32768, 128
    g = torch.Generator(device="cuda").manual_seed(1000 * Hq + ps)
    pages_per_seq = L // ps
    n_pages = B * pages_per_seq
    k_cache = torch.randn((n_pages * ps, Hkv, D), device="cuda", dtype=torch.bfloat16, generator=g)
    v_cache = torch.randn((n_pages * ps, Hkv, D), device="cuda", dtype=torch.bfloat16, generator=g)
    q = torch.randn((B, Hq, D), device="cuda", dtype=torch.bfloat16, generator=g)
    indices = torch.randperm(n_pages, device="cuda", generator=g).to(torch.int32)
    indptr = torch.arange(B + 1, device="cuda", dtype=torch.int32) * pages_per_seq
    last = torch.full((B,), ps, device="cuda", dtype=torch.int32)
    ref = _fi_decode(fi, q, k_cache, v_cache, indptr, indices, last, Hq, Hkv, ps)
    ws = torch.empty(pkg.ops.decode_workspace_bytes(B, Hq, Hkv, D), dtype=torch.uint8, device="cuda")
    mine = pkg.ops.decode(q, k_cache, v_cache, indptr, indices, last, ws, ps, D ** -0.5)
    assert mine.shape == ref.shape == (B, Hq, D)
    _close(mine, ref, f"config 5 Hq{Hq}/Hkv{Hkv} page {ps}")
    # and two rows against a direct fp32 computation (pins both libraries to the oracle's arithmetic at this size)
    for b, h in ((0, 0), (B - 1, Hq - 1)):
        pg = indices[b * pages_per_seq:(b + 1) * pages_per_seq].long()
        slots = (pg[:, None] * ps + torch.arange(ps, device="cuda")[None]).view(-1)
        K, V = k_cache[slots, h // (Hq // Hkv)].float(), v_cache[slots, h // (Hq // Hkv)].float()
        o = torch.softmax((K @ q[b, h].float()) * D ** -0.5, 0) @ V
        assert (mine[b, h].float() - o).abs().max().item() <= ATOL


@pytest.mark.timeout(900)
@pytest.mark.parametrize("Hq,Hkv,ps", [(16, 8, 1), (32, 8, 1), (4, 1, 1), (8, 2, 16)])
def test_prefill_matches_flashinfer(pkg, Hq, Hkv, ps):
    fi = _flashinfer()
    q_lens, prefix = [8, 1, 93, 64, 300, 700], [32, 32, 7, 0, 0, 1100]
    case = make_paged_case([a + b for a, b in zip(q_lens, prefix)], Hq, Hkv, q_lens=q_lens, page_size=ps, seed=6)
    d = dev(case)
    try:
        ws = torch.empty(128 * 1024 * 1024, dtype=torch.uint8, device="cuda")
        w = fi.BatchPrefillWithPagedKVCacheWrapper(ws, "NHD", backend="auto")
        w.plan(qo_indptr=d["qo_indptr"], paged_kv_indices=d["kv_indices"], paged_kv_indptr=d["kv_indptr"],
               paged_kv_last_page_len=d["last_page_len"], num_qo_heads=Hq, num_kv_heads=Hkv, head_dim_qk=128,
               head_dim_vo=128, page_size=ps, causal=True, q_data_type=torch.bfloat16, kv_data_type=torch.bfloat16)
        w._sm_scale = 128 ** -0.5
        ref = w.run(d["q"], (d["k_cache"].view(-1, ps, Hkv, 128), d["v_cache"].view(-1, ps, Hkv, 128)))
    except Exception as e:
        pytest.fail(f"flashinfer prefill could not run on this CUDA box, the FP oracle is unpinned: {type(e).__name__}: {e}")
    mine = pkg.ops.prefill(d["q"], d["k_cache"], d["v_cache"], d["qo_indptr"], d["kv_indptr"], d["kv_indices"],
                           d["last_page_len"], torch.empty(pkg.ops.prefill_workspace_bytes(case["B"], sum(q_lens), Hq, Hkv, 128),
                                                               dtype=torch.uint8, device="cuda"), ps, 128 ** -0.5, True)
    _close(mine, ref, "prefill")
