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


def _fi_prefill(fi, d, Hq, Hkv, ps):
    """FlashInfer planned and run as the reference does for a PREFILL batch (layers/attention.py:49-53, 106-119)."""
    try:
        ws = torch.empty(256 * 1024 * 1024, dtype=torch.uint8, device="cuda")
        w = fi.BatchPrefillWithPagedKVCacheWrapper(ws, "NHD", backend="auto")
        w.plan(qo_indptr=d["qo_indptr"], paged_kv_indices=d["kv_indices"], paged_kv_indptr=d["kv_indptr"],
               paged_kv_last_page_len=d["last_page_len"], num_qo_heads=Hq, num_kv_heads=Hkv, head_dim_qk=128,
               head_dim_vo=128, page_size=ps, causal=True, q_data_type=torch.bfloat16, kv_data_type=torch.bfloat16)
        w._sm_scale = 128 ** -0.5
        return w.run(d["q"], (d["k_cache"].view(-1, ps, Hkv, 128), d["v_cache"].view(-1, ps, Hkv, 128)))
    except Exception as e:
        pytest.fail(f"flashinfer prefill could not run on this CUDA box, the FP oracle is unpinned: {type(e).__name__}: {e}")


def _every_row_close(mine, ref, what):
    """Per-row version of the suite's tolerances: no row may hide behind the global norm."""
    d = mine.float() - ref.float()
    T = d.shape[0]
    worst = d.abs().reshape(T, -1).max(dim=1).values
    assert worst.max().item() <= ATOL, f"{what}: row {int(worst.argmax())} max abs diff {worst.max().item():.4g}"
    rel = d.reshape(T, -1).norm(dim=1) / ref.float().reshape(T, -1).norm(dim=1).clamp_min(1e-6)
    # a single row has far fewer terms than the whole tensor: RTOL on the global norm, 3x RTOL on any one row
    assert rel.max().item() <= 3 * RTOL, f"{what}: row {int(rel.argmax())} rel diff {rel.max().item():.4g}"
    _close(mine, ref, what)


@pytest.mark.timeout(1800)
@pytest.mark.parametrize("Hq,Hkv", [(32, 8), (4, 1)])
def test_config4_shared_prefix_every_row_matches_flashinfer(pkg, Hq, Hkv):
    """BASELINE.json configs[3] / SURVEY.md 8(d) config 4 at the kernel: 256 sequences admitted with
    cached_kv_len = 2048 whose first 2048 KV slots are THE SAME 2048 slots (one radix-tree path, core/kv_cache.py:
    159-161 match_prefix; core/scheduler.py:67-74 admission) followed by a private suffix of randint(16, 256) new
    tokens, page size 1, full-head and one-KV-head-per-rank (TP8) shapes.  Every output row is compared with FlashInfer
    planned as the reference plans a prefill (layers/attention.py:49-53, 106-119); sampled rows with fp32 on the device."""
    import numpy as np
    fi = _flashinfer()
    rng = np.random.default_rng(4)
    B, P, D = 256, 2048, 128
    suffix = rng.integers(16, 257, size=B)
    T = int(suffix.sum())
    n_slots = P + T + 4096
    perm = rng.permutation(n_slots).astype(np.int32)
    shared, private = perm[:P], perm[P:P + T]
    qo = np.zeros(B + 1, np.int32)
    qo[1:] = np.cumsum(suffix)
    kv_indptr = np.zeros(B + 1, np.int32)
    kv_indptr[1:] = np.cumsum(P + suffix)
    kv_indices = np.concatenate([np.concatenate([shared, private[qo[i]:qo[i + 1]]]) for i in range(B)]).astype(np.int32)
    g = torch.Generator(device="cuda").manual_seed(44)
    rnd = lambda *shape: torch.randn(shape, device="cuda", dtype=torch.float32, generator=g).to(torch.bfloat16)
    d = dict(q=rnd(T, Hq, D), k_cache=rnd(n_slots, Hkv, D), v_cache=rnd(n_slots, Hkv, D),
             qo_indptr=torch.from_numpy(qo).cuda(), kv_indptr=torch.from_numpy(kv_indptr).cuda(),
             kv_indices=torch.from_numpy(kv_indices).cuda(), last_page_len=torch.ones(B, dtype=torch.int32, device="cuda"))
    ref = _fi_prefill(fi, d, Hq, Hkv, 1)
    ws = torch.empty(pkg.ops.prefill_workspace_bytes(B, T, Hq, Hkv, D), dtype=torch.uint8, device="cuda")
    mine = pkg.ops.prefill(d["q"], d["k_cache"], d["v_cache"], d["qo_indptr"], d["kv_indptr"], d["kv_indices"],
                           d["last_page_len"], ws, 1, D ** -0.5, True)
    assert mine.shape == ref.shape == (T, Hq, D) and torch.isfinite(mine.float()).all()
    _every_row_close(mine, ref, f"config 4 prefill Hq{Hq} Hkv{Hkv}")
    # sampled rows against a direct fp32 computation (first, middle and last new token of four sequences)
    for s in (0, 77, 200, 255):
        n = int(suffix[s])
        for j in (0, n // 2, n - 1):
            h = (s + j) % Hq
            slots = d["kv_indices"][int(kv_indptr[s]):int(kv_indptr[s]) + P + j + 1].long()
            K, V = d["k_cache"][slots, h // (Hq // Hkv)].float(), d["v_cache"][slots, h // (Hq // Hkv)].float()
            p = torch.softmax((K @ d["q"][int(qo[s]) + j, h].float()) * D ** -0.5, 0)
            assert (mine[int(qo[s]) + j, h].float() - p @ V).abs().max().item() <= ATOL


@pytest.mark.timeout(1800)
def test_full_bench_prefill_step_every_row_matches_flashinfer(pkg):
    """The whole prefill step of the bench workload (BASELINE.json configs[2]: 256 prompts, 142 827 tokens, Hq32/Hkv8,
    page 1, fragmented slots): every output row against FlashInfer.  tests/test_gpu_attention.py checks the same step
    through size-independent properties plus three spot rows; this closes the gap between 'properties' and 'every row'."""
    from oracle.paged_attention_oracle import bench_workload
    fi = _flashinfer()
    plens = [len(p) for p in bench_workload()[0]]
    Hq, Hkv, D = 32, 8, 128
    d = dev(make_paged_case(plens, Hq, Hkv, q_lens=plens, seed=11))
    ref = _fi_prefill(fi, d, Hq, Hkv, 1)
    ws = torch.empty(pkg.ops.prefill_workspace_bytes(len(plens), sum(plens), Hq, Hkv, D), dtype=torch.uint8, device="cuda")
    mine = pkg.ops.prefill(d["q"], d["k_cache"], d["v_cache"], d["qo_indptr"], d["kv_indptr"], d["kv_indices"],
                           d["last_page_len"], ws, 1, D ** -0.5, True)
    _every_row_close(mine, ref, "bench prefill step")
