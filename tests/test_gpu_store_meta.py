"""GPU: slot write and metadata build through the C ABI — bit-exact vs oracle and reference goldens."""
import numpy as np
import pytest
import torch

from oracle import paged_attention_oracle as orc
from tests.test_oracle_golden import _seqs

pytestmark = pytest.mark.gpu


def _bits(t):
    return t.detach().cpu().view(torch.int16)


@pytest.mark.parametrize("T,Hkv,D,pool", [(11, 2, 64, 40), (1, 8, 128, 16), (300, 8, 128, 1000), (77, 1, 128, 90),
                                          (5, 3, 40, 9), (4097, 4, 128, 5000)])
def test_store_kv_bit_exact(pkg, T, Hkv, D, pool):
    g = torch.Generator().manual_seed(T)
    Hq = 2 * Hkv
    qkv = torch.randn((T, (Hq + 2 * Hkv) * D), generator=g).to(torch.bfloat16)
    k = qkv[:, Hq * D:(Hq + Hkv) * D].contiguous().view(T, Hkv, D)
    v = qkv[:, (Hq + Hkv) * D:].view(T, Hkv, D)                       # strided view (models/qwen3.py:123)
    slots = torch.from_numpy(np.random.default_rng(T).permutation(pool)[:T].astype(np.int64))
    slots[::5] = -1                                                   # padding rows must be skipped
    init = torch.randn((pool, Hkv, D), generator=g).to(torch.bfloat16)
    kc_ref, vc_ref = init.clone(), init.clone()
    orc.store_kvcache(k, v, kc_ref, vc_ref, slots)
    qkv_d = qkv.cuda()
    kd = qkv_d[:, Hq * D:(Hq + Hkv) * D].contiguous().view(T, Hkv, D)
    vd = qkv_d[:, (Hq + Hkv) * D:].view(T, Hkv, D)
    kc, vc = init.clone().cuda(), init.clone().cuda()
    pkg.store_kvcache(kd, vd, kc, vc, slots.cuda())
    assert torch.equal(_bits(kc), _bits(kc_ref)) and torch.equal(_bits(vc), _bits(vc_ref))


def test_store_kv_reference_golden_and_pool(pkg, golden):
    seed, T, Hkv, D, pool = (int(x) for x in golden["store.seed_T_Hkv_D_pool"])
    g = torch.Generator().manual_seed(seed)
    qkv = torch.randn((T, (4 + 2 * Hkv) * D), generator=g).to(torch.bfloat16).cuda()
    k = qkv[:, 4 * D:(4 + Hkv) * D].contiguous().view(T, Hkv, D)
    v = qkv[:, (4 + Hkv) * D:].view(T, Hkv, D)
    kv = pkg.KVCachePool(torch.bfloat16, torch.device("cuda"), pool, 3, Hkv, D)
    kv.k_cache.zero_(); kv.v_cache.zero_()
    kv.set_kv_cache(1, torch.from_numpy(golden["store.slots"]).cuda(), k, v)
    kc, vc = kv.get_kv_cache(1)
    np.testing.assert_array_equal(_bits(kc).numpy(), golden["store.k_cache_bits"])
    np.testing.assert_array_equal(_bits(vc).numpy(), golden["store.v_cache_bits"])
    assert kv.k_cache[0].abs().sum() == 0 and kv.k_cache[2].abs().sum() == 0   # other layers untouched
    # empty step is a no-op
    pkg.store_kvcache(k[:0], v[:0], kc, vc, torch.zeros(0, dtype=torch.int64, device="cuda"))


class S:
    def __init__(self, sid, v):
        self.seq_id, self.token_ids, self.kv_indices, self.cached_kv_len = sid, [0] * v.num_token_ids, list(v.kv_indices), v.cached_kv_len


class Batch:
    def __init__(self, mode, seqs):
        self.forward_mode, self.seqs, self.num_seqs = mode, seqs, len(seqs)


def _backend(pkg, page_size=1, Hq=16, Hkv=8):
    pool = pkg.KVCachePool(torch.bfloat16, torch.device("cuda"), 64, 1, Hkv, 128)
    return pkg.AttentionBackend(pool, Hq, Hkv, 128, torch.bfloat16, torch.device("cuda"), page_size=page_size)


def test_build_metadata_matches_reference_goldens(pkg, golden):
    be = _backend(pkg)
    for name in golden["eager_case_names"]:
        name = str(name)
        views = _seqs(golden, name)
        mode = pkg.ForwardMode.PREFILL if name.startswith("prefill") else pkg.ForwardMode.DECODE
        md = be.build_metadata(Batch(mode, [S(f"{name}{i}", v) for i, v in enumerate(views)]))
        w = md.wrapper
        if f"{name}.qo_indptr" in golden:
            np.testing.assert_array_equal(w.qo_indptr.cpu().numpy(), golden[f"{name}.qo_indptr"])
        np.testing.assert_array_equal(w.paged_kv_indptr.cpu().numpy(), golden[f"{name}.paged_kv_indptr"])
        np.testing.assert_array_equal(w.paged_kv_indices.cpu().numpy(), golden[f"{name}.paged_kv_indices"])
        np.testing.assert_array_equal(w.paged_kv_last_page_len.cpu().numpy(), golden[f"{name}.paged_kv_last_page_len"])
        np.testing.assert_array_equal(md.output_kv_indices.cpu().numpy(), golden[f"{name}.output_kv_indices"])
        assert md.output_kv_indices.dtype == torch.int64 and w.paged_kv_indices.dtype == torch.int32


def test_graph_replay_buffers_match_reference_goldens(pkg, golden):
    be = _backend(pkg)
    graph = pkg.CUDAGraph()
    max_bs, ctx = (int(x) for x in golden["graph.max_bs_context_len"])
    be.prepare_for_cuda_graph_capture(graph, max_bs, ctx)
    names = ("paged_kv_indptr_buffer", "paged_kv_indices_buffer", "paged_kv_last_page_len_buffer",
             "output_kv_indices_buffer")
    for n in names:
        got = graph.get_input_buffer(n)
        np.testing.assert_array_equal(got.cpu().numpy(), golden[f"graph.prepare.{n}"])
        assert got.cpu().numpy().dtype == golden[f"graph.prepare.{n}"].dtype
    for bs in (8, 4, 2, 1):
        md = be.build_metadata_for_cuda_graph_capture(graph, bs, ctx)
        np.testing.assert_array_equal(md.wrapper.paged_kv_indptr.cpu().numpy(), golden[f"graph.capture{bs}.indptr"])
        assert md.wrapper.paged_kv_indices.numel() == int(golden[f"graph.capture{bs}.indices_len"][0])
        assert md.output_kv_indices.numel() == int(golden[f"graph.capture{bs}.out_len"][0])
    for tag in golden["graph.replay_names"]:
        tag = str(tag)
        views = _seqs(golden, f"graph.{tag}")
        padded = int(golden[f"graph.{tag}.padded_bs"][0])
        md = be.build_metadata_for_cuda_graph_replay(
            graph, Batch(pkg.ForwardMode.DECODE, [S(f"{tag}{i}", v) for i, v in enumerate(views)]), padded)
        for n in names:   # whole buffers, including the stale-data-wiping tails
            np.testing.assert_array_equal(graph.get_input_buffer(n).cpu().numpy(), golden[f"graph.{tag}.{n}"],
                                          err_msg=f"{tag}.{n}")
        np.testing.assert_array_equal(md.output_kv_indices.cpu().numpy(), golden[f"graph.{tag}.md_output_kv_indices"])


@pytest.mark.parametrize("page_size", [1, 16])
def test_build_metadata_random_vs_oracle_incremental(pkg, page_size):
    """Prefill then several decode steps of a changing batch: device arrays == oracle every step."""
    be = _backend(pkg, page_size)
    rng = np.random.default_rng(5)
    next_page = [0]

    def alloc(n_tokens_have, n_new, kv):
        for _ in range(n_new):
            if page_size == 1:
                kv.append(int(rng.integers(0, 1 << 20)))
            else:
                if len(kv) % page_size == 0:
                    next_page[0] += int(rng.integers(1, 5))
                    kv.append(next_page[0] * page_size)
                else:
                    kv.append(kv[-1] + 1)

    seqs = []
    for i, n in enumerate([1, 15, 16, 17, 333, 64, 1000, 2]):
        s = S(f"q{i}", orc.SeqView(n, [], 0))
        alloc(0, n, s.kv_indices)
        seqs.append(s)

    def check(batch, mode):
        md = be.build_metadata(Batch(mode, batch))
        ref = orc.build_metadata([orc.SeqView(len(s.token_ids), s.kv_indices, s.cached_kv_len) for s in batch], page_size)
        w = md.wrapper
        np.testing.assert_array_equal(w.qo_indptr.cpu().numpy(), ref.qo_indptr)
        np.testing.assert_array_equal(w.paged_kv_indptr.cpu().numpy(), ref.paged_kv_indptr)
        np.testing.assert_array_equal(w.paged_kv_indices.cpu().numpy(), ref.paged_kv_indices)
        np.testing.assert_array_equal(w.paged_kv_last_page_len.cpu().numpy(), ref.paged_kv_last_page_len)
        np.testing.assert_array_equal(md.output_kv_indices.cpu().numpy(), ref.output_kv_indices)

    check(seqs, pkg.ForwardMode.PREFILL)
    for step in range(20):
        for s in seqs:
            s.cached_kv_len = len(s.kv_indices)
            s.token_ids.append(0)
            alloc(0, 1, s.kv_indices)
        batch = seqs if step % 3 else seqs[::2]      # batch membership changes between steps
        check(batch, pkg.ForwardMode.DECODE)
    check([], pkg.ForwardMode.DECODE)                 # empty batch
    big = [S(f"b{i}", orc.SeqView(3, [int(x) for x in rng.integers(0, 1 << 20, 3 if page_size == 1 else 0)], 0))
           for i in range(1500)]                      # B > one scan block (1024)
    if page_size == 1:
        check(big, pkg.ForwardMode.PREFILL)
