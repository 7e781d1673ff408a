"""CPU: the oracle restatement against fixtures produced by RUNNING the reference's own code
(oracle/gen_golden.py).  Bit-exact on every integer array."""
import numpy as np
import torch

from oracle import paged_attention_oracle as orc


def _seqs(golden, prefix):
    ntok = golden[f"{prefix}.in_num_token_ids"] if f"{prefix}.in_num_token_ids" in golden else None
    cached = golden[f"{prefix}.in_cached_kv_len"]
    lens = golden[f"{prefix}.in_kv_lens"]
    flat = golden[f"{prefix}.in_kv_flat"]
    out, o = [], 0
    for i, n in enumerate(lens):
        kv = [int(x) for x in flat[o:o + n]]
        o += n
        out.append(orc.SeqView(num_token_ids=int(ntok[i]) if ntok is not None else int(n),
                               kv_indices=kv, cached_kv_len=int(cached[i])))
    return out


def test_eager_metadata_matches_reference(golden):
    for name in golden["eager_case_names"]:
        name = str(name)
        md = orc.build_metadata(_seqs(golden, name), page_size=1)
        if f"{name}.qo_indptr" in golden:
            np.testing.assert_array_equal(md.qo_indptr, golden[f"{name}.qo_indptr"])
            assert md.qo_indptr.dtype == golden[f"{name}.qo_indptr"].dtype == np.int32
        np.testing.assert_array_equal(md.paged_kv_indptr, golden[f"{name}.paged_kv_indptr"])
        np.testing.assert_array_equal(md.paged_kv_indices, golden[f"{name}.paged_kv_indices"])
        np.testing.assert_array_equal(md.paged_kv_last_page_len, golden[f"{name}.paged_kv_last_page_len"])
        np.testing.assert_array_equal(md.output_kv_indices, golden[f"{name}.output_kv_indices"])
        assert md.paged_kv_indices.dtype == golden[f"{name}.paged_kv_indices"].dtype == np.int32
        assert md.output_kv_indices.dtype == golden[f"{name}.output_kv_indices"].dtype == np.int64


def test_graph_buffers_match_reference(golden):
    max_bs, ctx = (int(x) for x in golden["graph.max_bs_context_len"])
    buf = orc.prepare_graph_buffers(max_bs, ctx)
    for n in ("paged_kv_indptr_buffer", "paged_kv_indices_buffer", "paged_kv_last_page_len_buffer",
              "output_kv_indices_buffer"):
        got = getattr(buf, n)
        np.testing.assert_array_equal(got, golden[f"graph.prepare.{n}"])
        assert got.dtype == golden[f"graph.prepare.{n}"].dtype
    for bs in (8, 4, 2, 1):
        np.testing.assert_array_equal(buf.paged_kv_indptr_buffer[:bs + 1], golden[f"graph.capture{bs}.indptr"])
        assert int(golden[f"graph.capture{bs}.indices_len"][0]) == bs * ctx
    # replays are applied in order on the SAME buffers, exactly as the reference run did
    for tag in golden["graph.replay_names"]:
        tag = str(tag)
        seqs = _seqs(golden, f"graph.{tag}")
        padded = int(golden[f"graph.{tag}.padded_bs"][0])
        orc.replay_graph_buffers(buf, seqs, padded)
        for n in ("paged_kv_indptr_buffer", "paged_kv_indices_buffer", "paged_kv_last_page_len_buffer",
                  "output_kv_indices_buffer"):
            np.testing.assert_array_equal(getattr(buf, n), golden[f"graph.{tag}.{n}"], err_msg=f"{tag}.{n}")
        np.testing.assert_array_equal(buf.output_kv_indices_buffer[:padded], golden[f"graph.{tag}.md_output_kv_indices"])
        np.testing.assert_array_equal(buf.paged_kv_indptr_buffer[:padded + 1], golden[f"graph.{tag}.plan_indptr"])


def test_store_kvcache_matches_reference_triton_kernel(golden):
    seed, T, Hkv, D, pool = (int(x) for x in golden["store.seed_T_Hkv_D_pool"])
    g = torch.Generator().manual_seed(seed)
    qkv = torch.randn((T, (4 + 2 * Hkv) * D), generator=g).to(torch.bfloat16)
    k = qkv[:, 4 * D:(4 + Hkv) * D].contiguous().view(T, Hkv, D)
    v = qkv[:, (4 + Hkv) * D:].view(T, Hkv, D)
    k_cache = torch.zeros((pool, Hkv, D), dtype=torch.bfloat16)
    v_cache = torch.zeros((pool, Hkv, D), dtype=torch.bfloat16)
    orc.store_kvcache(k, v, k_cache, v_cache, torch.from_numpy(golden["store.slots"]))
    np.testing.assert_array_equal(k_cache.view(torch.int16).numpy(), golden["store.k_cache_bits"])
    np.testing.assert_array_equal(v_cache.view(torch.int16).numpy(), golden["store.v_cache_bits"])


def test_known_answers(golden):
    # README.md:99-100 (133966 "output tokens") and kv_cache.py:108-114, :346-352
    prompts, max_tokens = orc.bench_workload()
    assert [len(prompts), sum(map(len, prompts)), sum(max_tokens)] == [256, 142827, 133966]
    np.testing.assert_array_equal(golden["bench.num_prompts_prompt_tokens_sum_max_tokens"], [256, 142827, 133966])
    np.testing.assert_array_equal(golden["alloc_fresh_first8"], np.arange(8))
    np.testing.assert_array_equal(golden["radix_dedup_match"], [10, 11, 12, 23])


def test_page_size_16_definition():
    # page-aligned allocation: 3 pages of 16 slots, last page holds 5 tokens
    pages = [7, 2, 9]
    kv = [p * 16 + j for p in pages for j in range(16)][:37]
    md = orc.build_metadata([orc.SeqView(37, kv, 36)], page_size=16)
    np.testing.assert_array_equal(md.paged_kv_indices, pages)
    np.testing.assert_array_equal(md.paged_kv_indptr, [0, 3])
    np.testing.assert_array_equal(md.paged_kv_last_page_len, [5])
    np.testing.assert_array_equal(md.output_kv_indices, [kv[36]])


def test_attention_oracle_against_dense_sdpa():
    """The paged fp32 oracle equals torch SDPA on the gathered, GQA-expanded tensors."""
    import torch.nn.functional as F
    from tests.util import make_paged_case, token_slots
    case = make_paged_case([33, 1, 80], Hq=4, Hkv=2, D=64, q_lens=[5, 1, 80], seed=3)
    out = orc.prefill_attention(case["q"], case["k_cache"], case["v_cache"], case["qo_indptr"], case["kv_indptr"],
                                case["kv_indices"], case["last_page_len"], return_fp32=True)
    for i in range(3):
        sl = torch.from_numpy(token_slots(case, i))
        K = case["k_cache"][sl].float().repeat_interleave(2, 1).transpose(0, 1)
        V = case["v_cache"][sl].float().repeat_interleave(2, 1).transpose(0, 1)
        q0, q1 = case["qo_indptr"][i], case["qo_indptr"][i + 1]
        Q = case["q"][q0:q1].float().transpose(0, 1)
        lq, lkv = Q.shape[1], K.shape[1]
        mask = torch.ones(lq, lkv, dtype=torch.bool).tril(diagonal=lkv - lq)
        ref = F.scaled_dot_product_attention(Q, K, V, attn_mask=mask).transpose(0, 1)
        torch.testing.assert_close(out[q0:q1], ref, atol=2e-5, rtol=1e-4)
