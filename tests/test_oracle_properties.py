"""CPU: size-independent properties of the attention oracle (numpy/torch and plain C).

The GPU parity tests use these same properties at BASELINE.json's full sizes, where the oracle itself is too slow
(tests/test_gpu_attention.py::test_prefill_properties_full_bench_step); here the oracle is held to them at sizes
it finishes in seconds, so a property failing on the GPU cannot be the oracle's own defect.
Semantics under test: layers/attention.py:355-369 (bottom-right causal prefill over prefix + new tokens, unmasked
decode), GQA head mapping h -> h // (Hq // Hkv), scale head_dim ** -0.5 (:324).
"""
import numpy as np
import pytest
import torch

from oracle import c_oracle
from oracle import paged_attention_oracle as orc
from tests.util import make_paged_case


def _py(case, **kw):
    return orc.paged_attention(case["q"], case["k_cache"], case["v_cache"], case["qo_indptr"], case["kv_indptr"],
                               case["kv_indices"], case["last_page_len"], case["page_size"], return_fp32=True, **kw)


def _c(case, **kw):
    return c_oracle.paged_attention(case["q"], case["k_cache"], case["v_cache"], case["qo_indptr"], case["kv_indptr"],
                                    case["kv_indices"], case["last_page_len"], case["page_size"], **kw)


IMPLS = [pytest.param(_py, id="python"), pytest.param(_c, id="c")]


@pytest.mark.parametrize("attn", IMPLS)
@pytest.mark.parametrize("ps", [1, 16])
def test_last_prefill_row_equals_decode(attn, ps):
    """The last new token of a prefill sees every KV position: its row is decode attention of that query."""
    kv_lens, q_lens = [37, 130, 5, 64], [37, 20, 1, 64]
    c = make_paged_case(kv_lens, 8, 2, D=64, page_size=ps, q_lens=q_lens, seed=3)
    o = attn(c, causal=True)
    last = c["qo_indptr"][1:] - 1
    d = dict(c, q=c["q"][last], qo_indptr=np.arange(len(kv_lens) + 1, dtype=np.int32))
    od = attn(d, causal=False)
    torch.testing.assert_close(o[last], od, atol=1e-5, rtol=1e-5)


@pytest.mark.parametrize("attn", IMPLS)
def test_constant_v_and_linearity_in_v(attn):
    """softmax weights sum to one => constant V gives that constant; the output is linear in V."""
    c = make_paged_case([50, 9, 140], 4, 2, D=64, q_lens=[50, 3, 17], seed=5)
    const = dict(c, v_cache=torch.full_like(c["v_cache"], -0.5))
    o = attn(const, causal=True)
    torch.testing.assert_close(o, torch.full_like(o, -0.5), atol=1e-5, rtol=0)
    g = torch.Generator().manual_seed(1)
    v2 = torch.randn(c["v_cache"].shape, generator=g).to(torch.bfloat16)
    vs = (c["v_cache"].float() + v2.float())
    # keep the sum exactly representable so the check is about attention, not about rounding V
    vs_b = vs.to(torch.bfloat16)
    v1_exact = (vs_b.float() - v2.float()).to(torch.bfloat16)
    keep = (v1_exact.float() + v2.float()) == vs_b.float()
    v1 = torch.where(keep, v1_exact, torch.zeros_like(v1_exact))
    v2 = torch.where(keep, v2, vs_b)
    o1, o2, o12 = (attn(dict(c, v_cache=v), causal=True) for v in (v1, v2, vs_b))
    torch.testing.assert_close(o12, o1 + o2, atol=2e-5, rtol=1e-5)


@pytest.mark.parametrize("attn", IMPLS)
@pytest.mark.parametrize("ps", [1, 16])
def test_prefix_split_invariance(attn, ps):
    """Rows of a full prefill == rows of a prefill of the suffix alone over (cached prefix + suffix) KV: what
    admission-time prefix reuse (RadixTree.match_prefix) relies on."""
    kv_lens = [96, 33, 200]
    full = make_paged_case(kv_lens, 8, 4, D=64, page_size=ps, q_lens=kv_lens, seed=7)
    o_full = attn(full, causal=True)
    cached = [64, 0, 199]
    q_lens = [L - p for L, p in zip(kv_lens, cached)]
    rows = np.concatenate([np.arange(full["qo_indptr"][i] + cached[i], full["qo_indptr"][i + 1]) for i in range(3)])
    qo = np.zeros(4, np.int32)
    qo[1:] = np.cumsum(q_lens)
    part = dict(full, q=full["q"][rows], qo_indptr=qo)
    o_part = attn(part, causal=True)
    torch.testing.assert_close(o_part, o_full[rows], atol=1e-5, rtol=1e-5)


@pytest.mark.parametrize("attn", IMPLS)
def test_empty_and_degenerate_sequences(attn):
    """Sequences with no new tokens are skipped, rows with no visible key are zeros, an empty batch is a no-op."""
    c = make_paged_case([12, 0, 7], 4, 2, D=64, q_lens=[0, 2, 7], seed=9)
    o = attn(c, causal=True)
    assert o.shape == (9, 4, 64)
    assert torch.count_nonzero(o[0:2]) == 0            # sequence 1: queries but Lkv == 0
    assert torch.isfinite(o).all() and torch.count_nonzero(o[2:]) > 0
    e = make_paged_case([], 4, 2, D=64, q_lens=[], seed=9)
    assert attn(e, causal=True).shape == (0, 4, 64)


@pytest.mark.parametrize("attn", IMPLS)
def test_gqa_head_mapping(attn):
    """Query head h reads KV head h // group: zeroing one KV head's V zeroes exactly its group of query heads."""
    c = make_paged_case([40, 11], 8, 4, D=64, q_lens=[40, 1], seed=2)
    v = c["v_cache"].clone()
    v[:, 2] = 0
    o = attn(dict(c, v_cache=v), causal=True)
    assert torch.count_nonzero(o[:, 4:6]) == 0
    assert all(torch.count_nonzero(o[:, h]) > 0 for h in (0, 1, 2, 3, 6, 7))


@pytest.mark.parametrize("Hq,Hkv,D", [(4, 4, 64), (16, 2, 64), (6, 3, 36), (8, 1, 128)])
def test_c_port_matches_python_oracle_across_group_sizes(Hq, Hkv, D):
    """The C port walks K/V once per (sequence, KV head) for the whole group of query heads, 8 partial sums per
    dot product plus a scalar tail when D % 8 != 0: group 1 / 2 / 8 and a ragged D against the torch restatement."""
    for ps, causal in [(1, True), (16, False)]:
        q_lens = [7, 1, 33] if causal else [1, 1, 1]
        c = make_paged_case([19, 64, 33], Hq, Hkv, D=D, page_size=ps, q_lens=q_lens, seed=13)
        torch.testing.assert_close(_c(c, causal=causal), _py(c, causal=causal), atol=2e-5, rtol=1e-4)
