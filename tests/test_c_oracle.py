"""CPU: the plain-C oracle agrees with the numpy/torch oracle and the reference goldens."""
import numpy as np
import torch

from oracle import c_oracle
from oracle import paged_attention_oracle as orc
from tests.util import make_paged_case


def test_c_metadata_matches_goldens(golden):
    for name in golden["eager_case_names"]:
        name = str(name)
        qo, kvp, idx, last, outs = c_oracle.build_metadata(golden[f"{name}.in_num_token_ids"], golden[f"{name}.in_cached_kv_len"],
                                                           golden[f"{name}.in_kv_lens"], golden[f"{name}.in_kv_flat"])
        if f"{name}.qo_indptr" in golden:
            np.testing.assert_array_equal(qo, golden[f"{name}.qo_indptr"])
        np.testing.assert_array_equal(kvp, golden[f"{name}.paged_kv_indptr"])
        np.testing.assert_array_equal(idx, golden[f"{name}.paged_kv_indices"])
        np.testing.assert_array_equal(last, golden[f"{name}.paged_kv_last_page_len"])
        np.testing.assert_array_equal(outs, golden[f"{name}.output_kv_indices"])


def test_c_attention_matches_python_oracle():
    for ps, q_lens, kv_lens in [(1, [5, 1, 40], [5, 9, 140]), (16, [1, 1], [33, 200]), (1, [3], [3])]:
        case = make_paged_case(kv_lens, 8, 2, D=64, page_size=ps, q_lens=q_lens, seed=9)
        a = orc.paged_attention(case["q"], case["k_cache"], case["v_cache"], case["qo_indptr"], case["kv_indptr"],
                                case["kv_indices"], case["last_page_len"], ps, causal=True, return_fp32=True)
        b = c_oracle.paged_attention(case["q"], case["k_cache"], case["v_cache"], case["qo_indptr"], case["kv_indptr"],
                                     case["kv_indices"], case["last_page_len"], ps, causal=True)
        torch.testing.assert_close(a, b, atol=2e-5, rtol=1e-4)


def test_c_graph_replay_matches_goldens(golden):
    import ctypes
    lib = c_oracle.load()
    max_bs, ctx = (int(x) for x in golden["graph.max_bs_context_len"])
    buf = orc.prepare_graph_buffers(max_bs, ctx)
    for tag in golden["graph.replay_names"]:
        tag = str(tag)
        lens = np.ascontiguousarray(golden[f"graph.{tag}.in_kv_lens"], np.int64)
        cached = np.ascontiguousarray(golden[f"graph.{tag}.in_cached_kv_len"], np.int64)
        flat = np.ascontiguousarray(golden[f"graph.{tag}.in_kv_flat"], np.int64)
        padded = int(golden[f"graph.{tag}.padded_bs"][0])
        P = lambda a: ctypes.c_void_p(a.ctypes.data)
        lib.oracle_replay_graph_buffers(P(cached), P(lens), P(flat), len(lens), padded, P(buf.paged_kv_indptr_buffer),
                                        P(buf.paged_kv_indices_buffer), ctypes.c_int64(buf.paged_kv_indices_buffer.size),
                                        P(buf.output_kv_indices_buffer), ctypes.c_int64(buf.output_kv_indices_buffer.size))
        for n in ("paged_kv_indptr_buffer", "paged_kv_indices_buffer", "output_kv_indices_buffer"):
            np.testing.assert_array_equal(getattr(buf, n), golden[f"graph.{tag}.{n}"], err_msg=f"{tag}.{n}")
