"""Generate tests/golden/index_golden.npz by RUNNING THE REFERENCE'S OWN CODE on CPU.

Run in the authoring container only (needs /root/reference, which does not exist on the
GPU box):

    TRITON_INTERPRET=1 python oracle/gen_golden.py

What runs unmodified from /root/reference/nanovllmx:
  * layers/attention.py  AttentionBackend.build_metadata, prepare_for_cuda_graph_capture,
    build_metadata_for_cuda_graph_capture, build_metadata_for_cuda_graph_replay
    (device="cpu").  The two FlashInfer wrapper classes it instantiates are replaced by a
    recording stub whose ``plan`` stores the arrays the reference hands to FlashInfer, so the
    captured qo_indptr / paged_kv_indptr / paged_kv_indices / last_page_len are exactly what
    the reference computes.
  * core/kv_cache.py     store_kvcache (the Triton kernel, under TRITON_INTERPRET=1) and
    KVCacheAllocator / RadixTree / KVCacheManager (to produce realistic fragmented slot ids
    and to record the worked de-dup example of kv_cache.py:346-352).

Attention outputs are NOT generated here: FlashInfer is CUDA-only (see the oracle header).
"""
import os
import sys

os.environ.setdefault("TRITON_INTERPRET", "1")
sys.path.insert(0, "/root/reference")
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))

import numpy as np
import torch


class _RecordingWrapper:
    """Stands in for flashinfer.Batch{Prefill,Decode}WithPagedKVCacheWrapper."""

    def __init__(self, *args, **kwargs):
        self.init_kwargs = kwargs
        self.plans = []

    def plan(self, **kw):
        rec = {}
        for name, val in kw.items():
            rec[name] = val.clone() if isinstance(val, torch.Tensor) else val
        self.plans.append(rec)


def _arr(t):
    return t.detach().cpu().numpy().copy()


def main():
    import nanovllmx.layers.attention as ref_attn
    import nanovllmx.core.kv_cache as ref_kv
    from nanovllmx.core.common import Sequence, ForwardBatch, ForwardMode
    from nanovllmx.core.cuda_graph import CUDAGraph

    ref_attn.BatchPrefillWithPagedKVCacheWrapper = _RecordingWrapper
    ref_attn.BatchDecodeWithPagedKVCacheWrapper = _RecordingWrapper

    out = {}
    rng = np.random.default_rng(1234)
    cpu = torch.device("cpu")

    # ---- slot ids from the reference allocator + radix tree (fragmented on purpose) -------
    alloc_total = 4096
    mgr = ref_kv.KVCacheManager(alloc_total)
    fresh = mgr.alloc_slots(8)
    out["alloc_fresh_first8"] = np.asarray(fresh, dtype=np.int64)      # kv_cache.py:108-114
    mgr.free_slots(fresh)

    # worked example of kv_cache.py:346-352
    m2 = ref_kv.KVCacheManager(64)
    s1 = Sequence(seq_id="a", token_ids=[1, 2, 3, 4], kv_indices=[10, 11, 12, 13])
    s2 = Sequence(seq_id="b", token_ids=[1, 2, 3, 5], kv_indices=[20, 21, 22, 23])
    m2.cache_sequence(s1)
    m2.cache_sequence(s2)
    matched, _node = m2.radix_tree.match_prefix([1, 2, 3, 5])
    out["radix_dedup_match"] = np.asarray(matched, dtype=np.int64)

    backend = ref_attn.AttentionBackend.__new__(ref_attn.AttentionBackend)
    backend.kv_cache = None
    backend.num_qo_heads, backend.num_kv_heads, backend.head_dim = 16, 8, 128
    backend.dtype, backend.device = torch.bfloat16, cpu
    backend.prefill_wrapper = _RecordingWrapper()
    backend.decode_wrapper = _RecordingWrapper()
    backend.decode_wrappers_cuda_graph = {}
    backend.workspace_buffer = torch.empty(16, dtype=torch.uint8)

    def make_batch(case, mode):
        seqs = []
        for j, (ntok, ncached) in enumerate(case):
            slots = mgr.alloc_slots(ntok)
            # shuffle a little so ids are not monotone
            seqs.append(Sequence(seq_id=str(j), token_ids=list(rng.integers(0, 10000, ntok)),
                                 kv_indices=[int(s) for s in slots], cached_kv_len=ncached))
        return ForwardBatch(forward_mode=mode, num_seqs=len(seqs), seqs=seqs)

    # burn + free some slots so later allocations are fragmented (LIFO free list)
    burn = mgr.alloc_slots(300)
    mgr.free_slots(burn[::3])

    cases = {
        "prefill_ragged": ([(5, 0), (1, 0), (130, 0), (17, 0)], ForwardMode.PREFILL),
        "prefill_prefix": ([(40, 32), (33, 32), (100, 7), (64, 0)], ForwardMode.PREFILL),
        "decode_small": ([(9, 8), (1, 0), (257, 256)], ForwardMode.DECODE),
        "decode_single": ([(77, 76)], ForwardMode.DECODE),
    }
    names = []
    for name, (case, mode) in cases.items():
        batch = make_batch(case, mode)
        md = backend.build_metadata(batch)
        if mode == ForwardMode.PREFILL:
            plan = backend.prefill_wrapper.plans[-1]
            out[f"{name}.qo_indptr"] = _arr(plan["qo_indptr"])
            out[f"{name}.paged_kv_indptr"] = _arr(plan["paged_kv_indptr"])
            out[f"{name}.paged_kv_indices"] = _arr(plan["paged_kv_indices"])
            out[f"{name}.paged_kv_last_page_len"] = _arr(plan["paged_kv_last_page_len"])
        else:
            plan = backend.decode_wrapper.plans[-1]
            out[f"{name}.paged_kv_indptr"] = _arr(plan["indptr"])
            out[f"{name}.paged_kv_indices"] = _arr(plan["indices"])
            out[f"{name}.paged_kv_last_page_len"] = _arr(plan["last_page_len"])
        out[f"{name}.output_kv_indices"] = _arr(md.output_kv_indices)
        out[f"{name}.in_num_token_ids"] = np.asarray([len(s.token_ids) for s in batch.seqs], np.int64)
        out[f"{name}.in_cached_kv_len"] = np.asarray([s.cached_kv_len for s in batch.seqs], np.int64)
        out[f"{name}.in_kv_lens"] = np.asarray([len(s.kv_indices) for s in batch.seqs], np.int64)
        out[f"{name}.in_kv_flat"] = np.asarray([x for s in batch.seqs for x in s.kv_indices], np.int64)
        names.append(name)
        for s in batch.seqs:
            mgr.free_slots(s.kv_indices)
    out["eager_case_names"] = np.asarray(names)

    # ---- CUDA-graph buffers: prepare, capture slices, two replays (second with fewer/shorter
    #      sequences, so stale contents from the first must be wiped by the fill rules) --------
    graph = CUDAGraph()
    max_bs, context_len = 8, 64
    backend.prepare_for_cuda_graph_capture(graph, max_bs=max_bs, context_len=context_len)
    for n in ("paged_kv_indptr_buffer", "paged_kv_indices_buffer",
              "paged_kv_last_page_len_buffer", "output_kv_indices_buffer"):
        out[f"graph.prepare.{n}"] = _arr(graph.get_input_buffer(n))
    for bs in (8, 4, 2, 1):
        md = backend.build_metadata_for_cuda_graph_capture(graph, bs=bs, context_len=context_len)
        plan = backend.decode_wrappers_cuda_graph[bs].plans[-1]
        out[f"graph.capture{bs}.indptr"] = _arr(plan["indptr"])
        out[f"graph.capture{bs}.indices_len"] = np.asarray([plan["indices"].numel()])
        out[f"graph.capture{bs}.out_len"] = np.asarray([md.output_kv_indices.numel()])
    replays = [
        ("r1", [(30, 29), (64, 63), (7, 6), (12, 11), (50, 49)], 8),
        ("r2", [(3, 2), (41, 40), (9, 8)], 4),
        ("r3", [(18, 17)], 1),
    ]
    rnames = []
    for tag, case, padded in replays:
        batch = make_batch(case, ForwardMode.DECODE)
        md = backend.build_metadata_for_cuda_graph_replay(graph, batch, padded_bs=padded)
        for n in ("paged_kv_indptr_buffer", "paged_kv_indices_buffer",
                  "paged_kv_last_page_len_buffer", "output_kv_indices_buffer"):
            out[f"graph.{tag}.{n}"] = _arr(graph.get_input_buffer(n))
        out[f"graph.{tag}.md_output_kv_indices"] = _arr(md.output_kv_indices)
        plan = backend.decode_wrappers_cuda_graph[padded].plans[-1]
        out[f"graph.{tag}.plan_indptr"] = _arr(plan["indptr"])
        out[f"graph.{tag}.plan_indices"] = _arr(plan["indices"])
        out[f"graph.{tag}.plan_last_page_len"] = _arr(plan["last_page_len"])
        out[f"graph.{tag}.in_kv_lens"] = np.asarray([len(s.kv_indices) for s in batch.seqs], np.int64)
        out[f"graph.{tag}.in_cached_kv_len"] = np.asarray([s.cached_kv_len for s in batch.seqs], np.int64)
        out[f"graph.{tag}.in_kv_flat"] = np.asarray([x for s in batch.seqs for x in s.kv_indices], np.int64)
        out[f"graph.{tag}.padded_bs"] = np.asarray([padded])
        rnames.append(tag)
        for s in batch.seqs:
            mgr.free_slots(s.kv_indices)
    out["graph.replay_names"] = np.asarray(rnames)
    out["graph.max_bs_context_len"] = np.asarray([max_bs, context_len])

    # ---- slot write: the reference Triton kernel under the interpreter --------------------
    T, Hkv, D, pool = 11, 2, 64, 40
    g = torch.Generator().manual_seed(7)
    qkv = torch.randn((T, (4 + 2 * Hkv) * D), generator=g).to(torch.bfloat16)
    k = qkv[:, 4 * D:(4 + Hkv) * D].contiguous().view(T, Hkv, D)        # contiguous like post-RoPE k
    v = qkv[:, (4 + Hkv) * D:].view(T, Hkv, D)                          # strided view like models/qwen3.py:123
    k_cache = torch.zeros((pool, Hkv, D), dtype=torch.bfloat16)
    v_cache = torch.zeros((pool, Hkv, D), dtype=torch.bfloat16)
    slots = torch.tensor([5, 39, -1, 0, 17, 18, -1, 2, 30, 31, 7], dtype=torch.int64)
    ref_kv.store_kvcache(k, v, k_cache, v_cache, slots)
    out["store.seed_T_Hkv_D_pool"] = np.asarray([7, T, Hkv, D, pool])
    out["store.slots"] = slots.numpy()
    out["store.k_cache_bits"] = k_cache.view(torch.int16).numpy().copy()
    out["store.v_cache_bits"] = v_cache.view(torch.int16).numpy().copy()

    # ---- bench.py workload known answers (README.md:99-100) ---------------------------------
    from oracle.paged_attention_oracle import bench_workload
    prompts, max_tokens = bench_workload()
    out["bench.num_prompts_prompt_tokens_sum_max_tokens"] = np.asarray(
        [len(prompts), sum(len(p) for p in prompts), sum(max_tokens)])

    dst = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))),
                       "tests", "golden", "index_golden.npz")
    os.makedirs(os.path.dirname(dst), exist_ok=True)
    np.savez_compressed(dst, **out)
    print("wrote", dst, "with", len(out), "arrays")
    print("bench known answer:", out["bench.num_prompts_prompt_tokens_sum_max_tokens"])
    print("radix dedup:", out["radix_dedup_match"], "fresh:", out["alloc_fresh_first8"])


if __name__ == "__main__":
    main()
